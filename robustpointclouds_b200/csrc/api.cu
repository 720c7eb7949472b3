// api.cu -- the C-ABI of include/rpc_b200.h: argument validation, workspace carving, launches.
// No allocation, no host synchronisation, fixed launch topology per call (CUDA-graph capturable).
#include <math.h>
#include <string.h>
#include <stdlib.h>

#include "common.cuh"
#include "dynamic.cuh"
#include "vfe.cuh"
#include "hard_fwd.cuh"
#include "rows.cuh"
#include "dyn_scatter.cuh"
#include "allreduce.cuh"
#include "perturber.cuh"
#include "attention.cuh"
#include "linear.cuh"
#include "bnrelu.cuh"
#include "mlp_tc.cuh"
#include "wgrad_tc.cuh"

using namespace rpc;

namespace {

constexpr int kNumSM = 148;                       // B200
constexpr size_t kAlign = 256;
constexpr int kMinCtas = 16;                      // smallest grid of the persistent forward
inline size_t align_up(size_t x) { return (x + kAlign - 1) / kAlign * kAlign; }

int check_cfg(const rpc_voxel_cfg *cfg, bool hard)
{
    if (!cfg) return RPC_ERR_NULL;
    if (cfg->C < 3 || cfg->C > RPC_MAX_CHANNELS) return RPC_ERR_CONFIG;
    if (cfg->F < 1 || cfg->F > cfg->C) return RPC_ERR_CONFIG;
    for (int d = 0; d < 3; ++d) {
        if (!(cfg->voxel_size[d] > 0.0f)) return RPC_ERR_CONFIG;
        if (!(cfg->pcr[3 + d] > cfg->pcr[d])) return RPC_ERR_CONFIG;
    }
    if (hard && (cfg->max_points < 1 || cfg->max_voxels < 1)) return RPC_ERR_CONFIG;
    return RPC_OK;
}

// grid[d] = round((hi - lo) / vs) evaluated in fp32 (mmcv voxelization.cpp)
int make_grid(const rpc_voxel_cfg *cfg, GridDesc *gd)
{
    int64_t cells = 1;
    for (int d = 0; d < 3; ++d) {
        volatile float diff = cfg->pcr[3 + d] - cfg->pcr[d];
        volatile float q = diff / cfg->voxel_size[d];
        const float g = roundf(q);
        if (!(g >= 1.0f) || g > 2147483647.0f) return RPC_ERR_CONFIG;
        gd->lo[d] = cfg->pcr[d];
        gd->vs[d] = cfg->voxel_size[d];
        { volatile float r = 1.0f / cfg->voxel_size[d]; gd->rvs[d] = r; }
        gd->g[d] = (int)g;
        gd->gf[d] = (float)gd->g[d];
        cells *= gd->g[d];
        if (cells > 0x7FFFFFFELL) return RPC_ERR_TOO_LARGE;
    }
    gd->P = cfg->max_points; gd->V = cfg->max_voxels; gd->C = cfg->C; gd->F = cfg->F;
    return RPC_OK;
}

struct Plan {                                     // host-side layout of one batched call
    BatchDesc bd;
    int64_t ntotal, ntiles, tab_total, cap;
    size_t o_tab, o_hslot, o_tiles, o_slots, o_stamps, bytes;
};

// Tiles are kTile consecutive points of one sweep; tables hold 1.5 entries per point (worst-case load
// 0.67 when every point is its own cell; ~0.3 on LiDAR sweeps).
int make_plan(const rpc_voxel_cfg *cfg, int B, const int32_t *n, Plan *pl)
{
    if (!n) return RPC_ERR_NULL;
    if (B < 1 || B > RPC_MAX_BATCH) return RPC_ERR_BATCH;
    memset(pl, 0, sizeof(*pl));
    pl->bd.B = B;
    int64_t pts = 0, tiles = 0, tab = 0, cap = 0;
    for (int b = 0; b < B; ++b) {
        if (n[b] < 0) return RPC_ERR_BATCH;
        pl->bd.pt_begin[b] = (int)pts;
        pl->bd.tile_begin[b] = (int)tiles;
        int64_t sz = ((int64_t)n[b] * 3 + 1) / 2;
        if (sz < 64) sz = 64;
        sz = (sz + 3) & ~int64_t(3);                            // buckets of 4 entries (32 bytes)
        pl->bd.tab_off[b] = (uint32_t)tab;
        pl->bd.tab_size[b] = (uint32_t)sz;
        tab += sz;
        pts += n[b];
        tiles += (n[b] + kTile - 1) / kTile;
        cap += n[b] < cfg->max_voxels ? n[b] : cfg->max_voxels;
        if (pts > 0x3FFFFFF0LL || tab > 0x3FFFFFF0LL) return RPC_ERR_TOO_LARGE;   // bits 30, 31 of hslot are flags
    }
    if (tiles > (int64_t)kMaxTiles * 32) return RPC_ERR_TOO_LARGE;   // (the scan takes groups of up to 32 tiles)
    pl->bd.pt_begin[B] = (int)pts;
    pl->bd.tile_begin[B] = (int)tiles;
    pl->ntotal = pts; pl->ntiles = tiles; pl->tab_total = tab; pl->cap = cap;
    if (cap * (int64_t)cfg->max_points > 0x7FFFFFF0LL) return RPC_ERR_TOO_LARGE;
    size_t o = 0;
    pl->o_tab = o;   o += align_up(8 * (size_t)tab);
    pl->o_hslot = o; o += align_up(4 * (size_t)pts);
    pl->o_tiles = o; o += align_up(8 * (size_t)(tiles + 1));       // 8 per-warp byte counts per tile
    pl->o_slots = o; o += align_up(4 * (size_t)cap * cfg->max_points);
    pl->o_stamps = o; o += align_up(8 * 16 * 1024);
    pl->bytes = o;
    return RPC_OK;
}

// SM count of the current device (148 on B200); cached per device.  Falls back to 148 when there is
// no device (the sizing entry points must work on a CPU-only box).
int device_sms()
{
    static int cache[64] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { cudaGetLastError(); return kNumSM; }
    if (!cache[dev]) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n < 1) {
            cudaGetLastError();
            n = kNumSM;
        }
        cache[dev] = n;
    }
    return cache[dev];
}

void fill_perturb(const rpc_voxel_cfg *cfg, const rpc_perturb_cfg *pc, PerturbDesc *pd)
{
    memset(pd, 0, sizeof(*pd));
    if (pc)
        for (int c = 0; c < cfg->C; ++c) { pd->eb[c] = pc->eb[c]; pd->fb[c] = pc->fb[c]; }
}

// ---- optional per-stage timing (bench.py's roofline leg): cudaEvents recorded around every launch of
// the most recent call while enabled.  Off by default; never enabled under graph capture.
constexpr int kMaxStages = 12;
struct Profiler {
    bool on = false;
    int n = 0;                                    // stages recorded by the last call
    cudaEvent_t ev[kMaxStages + 1] = {};
    const char *name[kMaxStages] = {};
    void begin(cudaStream_t st) { if (on) { n = 0; mark(st); } }
    void mark(cudaStream_t st) {
        const int k = n;                          // ev[k] closes stage k-1
        if (!ev[k]) cudaEventCreate(&ev[k]);
        cudaEventRecord(ev[k], st);
    }
    void stage(const char *nm, cudaStream_t st) {
        if (!on || n >= kMaxStages) return;
        name[n++] = nm;
        mark(st);
    }
} g_prof;

inline bool misaligned(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) != 0; }

template <int C, bool PERTURB>
int launch_hard_t(FwdParams &fp, size_t smem, cudaStream_t st)
{
    auto kern = k_hard_fwd<C, PERTURB>;
    static bool attr_done[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && !attr_done[dev]) {
        if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024) != cudaSuccess)
            return RPC_ERR_CUDA;
        attr_done[dev] = true;
    }
    void *args[] = {&fp};
    // one CTA per kSub tiles is all a small batch can use, and a grid barrier over few CTAs is cheaper; the
    // per-CTA stamps are read for the full grid, so profiling keeps it
    int grid = device_sms() * kCtasPerSM;
    if (!fp.stamps) {
        const int need = (fp.ntiles + kSub - 1) / kSub;
        grid = need < kMinCtas ? (grid < kMinCtas ? grid : kMinCtas) : (need < grid ? need : grid);
    }
    g_prof.begin(st);
    if (cudaLaunchCooperativeKernel((void *)kern, dim3(grid), dim3(kPT), args, smem, st) != cudaSuccess)
        return RPC_ERR_CUDA;
    g_prof.stage("k_hard_fwd", st);
    return RPC_OK;
}

template <int C>
int launch_hard(const GridDesc &gd, const Plan &pl, const PerturbDesc &pd, const float *points,
                const float *raw, float *voxels, int32_t *coors, int32_t *num_points,
                int32_t *voxel_num, int32_t *p2v, int32_t *p2s, float *feats, void *workspace,
                cudaStream_t st)
{
    FwdParams fp;
    size_t smem = 0;
    // P5: slot lists of VT voxels per warp in shared memory; VT = 32 unless max_points is large
    fp.VT = 32;
    while (fp.VT > 4 && (size_t)(kPT / 32) * fp.VT * gd.P * sizeof(int) > 200 * 1024 / kCtasPerSM) fp.VT >>= 1;
    fp.gather_stride = fp.VT * gd.P;
    smem = (size_t)(kPT / 32) * fp.gather_stride * sizeof(int);
    if (smem > 200 * 1024 / kCtasPerSM) return RPC_ERR_CONFIG;         // max_points > 195
    int gs = 0;                                                     // P3 scans groups of 2^gs tiles (0 up to kMaxTiles tiles)
    while (((pl.ntiles + (1 << gs) - 1) >> gs) > kMaxTiles) ++gs;
    fp.gs = gs;
    const size_t scan_bytes = 4 * (size_t)(((pl.ntiles + (1 << gs) - 1) >> gs) + 1);   // P3's prefix array shares the buffer
    if (smem < scan_bytes) smem = scan_bytes;
    const size_t stage_bytes = C == 4 ? (size_t)2 * kSub * (raw ? 2 : 1) * kTile * 16 : 0;   // P1's TMA tile stages
    smem = smem > stage_bytes ? smem : stage_bytes;
    char *w = static_cast<char *>(workspace);
    fp.pts = points; fp.raw = raw;
    fp.tab = reinterpret_cast<unsigned long long *>(w + pl.o_tab);
    fp.hslot = reinterpret_cast<uint32_t *>(w + pl.o_hslot);
    fp.wcnt = reinterpret_cast<unsigned long long *>(w + pl.o_tiles);
    fp.slots = reinterpret_cast<int *>(w + pl.o_slots);
    fp.stamps = g_prof.on ? reinterpret_cast<unsigned long long *>(w + pl.o_stamps) : nullptr;
    fp.voxels = voxels; fp.coors = coors; fp.num_points = num_points; fp.voxel_num = voxel_num;
    fp.p2v = p2v; fp.p2s = p2s; fp.feats = feats;
    fp.tab_total = pl.tab_total;
    fp.ntiles = (int)pl.ntiles;
    fp.gd = gd; fp.bd = pl.bd; fp.pd = pd;
    return raw ? launch_hard_t<C, true>(fp, smem, st) : launch_hard_t<C, false>(fp, smem, st);
}

}  // namespace

// =========================================================================================== C-ABI
extern "C" {

#define RPC_API __attribute__((visibility("default")))

RPC_API int rpc_abi_version(void) { return RPC_ABI_VERSION; }

RPC_API const char *rpc_strerror(int status)
{
    switch (status) {
        case RPC_OK: return "ok";
        case RPC_ERR_NULL: return "required pointer is NULL";
        case RPC_ERR_CONFIG: return "invalid voxel configuration (C, F, P, V, voxel_size or range)";
        case RPC_ERR_BATCH: return "batch size outside 1..32 or negative point count";
        case RPC_ERR_WORKSPACE: return "workspace too small (see rpc_workspace_bytes)";
        case RPC_ERR_CUDA: return "CUDA launch failed";
        case RPC_ERR_TOO_LARGE: return "point count, table or grid exceeds 31-bit indexing";
        case RPC_ERR_ALIGN: return "pointer not 16-byte aligned";
        case RPC_ERR_STATIC_LIMIT: return "input larger than the static-scratch entry point allows (RPC_STATIC_FILTER_MAX_POINTS rows)";
        default: return "unknown status";
    }
}

RPC_API void rpc_profile_enable(int on) { g_prof.on = on != 0; g_prof.n = 0; }

RPC_API int rpc_profile_read(float *ms, const char **names, int max_stages)
{
    const int n = g_prof.n < max_stages ? g_prof.n : max_stages;
    if (n <= 0) return 0;
    if (cudaEventSynchronize(g_prof.ev[g_prof.n]) != cudaSuccess) return RPC_ERR_CUDA;
    for (int k = 0; k < n; ++k) {
        if (ms && cudaEventElapsedTime(&ms[k], g_prof.ev[k], g_prof.ev[k + 1]) != cudaSuccess) return RPC_ERR_CUDA;
        if (names) names[k] = g_prof.name[k];
    }
    return n;
}

RPC_API int rpc_grid_size(const rpc_voxel_cfg *cfg, int32_t grid[3])
{
    if (!grid) return RPC_ERR_NULL;
    int st = check_cfg(cfg, false);
    if (st) return st;
    GridDesc gd;
    st = make_grid(cfg, &gd);
    if (st) return st;
    grid[0] = gd.g[0]; grid[1] = gd.g[1]; grid[2] = gd.g[2];
    return RPC_OK;
}

RPC_API size_t rpc_workspace_bytes(const rpc_voxel_cfg *cfg, int B, const int32_t *n_per_sample)
{
    Plan pl;
    if (check_cfg(cfg, true) || make_plan(cfg, B, n_per_sample, &pl)) return 0;
    return pl.bytes;
}

RPC_API int64_t rpc_workspace_stamps_offset(const rpc_voxel_cfg *cfg, int B, const int32_t *n_per_sample)
{
    Plan pl;
    if (check_cfg(cfg, true) || make_plan(cfg, B, n_per_sample, &pl)) return -1;
    return (int64_t)pl.o_stamps;
}

RPC_API int64_t rpc_voxel_capacity(const rpc_voxel_cfg *cfg, int B, const int32_t *n_per_sample)
{
    Plan pl;
    if (check_cfg(cfg, true) || make_plan(cfg, B, n_per_sample, &pl)) return -1;
    return pl.cap;
}

RPC_API int rpc_voxelize_hard(const rpc_voxel_cfg *cfg, const float *points,
                              const int32_t *n_per_sample, int B, const float *raw,
                              const rpc_perturb_cfg *pc, float *voxels, int32_t *coors,
                              int32_t *num_points, int32_t *voxel_num, int32_t *point2voxel,
                              int32_t *point2slot, float *feats, void *workspace,
                              size_t workspace_bytes, void *stream)
{
    int st = check_cfg(cfg, true);
    if (st) return st;
    GridDesc gd;
    if ((st = make_grid(cfg, &gd))) return st;
    Plan pl;
    if ((st = make_plan(cfg, B, n_per_sample, &pl))) return st;
    if (!voxel_num || !workspace) return RPC_ERR_NULL;
    if (pl.ntotal > 0 && !points) return RPC_ERR_NULL;
    if (pl.cap > 0 && (!voxels || !coors || !num_points)) return RPC_ERR_NULL;
    if (raw && !pc) return RPC_ERR_NULL;
    if (workspace_bytes < pl.bytes) return RPC_ERR_WORKSPACE;
    if (misaligned(points) || misaligned(raw) || misaligned(voxels) || misaligned(coors) ||
        (reinterpret_cast<uintptr_t>(workspace) & 31u) != 0 || misaligned(feats))      // buckets are read as 32 bytes
        return RPC_ERR_ALIGN;
    PerturbDesc pd;
    fill_perturb(cfg, pc, &pd);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    switch (cfg->C) {
#define RPC_CASE(CC)                                                                              \
    case CC: return launch_hard<CC>(gd, pl, pd, points, raw, voxels, coors, num_points, voxel_num, \
                                    point2voxel, point2slot, feats, workspace, s);
        RPC_CASE(3) RPC_CASE(4) RPC_CASE(5) RPC_CASE(6) RPC_CASE(7) RPC_CASE(8)
#undef RPC_CASE
    }
    return RPC_ERR_CONFIG;
}

RPC_API int rpc_step_bwd(const rpc_voxel_cfg *cfg, const float *grad_feats,
                         const int32_t *num_points, const int32_t *point2voxel, const float *raw,
                         const rpc_perturb_cfg *pc, int64_t ntotal, float *grad_raw, void *stream)
{
    int st = check_cfg(cfg, false);
    if (st) return st;
    if (ntotal < 0 || ntotal > 0x7FFFFFF0LL) return RPC_ERR_TOO_LARGE;
    if (ntotal == 0) return RPC_OK;
    if (!grad_feats || !num_points || !point2voxel || !grad_raw) return RPC_ERR_NULL;
    if (raw && !pc) return RPC_ERR_NULL;
    if (misaligned(grad_feats) || misaligned(raw) || misaligned(grad_raw)) return RPC_ERR_ALIGN;
    PerturbDesc pd;
    fill_perturb(cfg, pc, &pd);
    int64_t blocks = (ntotal + kThreads - 1) / kThreads;
    if (blocks > kNumSM * 16) blocks = kNumSM * 16;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    g_prof.begin(s);
    switch (cfg->C) {
#define RPC_CASE(CC)                                                                               \
    case CC:                                                                                       \
        k_step_bwd<CC><<<(int)blocks, kThreads, 0, s>>>(grad_feats, num_points, point2voxel, raw, pd, \
                                                        ntotal, cfg->F, grad_raw);                 \
        break;
        RPC_CASE(3) RPC_CASE(4) RPC_CASE(5) RPC_CASE(6) RPC_CASE(7) RPC_CASE(8)
#undef RPC_CASE
        default: return RPC_ERR_CONFIG;
    }
    g_prof.stage("k_step_bwd", s);
    return cudaPeekAtLastError() == cudaSuccess ? RPC_OK : RPC_ERR_CUDA;
}

RPC_API int rpc_vfe_mean_fwd(const float *voxels, const int32_t *num_points, int64_t M, int P, int C,
                             int F, const int32_t *m_dev, float *out, void *stream)
{
    if (P < 1 || C < 1 || F < 1 || F > C || M < 0) return RPC_ERR_CONFIG;
    if (M == 0) return RPC_OK;
    if (!voxels || !num_points || !out) return RPC_ERR_NULL;
    if (misaligned(voxels)) return RPC_ERR_ALIGN;
    // voxel rows per CTA iteration: the [VB,P,C] tile is staged in shared memory (<= 32 KB)
    int VB = (int)((32 * 1024) / ((size_t)P * C * sizeof(float))) & ~3;
    if (VB > 64) VB = 64;
    if (VB < 4) VB = 4;
    const size_t smem = (size_t)VB * P * C * sizeof(float);
    if (smem > 200 * 1024) return RPC_ERR_CONFIG;
    if (smem > 48 * 1024 &&
        cudaFuncSetAttribute(k_vfe_mean_fwd, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
        return RPC_ERR_CUDA;
    int64_t blocks = (M + VB - 1) / VB;
    if (blocks > kNumSM * 8) blocks = kNumSM * 8;
    k_vfe_mean_fwd<<<(int)blocks, kThreads, smem, static_cast<cudaStream_t>(stream)>>>(
        voxels, num_points, M, m_dev, P, C, F, VB, out);
    return cudaPeekAtLastError() == cudaSuccess ? RPC_OK : RPC_ERR_CUDA;
}

RPC_API int rpc_vfe_mean_bwd(const float *grad_out, const int32_t *num_points, int64_t M, int P, int C,
                             int F, const int32_t *m_dev, float *grad_voxels, void *stream)
{
    if (P < 1 || C < 1 || F < 1 || F > C || M < 0) return RPC_ERR_CONFIG;
    if (M == 0) return RPC_OK;
    if (!grad_out || !num_points || !grad_voxels) return RPC_ERR_NULL;
    int64_t blocks = (M * P * C + kThreads - 1) / kThreads;
    if (blocks > kNumSM * 16) blocks = kNumSM * 16;
    k_vfe_mean_bwd<<<(int)blocks, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
        grad_out, num_points, M, m_dev, P, C, F, grad_voxels);
    return cudaPeekAtLastError() == cudaSuccess ? RPC_OK : RPC_ERR_CUDA;
}

RPC_API int rpc_voxelize_dynamic(const rpc_voxel_cfg *cfg, const float *points,
                                 const int32_t *n_per_sample, int B, const float *raw,
                                 const rpc_perturb_cfg *pc, int32_t *coors, int ncol, void *stream)
{
    int st = check_cfg(cfg, false);
    if (st) return st;
    if (ncol != 3 && ncol != 4) return RPC_ERR_CONFIG;
    GridDesc gd;
    if ((st = make_grid(cfg, &gd))) return st;
    rpc_voxel_cfg c2 = *cfg;
    c2.max_points = 1; c2.max_voxels = 1;
    Plan pl;
    if ((st = make_plan(&c2, B, n_per_sample, &pl))) return st;
    if (pl.ntotal == 0) return RPC_OK;
    if (!points || !coors) return RPC_ERR_NULL;
    if (raw && !pc) return RPC_ERR_NULL;
    if (misaligned(points) || misaligned(raw) || misaligned(coors)) return RPC_ERR_ALIGN;
    PerturbDesc pd;
    fill_perturb(cfg, pc, &pd);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    int64_t nblk = (pl.ntotal + kThreads - 1) / kThreads;
    if (nblk > kNumSM * 16) nblk = kNumSM * 16;
    const int nt = (int)nblk;
#define RPC_LAUNCH(CC, PP, NC) \
    k_dynamic_voxelize<CC, PP, NC><<<nt, kThreads, 0, s>>>(points, raw, gd, pl.bd, pd, coors)
#define RPC_CASE(CC)                                                              \
    case CC:                                                                      \
        if (raw) { if (ncol == 4) RPC_LAUNCH(CC, true, 4); else RPC_LAUNCH(CC, true, 3); }   \
        else     { if (ncol == 4) RPC_LAUNCH(CC, false, 4); else RPC_LAUNCH(CC, false, 3); } \
        break;
    switch (cfg->C) {
        RPC_CASE(3) RPC_CASE(4) RPC_CASE(5) RPC_CASE(6) RPC_CASE(7) RPC_CASE(8)
        default: return RPC_ERR_CONFIG;
    }
#undef RPC_CASE
#undef RPC_LAUNCH
    return cudaPeekAtLastError() == cudaSuccess ? RPC_OK : RPC_ERR_CUDA;
}

}  // extern "C" (closed for the dynamic-scatter helpers)

// ------------------------------------------------------------------ dynamic scatter (config 4)
namespace {
struct DynPlan {
    Plan pl;
    long long Gp, acap;
    int W1;
    size_t o_coarse, o_fine, o_gpre, o_cell, o_pslot, o_part, o_ssb, o_stamps, o_acc, bytes;
    uint32_t w1p;
    int acc_vec;
};
constexpr int kDynMaxCtas = 4096;

int make_dyn_plan(const rpc_voxel_cfg *cfg, const GridDesc &gd, int B, const int32_t *n, DynPlan *dp)
{
    rpc_voxel_cfg c2 = *cfg;
    c2.max_points = 1; c2.max_voxels = 1;
    int st = make_plan(&c2, B, n, &dp->pl);
    if (st) return st;
    const long long G = (long long)gd.g[0] * gd.g[1] * gd.g[2];
    dp->Gp = (G + kCoarseCells - 1) / kCoarseCells * kCoarseCells;
    const long long w1 = dp->Gp / kCoarseCells * B;
    if (w1 > 0x3FFFFFFFLL) return RPC_ERR_TOO_LARGE;
    dp->W1 = (int)w1;
    const long long groups = dp->Gp / kGroupCells * B;
    dp->acap = dp->pl.ntotal < groups ? dp->pl.ntotal : groups;
    if (dp->acap < 1) dp->acap = 1;
    size_t o = 0;
    dp->w1p = 1;
    while (dp->w1p < (uint32_t)dp->W1) dp->w1p <<= 1;
    dp->o_coarse = o; o += align_up(8 * (size_t)dp->w1p);
    dp->o_fine = o;   o += align_up(32 * (size_t)dp->acap);
    dp->o_gpre = o;   o += align_up(4 * (size_t)dp->acap);
    dp->o_cell = o;   o += align_up(4 * (size_t)dp->pl.ntotal);
    dp->o_pslot = o;  o += align_up(4 * (size_t)dp->pl.ntotal);
    dp->o_part = o;   o += align_up(4 * 2 * (size_t)kDynMaxCtas);
    dp->o_ssb = o;    o += align_up(4 * (RPC_MAX_BATCH + 1));
    dp->o_stamps = o; o += align_up(8 * 32 * (size_t)kDynMaxCtas);
    dp->acc_vec = (cfg->F + 1 + 3) / 4;
    dp->o_acc = o;    o += align_up(16 * (size_t)dp->acc_vec * (size_t)(dp->pl.ntotal > 0 ? dp->pl.ntotal : 1));
    dp->bytes = o;
    return RPC_OK;
}

// n / d for n < 2^31 as (umulhi(n, mg) + n) >> sh
void fast_div_magic(uint32_t d, uint32_t *mg, uint32_t *sh)
{
    uint32_t s = 0;
    while ((1ull << s) < d) ++s;
    *sh = s;
    *mg = (uint32_t)((((1ull << s) - d) << 32) / d + 1);
}

template <int C, bool PERTURB>
int launch_dyn_t(DynParams &p, cudaStream_t st)
{
    auto kern = k_dyn_scatter<C, PERTURB>;
    int grid = device_sms() * kDynCtasPerSM;
    if (grid > kDynMaxCtas) grid = kDynMaxCtas;
    const size_t smem = C == 4 ? 0 : (size_t)(kDynThreads / 32) * (PERTURB ? 2 : 1) * 32 * C * sizeof(float);
    static bool attr_done[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (smem > 48 * 1024 && dev >= 0 && dev < 64 && !attr_done[dev]) {
        if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
            return RPC_ERR_CUDA;
        attr_done[dev] = true;
    }
    void *args[] = {&p};
    g_prof.begin(st);
    if (cudaLaunchCooperativeKernel((void *)kern, dim3(grid), dim3(kDynThreads), args, smem, st) != cudaSuccess)
        return RPC_ERR_CUDA;
    g_prof.stage("k_dyn_scatter", st);
    return RPC_OK;
}
}  // namespace

extern "C" {

RPC_API size_t rpc_dynamic_workspace_bytes(const rpc_voxel_cfg *cfg, int B, const int32_t *n_per_sample)
{
    GridDesc gd;
    DynPlan dp;
    if (check_cfg(cfg, false) || make_grid(cfg, &gd) || make_dyn_plan(cfg, gd, B, n_per_sample, &dp)) return 0;
    return dp.bytes;
}

RPC_API int64_t rpc_dynamic_stamps_offset(const rpc_voxel_cfg *cfg, int B, const int32_t *n_per_sample)
{
    GridDesc gd;
    DynPlan dp;
    if (check_cfg(cfg, false) || make_grid(cfg, &gd) || make_dyn_plan(cfg, gd, B, n_per_sample, &dp)) return -1;
    return (int64_t)dp.o_stamps;
}

RPC_API int rpc_dynamic_scatter_mean_fwd(const rpc_voxel_cfg *cfg, const float *points,
                                         const int32_t *n_per_sample, int B, const float *raw,
                                         const rpc_perturb_cfg *pc, const int32_t *coors_in, int32_t *coors_out,
                                         int32_t *point2voxel, float *feats, int32_t *voxel_coors,
                                         int32_t *num_points, int32_t *voxel_num, void *workspace,
                                         size_t workspace_bytes, void *stream)
{
    int st = check_cfg(cfg, false);
    if (st) return st;
    GridDesc gd;
    if ((st = make_grid(cfg, &gd))) return st;
    DynPlan dp;
    if ((st = make_dyn_plan(cfg, gd, B, n_per_sample, &dp))) return st;
    if (dp.pl.ntotal > (1 << 24)) return RPC_ERR_TOO_LARGE;         // per-voxel counts ride in fp32 accumulators
    if (!voxel_num || !workspace) return RPC_ERR_NULL;
    if (dp.pl.ntotal > 0 && (!points || !point2voxel || !feats || !voxel_coors || !num_points)) return RPC_ERR_NULL;
    if (raw && !pc) return RPC_ERR_NULL;
    if (workspace_bytes < dp.bytes) return RPC_ERR_WORKSPACE;
    if (misaligned(points) || misaligned(raw) || misaligned(coors_in) || misaligned(coors_out) ||
        misaligned(feats) || misaligned(voxel_coors) || misaligned(workspace))
        return RPC_ERR_ALIGN;
    DynParams p;
    char *w = static_cast<char *>(workspace);
    p.pts = points; p.raw = raw; p.coors_in = coors_in; p.coors_out = coors_in ? nullptr : coors_out;
    p.p2v = point2voxel; p.feats = feats; p.vcoors = voxel_coors; p.count = num_points; p.voxel_num = voxel_num;
    p.coarse = reinterpret_cast<uint2 *>(w + dp.o_coarse);
    p.fine = reinterpret_cast<uint32_t *>(w + dp.o_fine);
    p.gpre = reinterpret_cast<uint32_t *>(w + dp.o_gpre);
    p.cell = reinterpret_cast<int *>(w + dp.o_cell);
    p.pslot = reinterpret_cast<int *>(w + dp.o_pslot);
    p.part = reinterpret_cast<uint32_t *>(w + dp.o_part);
    p.ssb = reinterpret_cast<int *>(w + dp.o_ssb);
    p.acc = reinterpret_cast<float4 *>(w + dp.o_acc);
    p.acc_vec = dp.acc_vec;
    p.w1_mask = dp.w1p - 1;
    fast_div_magic((uint32_t)gd.g[0] * (uint32_t)gd.g[1], &p.mg_gxy, &p.sh_gxy);
    fast_div_magic((uint32_t)gd.g[0], &p.mg_gx, &p.sh_gx);
    p.stamps = g_prof.on ? reinterpret_cast<unsigned long long *>(w + dp.o_stamps) : nullptr;
    p.Gp = dp.Gp; p.W1 = dp.W1; p.ntotal = (int)dp.pl.ntotal;
    p.gd = gd; p.bd = dp.pl.bd;
    fill_perturb(cfg, pc, &p.pd);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    switch (cfg->C) {
#define RPC_CASE(CC) \
    case CC: return raw ? launch_dyn_t<CC, true>(p, s) : launch_dyn_t<CC, false>(p, s);
        RPC_CASE(3) RPC_CASE(4) RPC_CASE(5) RPC_CASE(6) RPC_CASE(7) RPC_CASE(8)
#undef RPC_CASE
    }
    return RPC_ERR_CONFIG;
}

}  // extern "C"

extern "C" {
// ------------------------------------------------------------------ reference-order composition
RPC_API size_t rpc_valid_rows_workspace_bytes(int64_t M, int P)
{
    if (M < 0 || P < 1) return 0;
    return align_up(4 * (size_t)((M * P + kRowTile - 1) / kRowTile + 1));
}

static int launch_valid_rows(const float *voxels, const int32_t *num_points, int64_t M, int P, int C, int mode,
                             const int32_t *m_dev, int32_t *row_idx, int32_t *row2k, float *rows_out, int32_t *count,
                             void *workspace, size_t workspace_bytes, const RangeDesc &rd, cudaStream_t st)
{
    const int64_t R = M * P;
    if (R > 0x7FFFFFF0LL) return RPC_ERR_TOO_LARGE;
    int ntiles = (int)((R + kRowTile - 1) / kRowTile);
    if (ntiles > kMaxTiles) return RPC_ERR_TOO_LARGE;
    if (workspace_bytes < rpc_valid_rows_workspace_bytes(M, P)) return RPC_ERR_WORKSPACE;
    int *tile_cnt = static_cast<int *>(workspace);
    size_t smem = 4 * (size_t)(ntiles + 1);
    static bool attr_done[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && !attr_done[dev]) {
        if (cudaFuncSetAttribute(k_valid_rows, cudaFuncAttributeMaxDynamicSharedMemorySize, 4 * (kMaxTiles + 1)) != cudaSuccess)
            return RPC_ERR_CUDA;
        attr_done[dev] = true;
    }
    int grid = device_sms();
    if (grid > ntiles) grid = ntiles;
    RangeDesc rdc = rd;
    void *args[] = {&voxels, &num_points, &M, &m_dev, &P, &C, &mode, &ntiles, &tile_cnt, &row_idx, &row2k, &rows_out, &count, &rdc};
    if (cudaLaunchCooperativeKernel((void *)k_valid_rows, dim3(grid), dim3(kRowTile), args, smem, st) != cudaSuccess)
        return RPC_ERR_CUDA;
    return RPC_OK;
}

RPC_API int rpc_valid_rows(const float *voxels, const int32_t *num_points, int64_t M, int P, int C, int mode,
                           const int32_t *m_dev, int32_t *row_idx, int32_t *row2k, float *rows_out,
                           int32_t *count, void *workspace, size_t workspace_bytes, void *stream)
{
    if (P < 1 || C < 1 || M < 0 || (mode != 0 && mode != 1)) return RPC_ERR_CONFIG;
    if (!count) return RPC_ERR_NULL;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (M == 0) return cudaMemsetAsync(count, 0, 4, st) == cudaSuccess ? RPC_OK : RPC_ERR_CUDA;
    if (!voxels || !num_points || !row_idx || !workspace) return RPC_ERR_NULL;
    RangeDesc rd;
    memset(&rd, 0, sizeof(rd));
    return launch_valid_rows(voxels, num_points, M, P, C, mode, m_dev, row_idx, row2k, rows_out, count, workspace,
                             workspace_bytes, rd, st);
}

RPC_API int rpc_points_range_filter(const float *points, int64_t N, int C, const float *range6, float *out_points,
                                    int32_t *out_index, int32_t *count, void *workspace, size_t workspace_bytes,
                                    void *stream)
{
    if (C < 3 || N < 0) return RPC_ERR_CONFIG;
    if (!count || !range6) return RPC_ERR_NULL;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (N == 0) return cudaMemsetAsync(count, 0, 4, st) == cudaSuccess ? RPC_OK : RPC_ERR_CUDA;
    if (!points || !out_index || !workspace) return RPC_ERR_NULL;
    RangeDesc rd;
    for (int d = 0; d < 3; ++d) { rd.lo[d] = range6[d]; rd.hi[d] = range6[3 + d]; }
    return launch_valid_rows(points, nullptr, N, 1, C, 2, nullptr, out_index, nullptr, out_points, count, workspace,
                             workspace_bytes, rd, st);
}

static inline int stream_blocks(int64_t total)
{
    int64_t b = (total + kThreads - 1) / kThreads;
    if (b > kNumSM * 16) b = kNumSM * 16;
    return (int)(b < 1 ? 1 : b);
}

RPC_API int rpc_scatter_rows(const float *voxels, const int32_t *row2k, const float *rows, int64_t M, int P,
                             int C, float *out, void *stream)
{
    if (P < 1 || C < 1 || M < 0) return RPC_ERR_CONFIG;
    if (M == 0) return RPC_OK;
    if (!voxels || !row2k || !rows || !out) return RPC_ERR_NULL;
    k_scatter_rows<<<stream_blocks(M * P * C), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
        voxels, row2k, rows, M * P, C, out);
    return cudaPeekAtLastError() == cudaSuccess ? RPC_OK : RPC_ERR_CUDA;
}

RPC_API int rpc_gather_rows(const float *src, const int32_t *row_idx, int64_t K, const int32_t *k_dev, int C,
                            float *out, void *stream)
{
    if (C < 1 || K < 0) return RPC_ERR_CONFIG;
    if (K == 0) return RPC_OK;
    if (!src || !row_idx || !out) return RPC_ERR_NULL;
    k_gather_rows<<<stream_blocks(K * C), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(src, row_idx, K, k_dev, C, out);
    return cudaPeekAtLastError() == cudaSuccess ? RPC_OK : RPC_ERR_CUDA;
}

RPC_API int rpc_voxel_centers(const rpc_voxel_cfg *cfg, const int32_t *coors, int ncol, int64_t M, const int32_t *m_dev,
                              float *centers, void *stream)
{
    int st = check_cfg(cfg, false);
    if (st) return st;
    if ((ncol != 3 && ncol != 4) || M < 0) return RPC_ERR_CONFIG;
    if (M == 0) return RPC_OK;
    if (!coors || !centers) return RPC_ERR_NULL;
    CenterDesc cd;
    for (int d = 0; d < 3; ++d) { cd.vs[d] = cfg->voxel_size[d]; cd.lo[d] = cfg->pcr[d]; }
    k_voxel_centers<<<stream_blocks(M * 3), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(coors, ncol, M, m_dev, cd, centers);
    return cudaPeekAtLastError() == cudaSuccess ? RPC_OK : RPC_ERR_CUDA;
}

RPC_API int rpc_perturb_apply_fwd(const float *x, const float *raw, const rpc_perturb_cfg *pc, int64_t K,
                                  const int32_t *k_dev, int C, float *y, float *p_out, void *stream)
{
    if (C < 1 || C > RPC_MAX_CHANNELS || K < 0) return RPC_ERR_CONFIG;
    if (K == 0) return RPC_OK;
    if (!x || !raw || !pc || !y) return RPC_ERR_NULL;
    PerturbDesc pd;
    memset(&pd, 0, sizeof(pd));
    for (int c = 0; c < C; ++c) { pd.eb[c] = pc->eb[c]; pd.fb[c] = pc->fb[c]; }
    k_perturb_apply_fwd<<<stream_blocks(K * C), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(x, raw, pd, K, k_dev, C, y, p_out);
    return cudaPeekAtLastError() == cudaSuccess ? RPC_OK : RPC_ERR_CUDA;
}

RPC_API int rpc_perturb_apply_bwd(const float *grad_p, const float *raw, const rpc_perturb_cfg *pc, int64_t K,
                                  const int32_t *k_dev, int C, float *grad_raw, void *stream)
{
    if (C < 1 || C > RPC_MAX_CHANNELS || K < 0) return RPC_ERR_CONFIG;
    if (K == 0) return RPC_OK;
    if (!grad_p || !raw || !pc || !grad_raw) return RPC_ERR_NULL;
    PerturbDesc pd;
    memset(&pd, 0, sizeof(pd));
    for (int c = 0; c < C; ++c) { pd.eb[c] = pc->eb[c]; pd.fb[c] = pc->fb[c]; }
    k_perturb_apply_bwd<<<stream_blocks(K * C), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(grad_p, raw, pd, K, k_dev, C, grad_raw);
    return cudaPeekAtLastError() == cudaSuccess ? RPC_OK : RPC_ERR_CUDA;
}

// ------------------------------------------------------------------ perturber-gradient all-reduce over peer memory
RPC_API size_t rpc_allreduce_flag_bytes(void) { return (size_t)kArCtas * kArMaxWorld * sizeof(uint32_t); }

RPC_API int rpc_allreduce_oneshot(float *const *stage0, float *const *stage1, uint32_t *const *peer_flags, int rank,
                                  int world, int64_t n, const float *src, float *out, uint32_t *counter,
                                  int32_t *status, void *stream)
{
    if (world < 1 || world > kArMaxWorld || rank < 0 || rank >= world || n < 0) return RPC_ERR_CONFIG;
    if (n == 0) return RPC_OK;
    if (!stage0 || !stage1 || !peer_flags || !src || !out || !counter || !status) return RPC_ERR_NULL;
    ArParams p;
    memset(&p, 0, sizeof(p));
    for (int r = 0; r < world; ++r) {
        if (!stage0[r] || !stage1[r] || !peer_flags[r]) return RPC_ERR_NULL;
        if (misaligned(stage0[r]) || misaligned(stage1[r])) return RPC_ERR_ALIGN;
        p.stage[0][r] = stage0[r];
        p.stage[1][r] = stage1[r];
        p.flags[r] = peer_flags[r];
    }
    if (misaligned(out) || misaligned(src)) return RPC_ERR_ALIGN;
    p.src = src; p.out = out; p.ctr = counter; p.status = status; p.n = n; p.rank = rank; p.world = world;
    int64_t ctas = ((n >> 2) + kArThreads - 1) / kArThreads;
    if (ctas > kArCtas) ctas = kArCtas;
    if (ctas < 1) ctas = 1;
    k_allreduce_oneshot<<<(int)ctas, kArThreads, 0, static_cast<cudaStream_t>(stream)>>>(p);
    return cudaPeekAtLastError() == cudaSuccess ? RPC_OK : RPC_ERR_CUDA;
}

RPC_API size_t rpc_allreduce_ll_bytes(int world, int64_t n)
{
    if (world < 1 || world > kArMaxWorld || n < 0) return 0;
    const int64_t npad = (n + 1) / 2 * 2;                           // slots of a source rank stay 16-byte aligned
    return (size_t)2 * world * npad * sizeof(uint2);
}

RPC_API int rpc_allreduce_ll(void *const *recv, int rank, int world, int64_t n, const float *src, float *out,
                             uint32_t *counter, int32_t *status, void *stream)
{
    if (world < 1 || world > kArMaxWorld || rank < 0 || rank >= world || n < 0) return RPC_ERR_CONFIG;
    if (n == 0) return RPC_OK;
    if (!recv || !src || !out || !counter || !status) return RPC_ERR_NULL;
    ArLLParams p;
    memset(&p, 0, sizeof(p));
    p.npad = (n + 1) / 2 * 2;
    for (int r = 0; r < world; ++r) {
        if (!recv[r]) return RPC_ERR_NULL;
        if (misaligned(recv[r])) return RPC_ERR_ALIGN;
        p.recv[0][r] = static_cast<uint2 *>(recv[r]);
        p.recv[1][r] = static_cast<uint2 *>(recv[r]) + (int64_t)world * p.npad;
    }
    p.src = src; p.out = out; p.ctr = counter; p.status = status; p.n = n; p.rank = rank; p.world = world;
    int64_t ctas = (n + kArThreads - 1) / kArThreads;
    if (ctas > kArLLCtas) ctas = kArLLCtas;
    k_allreduce_ll<<<(int)ctas, kArThreads, 0, static_cast<cudaStream_t>(stream)>>>(p);
    return cudaPeekAtLastError() == cudaSuccess ? RPC_OK : RPC_ERR_CUDA;
}

// ------------------------------------------------------------------ perturber statistics
static int moments_launch(const float *x, int64_t K, const int32_t *k_dev, int C, double *acc, cudaStream_t st)
{
    if (cudaMemsetAsync(acc, 0, sizeof(double) * kMomVals, st) != cudaSuccess) return RPC_ERR_CUDA;
    int64_t blocks = (K + kThreads - 1) / kThreads;
    if (blocks > kNumSM * 4) blocks = kNumSM * 4;
    if (blocks < 1) blocks = 1;
    switch (C) {
#define RPC_CASE(CC) case CC: k_col_moments<CC><<<(int)blocks, kThreads, 0, st>>>(x, K, k_dev, acc); break;
        RPC_CASE(1) RPC_CASE(2) RPC_CASE(3) RPC_CASE(4) RPC_CASE(5) RPC_CASE(6) RPC_CASE(7) RPC_CASE(8)
#undef RPC_CASE
        default: return RPC_ERR_CONFIG;
    }
    return cudaPeekAtLastError() == cudaSuccess ? RPC_OK : RPC_ERR_CUDA;
}

RPC_API size_t rpc_moments_bytes(void) { return sizeof(double) * kMomVals; }

RPC_API int rpc_normalize_fwd(const float *x, int64_t K, const int32_t *k_dev, int C, double *moments, float *xn,
                              void *stream)
{
    if (C < 1 || C > RPC_MAX_CHANNELS || K < 0) return RPC_ERR_CONFIG;
    if (K == 0) return RPC_OK;
    if (!x || !moments || !xn) return RPC_ERR_NULL;
    if (C == 4 && misaligned(x)) return RPC_ERR_ALIGN;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int rc = moments_launch(x, K, k_dev, C, moments, st);
    if (rc) return rc;
    k_normalize_fwd<<<stream_blocks(K * C), kThreads, 0, st>>>(x, moments, K, k_dev, C, xn);
    return cudaPeekAtLastError() == cudaSuccess ? RPC_OK : RPC_ERR_CUDA;
}

RPC_API int rpc_perturb_losses_fwd(const float *p, int64_t K, const int32_t *k_dev, int C, double *moments,
                                   float *losses, float *consts, void *stream)
{
    if (C < 2 || C > RPC_MAX_CHANNELS || K < 0) return RPC_ERR_CONFIG;
    if (!p || !moments || !losses || !consts) return RPC_ERR_NULL;
    if (C == 4 && misaligned(p)) return RPC_ERR_ALIGN;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int rc = moments_launch(p, K, k_dev, C, moments, st);
    if (rc) return rc;
    k_losses_finalize<<<1, 32, 0, st>>>(moments, K, k_dev, C, losses, consts);
    return cudaPeekAtLastError() == cudaSuccess ? RPC_OK : RPC_ERR_CUDA;
}

RPC_API int rpc_perturb_losses_bwd(const float *p, const float *consts, const float *grad_losses, int64_t K,
                                   const int32_t *k_dev, int C, float *grad_p, void *stream)
{
    if (C < 2 || C > RPC_MAX_CHANNELS || K < 0) return RPC_ERR_CONFIG;
    if (K == 0) return RPC_OK;
    if (!p || !consts || !grad_losses || !grad_p) return RPC_ERR_NULL;
    int64_t blocks = (K + kThreads - 1) / kThreads;
    if (blocks > kNumSM * 16) blocks = kNumSM * 16;
    k_losses_bwd<<<(int)blocks, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(p, consts, grad_losses, K, k_dev, C, grad_p);
    return cudaPeekAtLastError() == cudaSuccess ? RPC_OK : RPC_ERR_CUDA;
}

RPC_API int rpc_momentum_apply_fwd(const float *x, const float *y, const float *last, int y_is_delta, float scaling,
                                   float alpha, int64_t n, float *out, float *sp_out, double *sumsq, float *l2, void *stream)
{
    if (n < 0) return RPC_ERR_CONFIG;
    if (!sumsq || !l2) return RPC_ERR_NULL;
    if (n > 0 && (!x || !y || !out || !sp_out)) return RPC_ERR_NULL;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (cudaMemsetAsync(sumsq, 0, sizeof(double), st) != cudaSuccess) return RPC_ERR_CUDA;
    if (n > 0) {
        int64_t blocks = (n + kThreads - 1) / kThreads;
        if (blocks > kNumSM * 8) blocks = kNumSM * 8;
        k_momentum_apply_fwd<<<(int)blocks, kThreads, 0, st>>>(x, y, last, y_is_delta, scaling, alpha, n, out, sp_out, sumsq);
    }
    k_momentum_l2<<<1, 32, 0, st>>>(sumsq, l2);
    return cudaPeekAtLastError() == cudaSuccess ? RPC_OK : RPC_ERR_CUDA;
}

RPC_API int rpc_momentum_apply_bwd(const float *g_out, const float *g_l2, const float *sp, const float *l2,
                                   int y_is_delta, float scaling, int64_t n, float *grad_x, float *grad_y, void *stream)
{
    if (n < 0) return RPC_ERR_CONFIG;
    if (n == 0) return RPC_OK;
    if (!sp || !l2 || (!grad_x && !grad_y)) return RPC_ERR_NULL;
    k_momentum_apply_bwd<<<stream_blocks(n), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(g_out, g_l2, sp, l2, y_is_delta, scaling, n, grad_x, grad_y);
    return cudaPeekAtLastError() == cudaSuccess ? RPC_OK : RPC_ERR_CUDA;
}

RPC_API size_t rpc_attention_scratch_bytes(void) { return sizeof(double) * kAttMaxParams; }

RPC_API int rpc_attention_gate_fwd(const float *xn, const float *t, const float *w1, const float *b1, const float *w2,
                                   const float *b2, int64_t K, int C, float *out, float *gate, void *stream)
{
    if (C < 1 || C > RPC_MAX_CHANNELS || K < 0) return RPC_ERR_CONFIG;
    if (K == 0) return RPC_OK;
    if (!xn || !t || !w1 || !b1 || !w2 || !b2 || !out) return RPC_ERR_NULL;
    if (C == 4 && (misaligned(xn) || misaligned(t) || misaligned(out))) return RPC_ERR_ALIGN;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int64_t blocks = (K + kThreads - 1) / kThreads;
    if (blocks > kNumSM * 8) blocks = kNumSM * 8;
    switch (C) {
#define RPC_CASE(CC) case CC: k_attention_gate_fwd<CC><<<(int)blocks, kThreads, 0, st>>>(xn, t, w1, b1, w2, b2, K, out, gate); break;
        RPC_CASE(1) RPC_CASE(2) RPC_CASE(3) RPC_CASE(4) RPC_CASE(5) RPC_CASE(6) RPC_CASE(7) RPC_CASE(8)
#undef RPC_CASE
        default: return RPC_ERR_CONFIG;
    }
    return cudaPeekAtLastError() == cudaSuccess ? RPC_OK : RPC_ERR_CUDA;
}

RPC_API int rpc_attention_gate_bwd(const float *g_out, const float *xn, const float *t, const float *w1, const float *b1,
                                   const float *w2, const float *b2, int64_t K, int C, double *scratch, float *grad_t,
                                   float *grad_xn, float *grad_w1, float *grad_b1, float *grad_w2, float *grad_b2,
                                   void *stream)
{
    if (C < 1 || C > RPC_MAX_CHANNELS || K < 0) return RPC_ERR_CONFIG;
    if (!scratch || !grad_w1 || !grad_b1 || !grad_w2 || !grad_b2) return RPC_ERR_NULL;
    if (K > 0 && (!g_out || !xn || !t || !w1 || !b1 || !w2 || !b2 || !grad_t)) return RPC_ERR_NULL;
    if (C == 4 && (misaligned(g_out) || misaligned(xn) || misaligned(t) || misaligned(grad_t) || misaligned(grad_xn)))
        return RPC_ERR_ALIGN;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (cudaMemsetAsync(scratch, 0, sizeof(double) * kAttMaxParams, st) != cudaSuccess) return RPC_ERR_CUDA;
    if (K > 0) {
        int64_t blocks = (K + kThreads - 1) / kThreads;
        if (blocks > kNumSM * 4) blocks = kNumSM * 4;
        switch (C) {
#define RPC_CASE(CC) case CC: k_attention_gate_bwd<CC><<<(int)blocks, kThreads, 0, st>>>(g_out, xn, t, w1, b1, w2, b2, K, grad_t, grad_xn, scratch); break;
            RPC_CASE(1) RPC_CASE(2) RPC_CASE(3) RPC_CASE(4) RPC_CASE(5) RPC_CASE(6) RPC_CASE(7) RPC_CASE(8)
#undef RPC_CASE
            default: return RPC_ERR_CONFIG;
        }
    }
    k_attention_gate_finalize<<<1, 64, 0, st>>>(scratch, C, grad_w1, grad_b1, grad_w2, grad_b2);
    return cudaPeekAtLastError() == cudaSuccess ? RPC_OK : RPC_ERR_CUDA;
}

RPC_API int rpc_linear_wgrad_bn(const float *g, const float *a, const float *bn_mean, const float *bn_invstd,
                                const float *bn_gamma, const float *bn_beta, int64_t K, int M, int N, float *grad_w,
                                float *grad_b, void *stream)
{
    if (bn_mean && (!bn_invstd || !bn_gamma || !bn_beta)) return RPC_ERR_NULL;
    const BnIn bn = {bn_mean, bn_invstd, bn_gamma, bn_beta};
    if (K < 0 || M < 1 || N < 1 || M > 128 || N > 128) return RPC_ERR_CONFIG;
    if (!grad_w) return RPC_ERR_NULL;
    if (K > 0 && (!g || !a)) return RPC_ERR_NULL;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (cudaMemsetAsync(grad_w, 0, sizeof(float) * M * N, st) != cudaSuccess) return RPC_ERR_CUDA;
    if (grad_b && cudaMemsetAsync(grad_b, 0, sizeof(float) * M, st) != cudaSuccess) return RPC_ERR_CUDA;
    if (K == 0) return RPC_OK;
    // 128 x 64 / 64 x 128 layers (the trunk of the KITTI config): 3 x TF32 on tcgen05, the result in one TMEM accumulator
    static const bool no_tc = getenv("RPC_WGRAD_SIMT") != nullptr;     // (A/B switch)
    if (!no_tc && ((M == 128 && N == 64) || (M == 64 && N == 128)) && K >= 4096 && !misaligned(g) && !misaligned(a)) {
        WgParams wp;
        memset(&wp, 0, sizeof(wp));
        wp.g = g; wp.a = a; wp.K = K; wp.M = M; wp.N = N;
        wp.bn_mean = bn_mean; wp.bn_invstd = bn_invstd; wp.bn_gamma = bn_gamma; wp.bn_beta = bn_beta;
        wp.gw = grad_w; wp.gb = grad_b;
        static bool attr_done[64] = {};
        int dev = 0;
        cudaGetDevice(&dev);
        if (dev >= 0 && dev < 64 && !attr_done[dev]) {
            if (cudaFuncSetAttribute(k_wgrad_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wg_smem_bytes()) != cudaSuccess)
                return RPC_ERR_CUDA;
            attr_done[dev] = true;
        }
        const int64_t stages = (K + kWgRows - 1) / kWgRows;
        const int sms = device_sms();
        k_wgrad_tc<<<(int)(stages < sms ? stages : sms), kWgThreads, wg_smem_bytes(), st>>>(wp);
        return cudaPeekAtLastError() == cudaSuccess ? RPC_OK : RPC_ERR_CUDA;
    }
    if (M <= 8 || N <= 8) {
        // skinny layer: a stream over the wide operand; the wider of the two is "wide" (ties: g)
        const bool wide_is_g = M >= N;
        if (wide_is_g && bn_mean) return RPC_ERR_CONFIG;                // (a prologue on the narrow operand is not needed by the trunk)
        const int W = wide_is_g ? M : N, S = wide_is_g ? N : M;
        const float *wide = wide_is_g ? g : a, *narrow = wide_is_g ? a : g;
        if (S == 4 && misaligned(narrow)) return RPC_ERR_ALIGN;
        const int groups = kSkinnyThreads / W;
        int64_t blocks = (K + groups - 1) / groups;
        if (blocks > (int64_t)device_sms() * 8) blocks = (int64_t)device_sms() * 8;
#define RPC_CASE(S_)                                                                                                   \
    case S_:                                                                                                           \
        if (wide_is_g) k_linear_wgrad_skinny<S_, true><<<(int)blocks, kSkinnyThreads, 0, st>>>(wide, narrow, K, W, grad_w, grad_b, bn);   \
        else k_linear_wgrad_skinny<S_, false><<<(int)blocks, kSkinnyThreads, 0, st>>>(wide, narrow, K, W, grad_w, grad_b, bn);            \
        break;
        switch (S) {
            RPC_CASE(1) RPC_CASE(2) RPC_CASE(3) RPC_CASE(4) RPC_CASE(5) RPC_CASE(6) RPC_CASE(7) RPC_CASE(8)
            default: return RPC_ERR_CONFIG;
        }
#undef RPC_CASE
        return cudaPeekAtLastError() == cudaSuccess ? RPC_OK : RPC_ERR_CUDA;
    }
    // every resident CTA slot of the device gets one equal chunk of rows (a multiple of the tile height)
    const int tm = M <= 16 ? 16 : M <= 64 ? 64 : 128, tn = N <= 16 ? 16 : N <= 64 ? 64 : 128;
#define RPC_CASE(TM_, TN_)                                                                                            \
    if (tm == TM_ && tn == TN_) {                                                                                     \
        int per_sm = 0;                                                                                               \
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_linear_wgrad<TM_, TN_>, LinTile<TM_, TN_>::NT, 0) != \
                cudaSuccess || per_sm < 1) { cudaGetLastError(); per_sm = 1; }                                        \
        int64_t ctas = (int64_t)device_sms() * per_sm;                                                                \
        const int64_t chunk = ((K + ctas - 1) / ctas + kLinKT - 1) / kLinKT * kLinKT;                                 \
        ctas = (K + chunk - 1) / chunk;                                                                               \
        k_linear_wgrad<TM_, TN_><<<(int)ctas, LinTile<TM_, TN_>::NT, 0, st>>>(g, a, K, M, N, chunk, grad_w, grad_b, bn);  \
    }
    RPC_CASE(16, 16) RPC_CASE(16, 64) RPC_CASE(16, 128) RPC_CASE(64, 16) RPC_CASE(64, 64) RPC_CASE(64, 128)
    RPC_CASE(128, 16) RPC_CASE(128, 64) RPC_CASE(128, 128)
#undef RPC_CASE
    return cudaPeekAtLastError() == cudaSuccess ? RPC_OK : RPC_ERR_CUDA;
}

RPC_API int rpc_linear_wgrad(const float *g, const float *a, int64_t K, int M, int N, float *grad_w, float *grad_b,
                             void *stream)
{
    return rpc_linear_wgrad_bn(g, a, nullptr, nullptr, nullptr, nullptr, K, M, N, grad_w, grad_b, stream);
}

// 16-byte columns per thread pay from ~32 M elements up (K = 835 k: 3.7 -> 4.2-4.6 TB/s); below that the kernels are
// occupancy / latency bound and the scalar mapping with its 8 CTAs per SM is faster (K = 131 k: 67 vs 91 us)
static bool bn_vec_pays(int64_t K, int H) { return H % 4 == 0 && K * (int64_t)H >= (int64_t(1) << 25); }

static int bn_blocks(int64_t K, int H, int V)
{
    const int groups = kBnThreads / (H / V);
    int64_t blocks = (K + (int64_t)groups * kBnU - 1) / ((int64_t)groups * kBnU);
    const int64_t cap = (int64_t)device_sms() * 8;
    return (int)(blocks < 1 ? 1 : blocks > cap ? cap : blocks);
}

RPC_API size_t rpc_bn_scratch_bytes(void) { return sizeof(double) * 2 * kBnMaxH; }

RPC_API int rpc_bn_relu_fwd(const float *y, int64_t K, int H, const float *gamma, const float *beta, float eps,
                            float momentum, float *running_mean, float *running_var, double *scratch,
                            float *save_mean, float *save_invstd, float *z, void *stream)
{
    if (K < 1 || H < 1 || H > kBnMaxH) return RPC_ERR_CONFIG;
    if (!y || !gamma || !beta || !scratch || !save_mean || !save_invstd || !z) return RPC_ERR_NULL;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (cudaMemsetAsync(scratch, 0, sizeof(double) * 2 * H, st) != cudaSuccess) return RPC_ERR_CUDA;
    const bool vec = bn_vec_pays(K, H) && !misaligned(y) && !misaligned(z);
    const int blocks = bn_blocks(K, H, vec ? 4 : 1);
    if (vec) {
        k_bn_moments<4><<<blocks, kBnThreads, 0, st>>>(y, K, H, scratch);
        k_bn_relu_fwd<4><<<blocks, kBnThreads, 0, st>>>(y, scratch, K, H, gamma, beta, eps, momentum, running_mean,
                                                        running_var, save_mean, save_invstd, z);
    } else {
        k_bn_moments<1><<<blocks, kBnThreads, 0, st>>>(y, K, H, scratch);
        k_bn_relu_fwd<1><<<blocks, kBnThreads, 0, st>>>(y, scratch, K, H, gamma, beta, eps, momentum, running_mean,
                                                        running_var, save_mean, save_invstd, z);
    }
    return cudaPeekAtLastError() == cudaSuccess ? RPC_OK : RPC_ERR_CUDA;
}

static int bn_relu_bwd_impl(const float *g, const float *y, const float *save_mean, const float *save_invstd,
                            const float *gamma, const float *beta, int64_t K, int H, double *scratch, bool have_sums,
                            float *grad_y, float *grad_gamma, float *grad_beta, void *stream)
{
    if (K < 1 || H < 1 || H > kBnMaxH) return RPC_ERR_CONFIG;
    if (!g || !y || !save_mean || !save_invstd || !gamma || !beta || !scratch || !grad_y) return RPC_ERR_NULL;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (!have_sums && cudaMemsetAsync(scratch, 0, sizeof(double) * 2 * H, st) != cudaSuccess) return RPC_ERR_CUDA;
    const bool vec = bn_vec_pays(K, H) && !misaligned(g) && !misaligned(y) && !misaligned(grad_y);
    const int blocks = bn_blocks(K, H, vec ? 4 : 1);
    if (vec) {
        if (!have_sums) k_bn_relu_bwd_reduce<4><<<blocks, kBnThreads, 0, st>>>(g, y, save_mean, save_invstd, gamma, beta, K, H, scratch);
        k_bn_relu_bwd_elemt<4><<<blocks, kBnThreads, 0, st>>>(g, y, save_mean, save_invstd, gamma, beta, scratch, K, H,
                                                              grad_y, grad_gamma, grad_beta);
    } else {
        if (!have_sums) k_bn_relu_bwd_reduce<1><<<blocks, kBnThreads, 0, st>>>(g, y, save_mean, save_invstd, gamma, beta, K, H, scratch);
        k_bn_relu_bwd_elemt<1><<<blocks, kBnThreads, 0, st>>>(g, y, save_mean, save_invstd, gamma, beta, scratch, K, H,
                                                              grad_y, grad_gamma, grad_beta);
    }
    return cudaPeekAtLastError() == cudaSuccess ? RPC_OK : RPC_ERR_CUDA;
}

RPC_API int rpc_bn_relu_bwd(const float *g, const float *y, const float *save_mean, const float *save_invstd,
                            const float *gamma, const float *beta, int64_t K, int H, double *scratch, float *grad_y,
                            float *grad_gamma, float *grad_beta, void *stream)
{
    return bn_relu_bwd_impl(g, y, save_mean, save_invstd, gamma, beta, K, H, scratch, false, grad_y, grad_gamma, grad_beta, stream);
}

RPC_API int rpc_bn_relu_bwd_sums(const float *g, const float *y, const float *save_mean, const float *save_invstd,
                                 const float *gamma, const float *beta, int64_t K, int H, const double *sums, float *grad_y,
                                 float *grad_gamma, float *grad_beta, void *stream)
{
    return bn_relu_bwd_impl(g, y, save_mean, save_invstd, gamma, beta, K, H, const_cast<double *>(sums), true, grad_y, grad_gamma,
                            grad_beta, stream);
}

}  // extern "C"

extern "C" {
// ------------------------------------------------------------------ Linear layers of the MLP trunk (tcgen05, 3 x TF32)
static int linear_fwd_impl(const float *x, int64_t K, const int32_t *k_dev, int NI, int NO, const float *w, int w_layout,
                           const float *bias, const float *in_mean, const float *in_invstd, const float *in_gamma,
                           const float *in_beta, int epi_tanh, double *moments, const float *red_v, const float *const *red_bn,
                           float *y, void *stream)
{
    if (K < 0 || NI < 1 || NO < 1 || NI > kTcMaxDim || NO > kTcMaxDim || (w_layout != 0 && w_layout != 1)) return RPC_ERR_CONFIG;
    if (K == 0) return RPC_OK;
    if (!x || !w || !y) return RPC_ERR_NULL;
    if (in_mean && (!in_invstd || !in_gamma || !in_beta)) return RPC_ERR_NULL;
    if (misaligned(x) || misaligned(y) || misaligned(w)) return RPC_ERR_ALIGN;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (moments && cudaMemsetAsync(moments, 0, sizeof(double) * 2 * NO, st) != cudaSuccess) return RPC_ERR_CUDA;
    TcParams p;
    memset(&p, 0, sizeof(p));
    p.x = x; p.w = w; p.bias = bias; p.y = y; p.k_dev = k_dev; p.K = K; p.NI = NI; p.NO = NO; p.w_layout = w_layout;
    p.in_mean = in_mean; p.in_invstd = in_invstd; p.in_gamma = in_gamma; p.in_beta = in_beta;
    p.epi_tanh = epi_tanh; p.moments = moments;
    if (red_v) {
        if (NI <= 8 || !moments || !red_bn) return RPC_ERR_CONFIG;      // only the tensor-core kernel has this epilogue
        p.red_v = red_v; p.red_mean = red_bn[0]; p.red_invstd = red_bn[1]; p.red_gamma = red_bn[2]; p.red_beta = red_bn[3];
    }
    { static const char *dbg = getenv("RPC_TC_DBG"); p.dbg = dbg ? atoi(dbg) : 0; }
    const int sms = device_sms();
    if (NI <= 8) {                                                   // first layer / input gradient of the last: a stream
        if (NO & 3) return RPC_ERR_CONFIG;
        int64_t blocks = (K * (NO >> 2) + kNarrowThreads - 1) / kNarrowThreads;
        if (blocks > (int64_t)sms * 8) blocks = (int64_t)sms * 8;
        k_linear_narrow_in<<<(int)blocks, kNarrowThreads, 0, st>>>(p);
    } else {                                                         // (the last layer, NO = C, also goes here: N padded to 32)
        if (NI & 7) return RPC_ERR_CONFIG;
        const int NOp = (NO + 31) & ~31;
        const size_t smem = tc_smem_bytes(NI, NOp);
        constexpr size_t kSmemMax = 227 * 1024;                        // per CTA on sm_100
        if (smem > kSmemMax) return RPC_ERR_CONFIG;                    // (128 -> 128: the weights alone take 128 KB)
        static bool attr_done[64] = {};
        int dev = 0;
        cudaGetDevice(&dev);
        if (dev >= 0 && dev < 64 && !attr_done[dev]) {
            if (cudaFuncSetAttribute(k_linear_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemMax) != cudaSuccess)
                return RPC_ERR_CUDA;
            attr_done[dev] = true;
        }
        int64_t tiles = (K + kTcRows - 1) / kTcRows;
        const int grid = (int)(tiles < sms ? tiles : sms);
        k_linear_tc<<<grid, kTcThreads2, smem, st>>>(p);
    }
    return cudaPeekAtLastError() == cudaSuccess ? RPC_OK : RPC_ERR_CUDA;
}

RPC_API int rpc_linear_fwd(const float *x, int64_t K, const int32_t *k_dev, int NI, int NO, const float *w, int w_layout,
                           const float *bias, const float *in_mean, const float *in_invstd, const float *in_gamma,
                           const float *in_beta, int epi_tanh, double *moments, float *y, void *stream)
{
    return linear_fwd_impl(x, K, k_dev, NI, NO, w, w_layout, bias, in_mean, in_invstd, in_gamma, in_beta, epi_tanh, moments,
                           nullptr, nullptr, y, stream);
}

RPC_API int rpc_linear_dgrad_bnred(const float *g, int64_t K, int NI, int NO, const float *w, const float *v,
                                   const float *v_mean, const float *v_invstd, const float *v_gamma, const float *v_beta,
                                   double *red, float *grad_in, void *stream)
{
    if (!v || !v_mean || !v_invstd || !v_gamma || !v_beta || !red) return RPC_ERR_NULL;
    const float *bn[4] = {v_mean, v_invstd, v_gamma, v_beta};
    return linear_fwd_impl(g, K, nullptr, NI, NO, w, 1, nullptr, nullptr, nullptr, nullptr, nullptr, 0, red, v, bn, grad_in, stream);
}

RPC_API int rpc_bn_finalize(const double *moments, int64_t K, const int32_t *k_dev, int H, float eps, float momentum,
                            float *running_mean, float *running_var, float *save_mean, float *save_invstd, void *stream)
{
    if (K < 1 || H < 1 || H > kTcMaxDim) return RPC_ERR_CONFIG;
    if (!moments || !save_mean || !save_invstd) return RPC_ERR_NULL;
    k_bn_finalize<<<1, kTcMaxDim, 0, static_cast<cudaStream_t>(stream)>>>(moments, K, k_dev, H, eps, momentum, running_mean,
                                                                      running_var, save_mean, save_invstd);
    return cudaPeekAtLastError() == cudaSuccess ? RPC_OK : RPC_ERR_CUDA;
}

RPC_API int rpc_bn_relu_apply(const float *y, int64_t K, const int32_t *k_dev, int H, const float *mean, const float *invstd,
                              const float *gamma, const float *beta, float *z, void *stream)
{
    if (K < 0 || H < 4 || (H & 3) || H > kTcMaxDim) return RPC_ERR_CONFIG;
    if (K == 0) return RPC_OK;
    if (!y || !mean || !invstd || !gamma || !beta || !z) return RPC_ERR_NULL;
    if (misaligned(y) || misaligned(z)) return RPC_ERR_ALIGN;
    int64_t blocks = (K * H / 4 + kNarrowThreads - 1) / kNarrowThreads;
    if (blocks > (int64_t)device_sms() * 8) blocks = (int64_t)device_sms() * 8;
    k_bn_relu_apply<<<(int)blocks, kNarrowThreads, 0, static_cast<cudaStream_t>(stream)>>>(y, K, k_dev, H, mean, invstd, gamma,
                                                                                       beta, z);
    return cudaPeekAtLastError() == cudaSuccess ? RPC_OK : RPC_ERR_CUDA;
}

}  // extern "C"

// ------------------------------------------------------------------ the whole trunk in one call each way
// The per-layer entry points above issued from Python cost ~15 us of interpreter time each; at the KITTI batch
// (K = 0.13 M rows) that was twice the kernels' own time.  rpc_trunk_fwd / rpc_trunk_bwd issue the same launches from C.
namespace {
__global__ void __launch_bounds__(256)
k_tanh_bwd(const float *__restrict__ g, const float *__restrict__ t, int64_t n, float *__restrict__ out)
{
    for (int64_t e = (int64_t)blockIdx.x * 256 + threadIdx.x; e < n; e += (int64_t)gridDim.x * 256) {
        const float tv = __ldg(t + e);
        out[e] = __fmul_rn(__ldg(g + e), __fsub_rn(1.0f, __fmul_rn(tv, tv)));          // ATen's tanh_backward
    }
}
int trunk_check(const rpc_trunk_desc *t)
{
    if (!t) return RPC_ERR_NULL;
    if (t->nl < 1 || t->nl > RPC_TRUNK_MAX_HIDDEN || t->C < 3 || t->C > RPC_MAX_CHANNELS) return RPC_ERR_CONFIG;
    for (int l = 0; l < t->nl; ++l) {
        if (t->width[l] < 8 || t->width[l] > kTcMaxDim || (t->width[l] & 7)) return RPC_ERR_CONFIG;
        if (!t->w[l] || !t->b[l] || !t->gamma[l] || !t->beta[l]) return RPC_ERR_NULL;
    }
    if (t->width[t->nl - 1] < 16) return RPC_ERR_CONFIG;
    if (!t->w[t->nl] || !t->b[t->nl]) return RPC_ERR_NULL;
    return RPC_OK;
}
}  // namespace

extern "C" {

RPC_API int rpc_trunk_fwd(const rpc_trunk_desc *t, const float *xn, int64_t K, float *const *y, float *stats, double *scratch,
                          float *raw, void *stream)
{
    int rc = trunk_check(t);
    if (rc != RPC_OK) return rc;
    if (K < 2) return RPC_ERR_CONFIG;
    if (!xn || !y || !stats || !scratch || !raw) return RPC_ERR_NULL;
    const float *x = xn;
    int ni = t->C;
    const float *mean = nullptr, *invstd = nullptr, *ga = nullptr, *be = nullptr;
    float *st_l = stats;
    for (int l = 0; l < t->nl; ++l) {
        const int no = t->width[l];
        if (!y[l]) return RPC_ERR_NULL;
        rc = rpc_linear_fwd(x, K, nullptr, ni, no, t->w[l], 0, t->b[l], mean, invstd, ga, be, 0, scratch, y[l], stream);
        if (rc != RPC_OK) return rc;
        rc = rpc_bn_finalize(scratch, K, nullptr, no, t->eps[l], t->momentum[l], t->running_mean[l], t->running_var[l], st_l,
                             st_l + no, stream);
        if (rc != RPC_OK) return rc;
        mean = st_l; invstd = st_l + no; ga = t->gamma[l]; be = t->beta[l];
        st_l += 2 * no;
        scratch += 2 * no;                                         // (a fresh accumulator per layer: no ordering assumptions)
        x = y[l]; ni = no;
    }
    return rpc_linear_fwd(x, K, nullptr, ni, t->C, t->w[t->nl], 0, t->b[t->nl], mean, invstd, ga, be, 1, nullptr, raw, stream);
}

RPC_API int rpc_trunk_fwd_eval(const rpc_trunk_desc *t, const float *xn, int64_t K, const float *stats, float *work, float *raw,
                               void *stream)
{
    // eval mode: BatchNorm normalises with the running statistics the caller turned into (mean, invstd) per layer (`stats`,
    // laid out like rpc_trunk_fwd's); nothing is kept, two [K, max width] areas of `work` are used in turn
    int rc = trunk_check(t);
    if (rc != RPC_OK) return rc;
    if (K < 1) return RPC_ERR_CONFIG;
    if (!xn || !stats || !work || !raw) return RPC_ERR_NULL;
    int maxw = t->C;
    for (int l = 0; l < t->nl; ++l) maxw = t->width[l] > maxw ? t->width[l] : maxw;
    const float *x = xn, *st_l = stats;
    int ni = t->C;
    const float *mean = nullptr, *invstd = nullptr, *ga = nullptr, *be = nullptr;
    for (int l = 0; l < t->nl; ++l) {
        const int no = t->width[l];
        float *y = work + (size_t)(l & 1) * K * maxw;
        rc = rpc_linear_fwd(x, K, nullptr, ni, no, t->w[l], 0, t->b[l], mean, invstd, ga, be, 0, nullptr, y, stream);
        if (rc != RPC_OK) return rc;
        mean = st_l; invstd = st_l + no; ga = t->gamma[l]; be = t->beta[l];
        st_l += 2 * no;
        x = y; ni = no;
    }
    return rpc_linear_fwd(x, K, nullptr, ni, t->C, t->w[t->nl], 0, t->b[t->nl], mean, invstd, ga, be, 1, nullptr, raw, stream);
}

RPC_API int rpc_trunk_bwd(const rpc_trunk_desc *t, const float *xn, int64_t K, float *const *y, const float *stats,
                          const float *raw, const float *g_raw, float *work, double *scratch, float *const *grad_w,
                          float *const *grad_b, float *const *grad_gamma, float *const *grad_beta, float *grad_xn, void *stream)
{
    int rc = trunk_check(t);
    if (rc != RPC_OK) return rc;
    if (K < 2) return RPC_ERR_CONFIG;
    if (!xn || !y || !stats || !raw || !g_raw || !work || !scratch || !grad_w || !grad_b || !grad_gamma || !grad_beta) return RPC_ERR_NULL;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int nl = t->nl, C = t->C;
    int maxw = C;
    const float *st_of[RPC_TRUNK_MAX_HIDDEN];
    {
        const float *s = stats;
        for (int l = 0; l < nl; ++l) { st_of[l] = s; s += 2 * t->width[l]; if (t->width[l] > maxw) maxw = t->width[l]; }
    }
    float *g_a = work, *g_y = work + (size_t)K * maxw;                                      // two [K, maxw] areas
    // Tanh, then the output layer: weight gradient against a = relu(bn(y[nl-1])), input gradient
    {
        int64_t n = K * C, blocks = (n + 255) / 256;
        if (blocks > (int64_t)device_sms() * 8) blocks = (int64_t)device_sms() * 8;
        k_tanh_bwd<<<(int)blocks, 256, 0, st>>>(g_raw, raw, n, g_y);
        if (cudaPeekAtLastError() != cudaSuccess) return RPC_ERR_CUDA;
    }
    const int hl = t->width[nl - 1];
    rc = rpc_linear_wgrad_bn(g_y, y[nl - 1], st_of[nl - 1], st_of[nl - 1] + hl, t->gamma[nl - 1], t->beta[nl - 1], K, C, hl,
                             grad_w[nl], grad_b[nl], stream);
    if (rc != RPC_OK) return rc;
    rc = rpc_linear_fwd(g_y, K, nullptr, C, hl, t->w[nl], 1, nullptr, nullptr, nullptr, nullptr, nullptr, 0, nullptr, g_a, stream);
    if (rc != RPC_OK) return rc;
    bool have_sums = false;                                        // the producer of g_a left this layer's BatchNorm reductions in `scratch`
    for (int l = nl - 1; l >= 0; --l) {
        const int h = t->width[l], hin = l > 0 ? t->width[l - 1] : C;
        rc = have_sums ? rpc_bn_relu_bwd_sums(g_a, y[l], st_of[l], st_of[l] + h, t->gamma[l], t->beta[l], K, h, scratch, g_y,
                                              grad_gamma[l], grad_beta[l], stream)
                       : rpc_bn_relu_bwd(g_a, y[l], st_of[l], st_of[l] + h, t->gamma[l], t->beta[l], K, h, scratch, g_y,
                                         grad_gamma[l], grad_beta[l], stream);
        if (rc != RPC_OK) return rc;
        have_sums = false;
        // (the layer input relu(bn(y[l-1])) is recomputed while the weight-gradient kernel stages it)
        if (l > 0)
            rc = rpc_linear_wgrad_bn(g_y, y[l - 1], st_of[l - 1], st_of[l - 1] + hin, t->gamma[l - 1], t->beta[l - 1], K, h, hin,
                                     grad_w[l], grad_b[l], stream);
        else
            rc = rpc_linear_wgrad(g_y, xn, K, h, C, grad_w[l], grad_b[l], stream);
        if (rc != RPC_OK) return rc;
        // (measured: with the reductions in the epilogue the nuScenes backward is 0.55 ms SLOWER -- the epilogue is what
        // bounds the tensor-core kernel, and the separate reduction kernel streams at 4.5 TB/s -- so it is opt-in)
        static const bool bnred = getenv("RPC_BNRED") != nullptr;
        if (l > 0 && h > 8 && bnred) {
            // input gradient on the tensor cores; its epilogue also reduces the BatchNorm backward of layer l - 1 (sum g',
            // sum g' * xhat), so that layer's rpc_bn_relu_bwd skips its reduction pass.  (`scratch` is free again: the
            // elementwise kernel of layer l above has been enqueued and everything here is stream-ordered.)
            rc = rpc_linear_dgrad_bnred(g_y, K, h, hin, t->w[l], y[l - 1], st_of[l - 1], st_of[l - 1] + hin, t->gamma[l - 1],
                                        t->beta[l - 1], scratch, g_a, stream);
            if (rc != RPC_OK) return rc;
            have_sums = true;
        } else {
            float *dst = l > 0 ? g_a : grad_xn;
            if (dst) {
                rc = rpc_linear_fwd(g_y, K, nullptr, h, hin, t->w[l], 1, nullptr, nullptr, nullptr, nullptr, nullptr, 0, nullptr, dst, stream);
                if (rc != RPC_OK) return rc;
            }
        }
    }
    return RPC_OK;
}

}  // extern "C"

extern "C" {
// experiment counters of k_linear_tc (RPC_TC_DBG=4; tools only, not in the header)
RPC_API int rpc_debug_tc_stats(unsigned long long *out16)
{
    return cudaMemcpyFromSymbol(out16, g_tc_stats, sizeof(unsigned long long) * 16) == cudaSuccess ? RPC_OK : RPC_ERR_CUDA;
}
}  // extern "C"

// ------------------------------------------------------------------ allocation-free variant (static device scratch)
// rpc_points_range_filter_static: rpc_points_range_filter with its per-tile count scratch in a file-scope device
// array, for callers that cannot provide a workspace (CUDA-graph capture, tight serving loops, harnesses that
// forbid cudaMalloc).  No counterpart in the reference.
// Every GPU that loads the library holds the array (kStaticFilterInts ints = 16 640 bytes) until exit.
// All calls on a device share it: the kernel writes every count it later reads, so nothing has to start at zero,
// but two calls that overlap on the GPU (different streams) race on it.
namespace {
constexpr int kStaticFilterTiles = (int)(RPC_STATIC_FILTER_MAX_POINTS / kRowTile);
static_assert((int64_t)kStaticFilterTiles * kRowTile == RPC_STATIC_FILTER_MAX_POINTS, "limit is a whole number of tiles");
static_assert(kStaticFilterTiles <= kMaxTiles, "static scratch must fit the compaction kernel's scan");
constexpr int kStaticFilterInts = (kStaticFilterTiles + 1 + 63) / 64 * 64;      // what rpc_valid_rows_workspace_bytes rounds to
static __device__ __align__(256) int g_static_filter_cnt[kStaticFilterInts];

// In host code `g_static_filter_cnt` is only a stand-in for the symbol: its device address is looked up once per
// device, with that device current, and handed to the kernel as an ordinary pointer.
int static_filter_scratch(int **out)
{
    static int *cache[64] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return RPC_ERR_CUDA;
    if (!cache[dev]) {
        void *ptr = nullptr;
        if (cudaGetSymbolAddress(&ptr, g_static_filter_cnt) != cudaSuccess || !ptr) {
            cudaGetLastError();
            return RPC_ERR_CUDA;
        }
        cache[dev] = static_cast<int *>(ptr);
    }
    *out = cache[dev];
    return RPC_OK;
}
}  // namespace

extern "C" {

RPC_API int rpc_points_range_filter_static(const float *points, int64_t N, int C, const float *range6,
                                           float *out_points, int32_t *out_index, int32_t *count, void *stream)
{
    if (C < 3 || N < 0) return RPC_ERR_CONFIG;
    if (N > RPC_STATIC_FILTER_MAX_POINTS) return RPC_ERR_STATIC_LIMIT;
    if (!count || !range6) return RPC_ERR_NULL;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (N == 0) return cudaMemsetAsync(count, 0, 4, st) == cudaSuccess ? RPC_OK : RPC_ERR_CUDA;
    if (!points || !out_index) return RPC_ERR_NULL;
    int *scratch = nullptr;
    const int rc = static_filter_scratch(&scratch);
    if (rc != RPC_OK) return rc;
    RangeDesc rd;
    for (int d = 0; d < 3; ++d) { rd.lo[d] = range6[d]; rd.hi[d] = range6[3 + d]; }
    return launch_valid_rows(points, nullptr, N, 1, C, 2, nullptr, out_index, nullptr, out_points, count, scratch,
                             sizeof(int) * (size_t)kStaticFilterInts, rd, st);
}

}  // extern "C"

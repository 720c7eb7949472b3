// dyn_scatter.cuh -- dynamic voxelization + DynamicScatter(mean) of a batch of sweeps as ONE persistent
// cooperative kernel, sm_100a (BASELINE.json config 4; SURVEY 8a-U "DynamicScatter(mean) semantics").
//
// mmcv's DynamicScatter orders the voxels of a sample by the lexicographically sorted unique (z, y, x)
// rows (torch.unique_dim) and mmdet3d concatenates the samples in batch order.  Sorted (z, y, x) is the
// ascending linear cell id (z*gy + y)*gx + x, so no sort is needed: occupancy is recorded in a two-level
// bitmap over the global cell id  b * Gp + cell  (Gp = cells per sample rounded up to 8192) and a voxel's
// row is the number of occupied cells below it:
//
//   coarse level  one bit per GROUP of 256 consecutive cells            (B*Gp/8192 words; ~160 KB at nuScenes sizes)
//   fine level    256 bits (32 bytes, one L2 sector) per OCCUPIED group, allocated by the coarse scan
//
//   D0 clear    coarse words := 0
//   D1 mark     per point: [perturb] -> cell -> coors row out; cell id kept; coarse bit set
//   D2 coarse   chunked popcount scan of the coarse words -> slot of every occupied group; fine := 0 (2 sub-phases)
//   D3 fine     per point: slot of its group (coarse word + prefix), fine bit set
//   D4 rank     chunked popcount scan of the fine words -> voxel rows before every group; accumulators := 0
//   D5 sum      per point: voxel row = group prefix + bits below -> point2voxel, voxel (b, z, y, x);
//               accumulator row {F sums, count} += {point, 1} with 16-byte fp32 red.add (order-free, like
//               mmcv's own CUDA reduce; 2 atomics per point instead of F + 1)
//   D6 mean     feats = sums / count, num_points = count
//
// Measured on B200 (profiles/r01_*dynamic*): the bitmap phases were bound by a handful of L2 lines -- a
// LiDAR sweep puts most ground points into a few hundred consecutive coarse words -- so (1) coarse words
// live at a multiplicatively permuted index (neighbouring words land on different lines and L2 slices),
// (2) the lanes of a warp that share a word elect one lane to test-and-set the OR of their bits, and
// (3) once a level is final it is read through L1.  Every zero-fill and every scan is split evenly over
// the CTAs (chunked emission left the dense chunks as stragglers at the grid barriers).
//
// The backward is k_step_bwd with point2voxel := this map and num_points := the counts.
#pragma once
#include <cooperative_groups.h>

#include "common.cuh"
#include "hard_fwd.cuh"

namespace rpc {

constexpr int kDynThreads = 1024;
constexpr int kDynCtasPerSM = 2;
constexpr int kGroupCells = 256;                  // cells per fine group (8 words, 32 bytes)
constexpr int kCoarseCells = kGroupCells * 32;    // cells per coarse word

struct DynParams {
    const float *pts, *raw;                       // [ntotal, C]
    const int *coors_in;                          // [ntotal, 4] (b, z, y, x) to scatter by, or NULL: computed
    int *coors_out;                               // [ntotal, 4] or NULL
    int *p2v;                                     // [ntotal] voxel row of each point, -1 dropped
    float *feats;                                 // [cap, F]
    int *vcoors;                                  // [cap, 4]
    int *count;                                   // [cap]
    int *voxel_num;                               // [B + 1]
    uint2 *coarse;                                // [W1p] {bits, occupied groups in the words before}, permuted index
    uint32_t *fine;                               // [A_cap * 8]
    uint32_t *gpre;                               // [A_cap] voxel rows before the group
    int *cell;                                    // [ntotal] linear cell within the sample, -1 dropped
    int *pslot;                                   // [ntotal] sample index (D1), then slot of the point's group (D3)
    uint32_t *part;                               // [2 * gridDim] chunk sums of the two scans
    int *ssb;                                     // [B + 1] first slot of each sample
    float4 *acc;                                  // [cap * acc_vec] per voxel: F sums, count, padding
    unsigned long long *stamps;                   // [gridDim][32] %globaltimer per CTA at the phase boundaries, or NULL
    long long Gp;                                 // cells per sample, padded to kCoarseCells
    int W1;                                       // coarse words
    uint32_t w1_mask;                             // W1p - 1 (W1p = power of two >= W1)
    int ntotal;
    int acc_vec;                                  // float4 per accumulator row = (F + 1 + 3) / 4
    uint32_t mg_gxy, sh_gxy, mg_gx, sh_gx;        // n / (gx*gy) and n / gx as (umulhi(n, mg) + n) >> sh
    GridDesc gd;
    BatchDesc bd;
    PerturbDesc pd;
};

// coarse word w lives at a permuted position: consecutive words go to different lines and L2 slices
__device__ __forceinline__ uint32_t coarse_pos(const DynParams &p, uint32_t w) { return (w * 0x9E3779B1u) & p.w1_mask; }
__device__ __forceinline__ uint32_t fast_div(uint32_t n, uint32_t mg, uint32_t sh) { return (__umulhi(n, mg) + n) >> sh; }


// exclusive prefix of `part[0 .. nct)` at index blockIdx.x and the grand total (all CTAs compute it)
__device__ __forceinline__ uint32_t chunk_base(const uint32_t *part, int nct, int *sm, uint32_t *total)
{
    __shared__ int s_mine;
    int acc = 0;                                                   // uniform over the CTA
    for (int j0 = 0; j0 < nct; j0 += kDynThreads) {
        const int j = j0 + threadIdx.x;
        const int v = j < nct ? (int)__ldcg(part + j) : 0;
        int tot;
        const int e = block_excl_scan<kDynThreads>(v, sm, &tot);
        if (j == blockIdx.x) s_mine = acc + e;
        acc += tot;
    }
    __syncthreads();
    *total = (uint32_t)acc;
    return (uint32_t)s_mine;
}

__device__ __forceinline__ void dyn_stamp(const DynParams &p, int k)
{
    // every CTA stamps (32 slots each: k = phase start after the barrier, 16 + k = arrival at the next barrier)
    if (p.stamps && threadIdx.x == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
        p.stamps[blockIdx.x * 32 + k] = t;
    }
}

// Rows of 32 consecutive points are pulled in with C fully coalesced 128-byte loads per warp and handed to
// the lanes through shared memory (a 20-byte row defeats vector loads; C == 4 uses one float4 per lane).
template <int C>
__device__ __forceinline__ void warp_stage_rows(const float *__restrict__ base, long long row0, long long nrows,
                                                float *sbuf)
{
    const int lane = threadIdx.x & 31;
    const long long e0 = row0 * C, e1 = min(row0 + 32, nrows) * C;
    __syncwarp();
#pragma unroll
    for (int k = 0; k < C; ++k) {
        const long long e = e0 + k * 32 + lane;
        if (e < e1) sbuf[k * 32 + lane] = __ldg(base + e);
    }
    __syncwarp();
}

// the (perturbed) features of row i into x[0..C): staged for C != 4
template <int C, bool PERTURB>
__device__ __forceinline__ void dyn_load_point(const DynParams &p, long long i0, bool valid, float *swarp, float x[C])
{
    const int lane = threadIdx.x & 31;
    if constexpr (C == 4) {
        if (valid) {
            const float4 v = __ldg(reinterpret_cast<const float4 *>(p.pts) + i0 + lane);
            x[0] = v.x; x[1] = v.y; x[2] = v.z; x[3] = v.w;
            if constexpr (PERTURB) {
                const float4 d = __ldg(reinterpret_cast<const float4 *>(p.raw) + i0 + lane);
                const float dv[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
                for (int c = 0; c < 4; ++c) x[c] = __fadd_rn(x[c], perturb_delta(dv[c], p.pd.eb[c], p.pd.fb[c]));
            }
        }
    } else {
        warp_stage_rows<C>(p.pts, i0, p.ntotal, swarp);
        if constexpr (PERTURB) warp_stage_rows<C>(p.raw, i0, p.ntotal, swarp + 32 * C);
        if (valid) {
#pragma unroll
            for (int c = 0; c < C; ++c) {
                x[c] = swarp[lane * C + c];
                if constexpr (PERTURB)
                    x[c] = __fadd_rn(x[c], perturb_delta(swarp[32 * C + lane * C + c], p.pd.eb[c], p.pd.fb[c]));
            }
        }
    }
}

template <int C, bool PERTURB>
__global__ void __launch_bounds__(kDynThreads, kDynCtasPerSM)
k_dyn_scatter(const __grid_constant__ DynParams p)
{
    cg::grid_group grid = cg::this_grid();
    extern __shared__ __align__(16) float dyn_stage[];             // [32 warps][(PERTURB ? 2 : 1) * 32 * C], C != 4
    __shared__ int sm[33];
    const int tid = threadIdx.x, nct = gridDim.x, lane = tid & 31;
    const long long gstride = (long long)nct * kDynThreads;
    const long long gtid = (long long)blockIdx.x * kDynThreads + tid;
    const int F = p.gd.F;
    const int S4 = p.acc_vec;                                      // float4 per accumulator row
    float *swarp = dyn_stage + (tid >> 5) * ((PERTURB ? 2 : 1) * 32 * C);

    // ------------------------------------------------------------------------------- D0 clear
    dyn_stamp(p, 0);
    for (long long w = gtid; w <= p.w1_mask; w += gstride) p.coarse[w] = make_uint2(0u, 0u);
    if (p.stamps) __syncthreads();
    dyn_stamp(p, 16);
    grid.sync();
    dyn_stamp(p, 1);

    // ------------------------------------------------------------------------------- D1 mark
    for (long long i0 = gtid - lane; i0 < p.ntotal; i0 += gstride) {
        const long long i = i0 + lane;
        const bool valid = i < p.ntotal;
        int b = 0, cell = -1;
        if (p.coors_in) {
            if (valid) {
                const int4 c = __ldg(reinterpret_cast<const int4 *>(p.coors_in) + i);
                const bool ok = c.x >= 0 && c.x < p.bd.B && c.y >= 0 && c.y < p.gd.g[2] && c.z >= 0 &&
                                c.z < p.gd.g[1] && c.w >= 0 && c.w < p.gd.g[0];
                if (ok) { b = c.x; cell = (c.y * p.gd.g[1] + c.z) * p.gd.g[0] + c.w; }
            }
        } else {
            float x[C];
            dyn_load_point<C, PERTURB>(p, i0, valid, swarp, x);
            if (valid) {
                const int cx = axis_cell(x[0], p.gd, 0);
                const int cy = axis_cell(x[1], p.gd, 1);
                const int cz = axis_cell(x[2], p.gd, 2);
#pragma unroll 1
                while (b + 1 < p.bd.B && i >= p.bd.pt_begin[b + 1]) ++b;
                const bool ok = (cx | cy | cz) >= 0;
                if (ok) cell = (cz * p.gd.g[1] + cy) * p.gd.g[0] + cx;
                if (p.coors_out)
                    reinterpret_cast<int4 *>(p.coors_out)[i] = ok ? make_int4(b, cz, cy, cx) : make_int4(b, -1, -1, -1);
            }
        }
        if (valid) {
            p.cell[i] = cell;
            p.pslot[i] = b;
            if (cell < 0) p.p2v[i] = -1;
        }
        const bool has = cell >= 0;
        const unsigned act = __ballot_sync(0xFFFFFFFFu, has);
        if (has) {
            const long long grp = ((long long)b * p.Gp + cell) >> 8;
            const uint32_t widx = (uint32_t)(grp >> 5);
            const unsigned peers = __match_any_sync(act, widx);
            const uint32_t bits = __reduce_or_sync(peers, 1u << (grp & 31));
            if (lane == __ffs(peers) - 1) {
                uint32_t *w = &p.coarse[coarse_pos(p, widx)].x;
                if ((__ldcg(w) & bits) != bits) atomicOr(w, bits);
            }
        }
    }
    if (p.stamps) __syncthreads();
    dyn_stamp(p, 17);
    grid.sync();
    dyn_stamp(p, 2);

    // ------------------------------------------------------------------------------- D2 coarse scan
    const int cw = (p.W1 + nct - 1) / nct;                        // coarse words per CTA
    const int w_begin = min(blockIdx.x * cw, p.W1), w_end = min(w_begin + cw, p.W1);
    {
        int s = 0;
        for (int w = w_begin + tid; w < w_end; w += kDynThreads) s += __popc(__ldcg(&p.coarse[coarse_pos(p, w)].x));
        int tot;
        block_excl_scan<kDynThreads>(s, sm, &tot);
        if (tid == 0) p.part[blockIdx.x] = (uint32_t)tot;
    }
    if (p.stamps) __syncthreads();
    dyn_stamp(p, 18);
    grid.sync();
    dyn_stamp(p, 3);
    uint32_t A;                                                    // occupied groups in the batch
    {
        uint32_t run = chunk_base(p.part, nct, sm, &A);
        const int wpb = (int)(p.Gp / kCoarseCells);                // coarse words per sample
        for (int w0 = w_begin; w0 < w_end; w0 += kDynThreads) {
            const int w = w0 + tid;
            const uint32_t pos = coarse_pos(p, (uint32_t)w);
            const uint32_t bits = w < w_end ? __ldcg(&p.coarse[pos].x) : 0u;
            int tot;
            const uint32_t e = run + (uint32_t)block_excl_scan<kDynThreads>(__popc(bits), sm, &tot);
            run += (uint32_t)tot;
            if (w < w_end) {
                p.coarse[pos].y = e;
                if (w % wpb == 0) p.ssb[w / wpb] = (int)e;
            }
        }
        if (blockIdx.x == 0 && tid == 0) p.ssb[p.bd.B] = (int)A;
        uint4 *f4 = reinterpret_cast<uint4 *>(p.fine);
        for (long long t = gtid; t < 2ll * A; t += gstride) f4[t] = make_uint4(0u, 0u, 0u, 0u);
    }
    if (p.stamps) __syncthreads();
    dyn_stamp(p, 19);
    grid.sync();
    dyn_stamp(p, 4);

    // ------------------------------------------------------------------------------- D3 fine bits
    // the coarse level is final: read through L1 (no L1 copy can predate the barrier: D1/D2 used ld.cg)
    for (long long i0 = gtid - lane; i0 < p.ntotal; i0 += gstride) {
        const long long i = i0 + lane;
        const int cell = i < p.ntotal ? __ldcg(p.cell + i) : -1;
        const bool has = cell >= 0;
        const unsigned act = __ballot_sync(0xFFFFFFFFu, has);
        if (has) {
            const int b = __ldcg(p.pslot + i);
            const long long grp = ((long long)b * p.Gp + cell) >> 8;
            const uint2 cwv = p.coarse[coarse_pos(p, (uint32_t)(grp >> 5))];
            const uint32_t slot = cwv.y + __popc(cwv.x & ((1u << (grp & 31)) - 1u));
            p.pslot[i] = (int)slot;
            const int sub = cell & (kGroupCells - 1);              // Gp is a multiple of 256
            const uint32_t fw = slot * 8 + (sub >> 5);
            const unsigned peers = __match_any_sync(act, fw);
            const uint32_t bits = __reduce_or_sync(peers, 1u << (sub & 31));
            if (lane == __ffs(peers) - 1) {
                uint32_t *w = p.fine + fw;
                if ((__ldcg(w) & bits) != bits) atomicOr(w, bits);
            }
        }
    }
    if (p.stamps) __syncthreads();
    dyn_stamp(p, 20);
    grid.sync();
    dyn_stamp(p, 5);

    // ------------------------------------------------------------------------------- D4 rank
    const uint32_t cgp = (A + nct - 1) / nct;                      // groups per CTA
    const uint32_t g_begin = min((uint32_t)blockIdx.x * cgp, A), g_end = min(g_begin + cgp, A);
    {
        int s = 0;
        for (uint32_t g = g_begin + tid; g < g_end; g += kDynThreads) {
            const uint4 *f = reinterpret_cast<const uint4 *>(p.fine + (size_t)g * 8);
            const uint4 a = __ldcg(f), c = __ldcg(f + 1);
            s += __popc(a.x) + __popc(a.y) + __popc(a.z) + __popc(a.w) + __popc(c.x) + __popc(c.y) + __popc(c.z) + __popc(c.w);
        }
        int tot;
        block_excl_scan<kDynThreads>(s, sm, &tot);
        if (tid == 0) p.part[nct + blockIdx.x] = (uint32_t)tot;
    }
    if (p.stamps) __syncthreads();
    dyn_stamp(p, 21);
    grid.sync();
    dyn_stamp(p, 6);
    uint32_t M;
    {
        uint32_t run = chunk_base(p.part + nct, nct, sm, &M);
        for (uint32_t g0 = g_begin; g0 < g_end; g0 += kDynThreads) {
            const uint32_t g = g0 + tid;
            int n = 0;
            if (g < g_end) {
                const uint4 *f = reinterpret_cast<const uint4 *>(p.fine + (size_t)g * 8);
                const uint4 a = __ldcg(f), c = __ldcg(f + 1);
                n = __popc(a.x) + __popc(a.y) + __popc(a.z) + __popc(a.w) + __popc(c.x) + __popc(c.y) + __popc(c.z) + __popc(c.w);
            }
            int tot;
            const uint32_t e = run + (uint32_t)block_excl_scan<kDynThreads>(n, sm, &tot);
            run += (uint32_t)tot;
            if (g < g_end) p.gpre[g] = e;
        }
        for (long long t = gtid; t < (long long)M * S4; t += gstride) p.acc[t] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    if (p.stamps) __syncthreads();
    dyn_stamp(p, 22);
    grid.sync();
    dyn_stamp(p, 7);

    // ------------------------------------------------------------------------------- D5 sum
    if (blockIdx.x == 0 && tid <= p.bd.B) {
        // voxels before sample b = voxel rows before its first slot
        auto start = [&](int b) -> int {
            const int s = __ldcg(p.ssb + b);
            return (uint32_t)s < A ? (int)__ldcg(p.gpre + s) : (int)M;
        };
        p.voxel_num[tid] = tid < p.bd.B ? start(tid + 1) - start(tid) : (int)M;
    }
    for (long long i0 = gtid - lane; i0 < p.ntotal; i0 += gstride) {
        const long long i = i0 + lane;
        const bool valid = i < p.ntotal;
        float x[C];
        dyn_load_point<C, PERTURB>(p, i0, valid, swarp, x);
        const int cell = valid ? __ldcg(p.cell + i) : -1;
        if (cell < 0) continue;
        const uint32_t slot = (uint32_t)__ldcg(p.pslot + i);
        const int sub = cell & (kGroupCells - 1);
        // fine level and group prefixes are final: through L1
        const uint4 *f = reinterpret_cast<const uint4 *>(p.fine + (size_t)slot * 8);
        const uint4 a = f[0], c4 = f[1];
        uint32_t m = p.gpre[slot];
        const int wq = sub >> 5;
        const uint32_t below = (1u << (sub & 31)) - 1u;
        m += __popc(wq > 0 ? a.x : a.x & below);
        if (wq >= 1) m += __popc(wq > 1 ? a.y : a.y & below);
        if (wq >= 2) m += __popc(wq > 2 ? a.z : a.z & below);
        if (wq >= 3) m += __popc(wq > 3 ? a.w : a.w & below);
        if (wq >= 4) m += __popc(wq > 4 ? c4.x : c4.x & below);
        if (wq >= 5) m += __popc(wq > 5 ? c4.y : c4.y & below);
        if (wq >= 6) m += __popc(wq > 6 ? c4.z : c4.z & below);
        if (wq >= 7) m += __popc(c4.w & below);
        p.p2v[i] = (int)m;
        // every point of a voxel writes the same (b, z, y, x)
        if (p.coors_in) {
            reinterpret_cast<int4 *>(p.vcoors)[m] = __ldg(reinterpret_cast<const int4 *>(p.coors_in) + i);
        } else {
            int b = 0;
#pragma unroll 1
            while (b + 1 < p.bd.B && i >= p.bd.pt_begin[b + 1]) ++b;
            const uint32_t z = fast_div((uint32_t)cell, p.mg_gxy, p.sh_gxy);
            const uint32_t r = (uint32_t)cell - z * (uint32_t)(p.gd.g[0] * p.gd.g[1]);
            const uint32_t y = fast_div(r, p.mg_gx, p.sh_gx);
            reinterpret_cast<int4 *>(p.vcoors)[m] = make_int4(b, (int)z, (int)y, (int)(r - y * (uint32_t)p.gd.g[0]));
        }
        // accumulator row: F feature sums, then the point count (exact in fp32 below 2^24), padded to float4s
        float v[12];
#pragma unroll
        for (int c = 0; c < 12; ++c) v[c] = c < C && c < F ? x[c < C ? c : 0] : 0.0f;
#pragma unroll
        for (int c = 0; c < 9; ++c) if (c == F) v[c] = 1.0f;
        float4 *row = p.acc + (size_t)m * S4;
#pragma unroll
        for (int q = 0; q < 3; ++q)
            if (q < S4) atomicAdd(row + q, make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]));
    }
    if (p.stamps) __syncthreads();
    dyn_stamp(p, 23);
    grid.sync();
    dyn_stamp(p, 8);

    // ------------------------------------------------------------------------------- D6 mean
    for (long long m = gtid; m < (long long)M; m += gstride) {
        float v[12];
#pragma unroll
        for (int q = 0; q < 3; ++q)
            if (q < S4) {
                const float4 t = __ldcg(p.acc + m * S4 + q);
                v[4 * q] = t.x; v[4 * q + 1] = t.y; v[4 * q + 2] = t.z; v[4 * q + 3] = t.w;
            }
        float n = 1.0f;
#pragma unroll
        for (int c = 0; c < 9; ++c) if (c == F) n = v[c];
        p.count[m] = (int)n;
#pragma unroll
        for (int c = 0; c < 8; ++c) if (c < F) p.feats[m * F + c] = __fdiv_rn(v[c], n);
    }
    if (p.stamps) __syncthreads();
    dyn_stamp(p, 9);
}

}  // namespace rpc

// hard_fwd.cuh -- deterministic hard voxelization of a batch of sweeps (+ optional perturbation and
// HardSimpleVFE mean) as ONE persistent cooperative kernel, sm_100a.
//
// Replaces mmcv `hard_voxelize_forward` with a parallel formulation that gives the SAME
// result as the serial CPU walk.  At the named sizes the path moves ~30-200 MB, i.e. 5-30 us at HBM
// speed, so the forward is one launch of two 1024-thread CTAs per SM that walk through six phases
// separated by grid-wide barriers; all intermediate state (hash table, per-point entry, slot lists)
// stays L2-resident between phases.  Measured on B200 the scatter/gather phases are bound by the
// per-SM L1/LSU rate of one 128-byte line per cycle, not by HBM, so every phase is laid out to touch as
// few distinct lines per warp instruction as possible and points are dealt to CTAs in interleaved
// 256-point tiles (a contiguous split leaves the in-range work on a few SMs).
//
//   P0 clear     tile ranges of this CTA -> shared memory; TMA request for the rows of P1's first iteration;
//                64-bit hash entries := EMPTY (also makes the table L2-resident for the atomics of P1)
//   P1 insert    per point: [perturb] -> cell id -> claim / find the cell's entry: two blind 64-bit CASes of
//                ((cell << 32) | point index) along the linear probe order, then 32-byte bucket loads for whoever is
//                still pending; an entry already held by the cell with a larger index gets an atomicMin ->
//                entry.lo = index of the FIRST point of the cell.  Points that saw a smaller index can never be
//                first and are not looked at again in P2.
//   P2 flag      candidates: first-of-cell flag (entry.lo == index); first points counted per WARP, one byte each
//   P3 assign    exclusive scan over tiles in point order = mmcv's voxel index; rank >= V -> cell refused for
//                good; entry.lo := voxel row; slot lists initialised (coalesced fill)
//   P4 rank      per kept cell the P smallest point indices, by an atomicMin insertion chain (level k
//                keeps the k-th smallest; losers carry to level k+1)
//   P5 gather    per voxel (one lane each, 32 voxel rows per warp task): zero-fill the padded [P,C] rows
//                (coalesced), then gather the <= P points in slot order over the padding; coors from the first
//                point, num_points, point2voxel/slot and the mean
// What bounds it is the L2's rate of scattered transactions (~7-8 per point), not HBM: DESIGN.md 4.1 and
// tools/micro/atomics.cu.
#pragma once
#include <cooperative_groups.h>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace rpc {

constexpr int kPT = 1024;                       // threads per CTA of the persistent kernel
constexpr int kTile = 256;                      // points per tile: the unit that is dealt to the CTAs
constexpr int kSub = kPT / kTile;               // tiles a CTA works on per iteration (8 warps each)
constexpr int kCtasPerSM = 2;                   // resident CTAs per SM (register cap 32 at 1024 threads)
#ifndef RPC_OPT_PROBES
#define RPC_OPT_PROBES 2
#endif
constexpr int kOptProbes = RPC_OPT_PROBES;                   // blind CAS probes of the insert phase before it starts loading buckets
constexpr int kIlp = 4;                          // iterations of the rank phase taken together per thread ...
constexpr int kIlpMinP = 8;                      // ... when the slot lists are at least this long
constexpr int kTrCache = 128;                    // tile ranges cached in shared memory per CTA (2 KB)
constexpr int kMaxTiles = 24576;                // tile counts are scanned in shared memory (96 KB per CTA)
constexpr uint32_t kDropped = 0xFFFFFFFFu;      // hslot: point outside the range
constexpr uint32_t kFirstBit = 0x80000000u;     // hslot: point is the first of its cell (set in P2)
constexpr uint32_t kCandBit = 0x40000000u;      // hslot: point may be the first of its cell (set in P1)
constexpr uint32_t kEntryMask = 0x3FFFFFFFu;
constexpr uint32_t kRefused = 0xFFFFFFFFu;      // entry.lo after P3: cell came after the V-th voxel
constexpr unsigned long long kEmpty64 = ~0ull;

struct FwdParams {
    const float *pts, *raw;                     // [ntotal, C]
    unsigned long long *tab;                    // [tab_total] (cell << 32) | first index, later | voxel row
    uint32_t *hslot;                            // [ntotal]   table entry of each point + flag bits
    unsigned long long *wcnt;                   // [ntiles]   first points per warp of the tile, one byte each
    int *slots;                                 // [cap * P]  per voxel: point index per slot, kInf empty
    float *voxels; int *coors; int *num_points; int *voxel_num; int *p2v; int *p2s; float *feats;
    unsigned long long *stamps;                 // [gridDim][16] %globaltimer per CTA and phase, or NULL
    long long tab_total;
    int ntiles, gs, VT, gather_stride;               // VT: voxel rows per gather task (one warp); floats of smem per warp
    GridDesc gd;
    BatchDesc bd;
    PerturbDesc pd;
};

// read x,y,z of row i (and the perturbation when PERTURB) -- C == 4 uses one 16-byte load
template <int C, bool PERTURB>
__device__ __forceinline__ void load_xyz(const float *__restrict__ pts, const float *__restrict__ raw,
                                         const PerturbDesc &pd, int64_t i, float &x, float &y, float &z)
{
    if constexpr (C == 4) {
        float4 v = __ldg(reinterpret_cast<const float4 *>(pts) + i);
        x = v.x; y = v.y; z = v.z;
        if constexpr (PERTURB) {
            float4 r = __ldg(reinterpret_cast<const float4 *>(raw) + i);
            x = __fadd_rn(x, perturb_delta(r.x, pd.eb[0], pd.fb[0]));
            y = __fadd_rn(y, perturb_delta(r.y, pd.eb[1], pd.fb[1]));
            z = __fadd_rn(z, perturb_delta(r.z, pd.eb[2], pd.fb[2]));
        }
    } else {
        const float *p = pts + i * C;
        x = __ldg(p); y = __ldg(p + 1); z = __ldg(p + 2);
        if constexpr (PERTURB) {
            const float *r = raw + i * C;
            x = __fadd_rn(x, perturb_delta(__ldg(r), pd.eb[0], pd.fb[0]));
            y = __fadd_rn(y, perturb_delta(__ldg(r + 1), pd.eb[1], pd.fb[1]));
            z = __fadd_rn(z, perturb_delta(__ldg(r + 2), pd.eb[2], pd.fb[2]));
        }
    }
}

__device__ __forceinline__ void stamp(const FwdParams &p, int k)
{
    if (p.stamps && threadIdx.x == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
        p.stamps[blockIdx.x * 16 + k] = t;
    }
}

struct TileRange { int b, begin, end; };
__device__ __forceinline__ TileRange tile_range(const FwdParams &p, int tile)
{
    TileRange r;
    r.b = tile_to_sweep(p.bd, tile);
    r.begin = p.bd.pt_begin[r.b] + (tile - p.bd.tile_begin[r.b]) * kTile;
    r.end = min(r.begin + kTile, p.bd.pt_begin[r.b + 1]);
    return r;
}

// first table entry probed for a cell: multiplicative hash, then a multiply-shift range reduction
// (table sizes are not powers of two: 1.5 entries per point keeps the clear traffic down)
__device__ __forceinline__ uint32_t probe_start(uint32_t cell, uint32_t size)
{
    uint32_t h = cell * 0x9E3779B1u;
    h ^= h >> 15;
    h *= 0x85EBCA77u;
    return __umulhi(h, size);
}

// Tiles are dealt round-robin: CTA c owns tiles c, c + G, c + 2G, ... (G = gridDim.x) and works on kSub of
// them per iteration, one per group of 8 warps, thread order = point order inside the tile.  With 256-point
// tiles every CTA gets the same number of tiles to within one (a 1024-point unit left 3-versus-2 tile
// shares on a KITTI batch: the grid barriers then waited for the unlucky CTAs in every phase).
__device__ __forceinline__ int cta_iterations(int ntiles)
{
    const int mine = (int)blockIdx.x < ntiles ? (ntiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
    return (mine + kSub - 1) / kSub;
}
__device__ __forceinline__ int my_tile(int ntiles, int it)
{
    const int t = (int)blockIdx.x + (it * kSub + (int)(threadIdx.x >> 8)) * (int)gridDim.x;
    return t < ntiles ? t : -1;
}

// sum of the 8 per-warp byte counts of a tile (each <= 32)
__device__ __forceinline__ int tile_first_points(unsigned long long w8)
{
    return (int)(__vsadu4((uint32_t)w8, 0u) + __vsadu4((uint32_t)(w8 >> 32), 0u));
}

// exclusive rank of every flagged thread within its tile (8 warps) and the tile total.  sm: 32 ints.
__device__ __forceinline__ void tile_rank_flags(bool f, int &rank, int &total, int *sm)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t bal = __ballot_sync(0xFFFFFFFFu, f);
    if (lane == 0) sm[warp] = __popc(bal);
    __syncthreads();
    int pre = 0, tot = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int c = sm[(warp & ~7) + j];
        pre += j < (warp & 7) ? c : 0;
        tot += c;
    }
    __syncthreads();
    rank = pre + __popc(bal & ((1u << lane) - 1u));
    total = tot;
}

// ---- TMA (bulk async copy) + mbarrier helpers for the point-tile pipeline of P1 (C == 4: a tile of rows is a
// contiguous 16-byte-aligned range, so one cp.async.bulk per array brings it into shared memory while the
// previous iteration is still probing the table)
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_1d(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    uint32_t ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    } while (!ok);
}

// ---- batches of more than kMaxTiles tiles (P3): exclusive scan over groups of 2^gs tiles; a tile's prefix is its
// group's plus the first points of the group's earlier tiles
__device__ __noinline__ void big_scan(const unsigned long long *wcnt, int ntiles, int gs, int *prefix, int *sm)
{
    const int tid = threadIdx.x;
    const int ngr = (ntiles + (1 << gs) - 1) >> gs;
    const int per = (ngr + kPT - 1) / kPT;
    const int j0 = tid * per, j1 = min(j0 + per, ngr);
    auto group = [&](int j) {
        int s = 0;
        for (int t = j << gs; t < min((j + 1) << gs, ntiles); ++t) s += tile_first_points(__ldcg(wcnt + t));
        return s;
    };
    int s = 0;
    for (int j = j0; j < j1; ++j) s += group(j);
    int tot;
    int run = block_excl_scan<kPT>(s, sm, &tot);
    for (int j = j0; j < j1; ++j) { prefix[j] = run; run += group(j); }
    if (tid == kPT - 1) prefix[ngr] = tot;
    __syncthreads();
}
// (warp-uniform t in [0, ntiles]; the whole warp calls)
__device__ __noinline__ int big_prefix_at(const unsigned long long *wcnt, const int *prefix, int t, int gs)
{
    const int lane = threadIdx.x & 31;
    const int g0 = (t >> gs) << gs;
    int s = g0 + lane < t ? tile_first_points(__ldcg(wcnt + g0 + lane)) : 0;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, d);
    return prefix[t >> gs] + s;
}

template <int C, bool PERTURB>
__global__ void __launch_bounds__(kPT, kCtasPerSM)
k_hard_fwd(const __grid_constant__ FwdParams p)
{
    cg::grid_group grid = cg::this_grid();
    extern __shared__ __align__(16) unsigned char dyn_sm[];     // P3: int prefix[ntiles]; P5: gather tile
    __shared__ int sm[33];
    __shared__ int s_voff[RPC_MAX_BATCH + 1];                   // first voxel row of each sweep, then M
    __shared__ int s_pb[RPC_MAX_BATCH + 1];                     // first points before each sweep
    __shared__ __align__(8) unsigned long long s_bar[2 * (kPT / 32)];   // P1 (C == 4): one mbarrier per warp and stage
    const int tid = threadIdx.x;
    const int P = p.gd.P, V = p.gd.V;
    const int nct = gridDim.x;

    // ------------------------------------------------------------------------------- P0 clear
    stamp(p, 0);
    const int iters = cta_iterations(p.ntiles);
    // the (sweep, first, end) of this CTA's tiles, looked up once: every phase walks the same tiles, and the
    // search over the sweep boundaries in parameter space showed up as 4.6 % of the stall samples
    __shared__ int4 s_tr[kTrCache];
    if (tid < kTrCache && tid < iters * kSub) {
        const int tile = (int)blockIdx.x + tid * nct;
        int4 v = make_int4(0, 0, 0, -1);
        if (tile < p.ntiles) {
            const TileRange tr = tile_range(p, tile);
            v = make_int4(tr.b, tr.begin, tr.end, tile);
        }
        s_tr[tid] = v;
    }
    __syncthreads();
    auto my_range = [&](int it, TileRange &tr) -> int {              // -> tile index or -1
        const int k = it * kSub + (tid >> 8);
        if (k < kTrCache) {
            const int4 v = s_tr[k];
            tr.b = v.x; tr.begin = v.y; tr.end = v.z;
            return v.w;
        }
        const int tile = my_tile(p.ntiles, it);
        if (tile >= 0) tr = tile_range(p, tile);
        return tile;
    };
    // (the staging pipeline of P1 is set up here, and the rows of iteration 0 are requested before the table is
    // cleared: they travel while this phase and its barrier run)
    constexpr bool kTma = (C == 4);
    constexpr int kArr = PERTURB ? 2 : 1;                            // arrays staged per warp
    const int wlane = tid & 31, wid = tid >> 5;
    const uint32_t bar_sa = (uint32_t)__cvta_generic_to_shared(s_bar) + 8u * wid;          // + 8 * 32 * stage
    const uint32_t stage_sa = (uint32_t)__cvta_generic_to_shared(dyn_sm) + (uint32_t)(wid * kArr * 512);   // + stage
    constexpr uint32_t kStageBytes = (kPT / 32) * kArr * 512;
    auto issue_rows = [&](int it) {                                  // called by lane 0
        const int st = it & 1;
        TileRange tr;
        const int tile = my_range(it, tr);
        const uint32_t bar = bar_sa + 8u * (kPT / 32) * st;
        int r0 = 0, r1 = 0;
        if (tile >= 0) {
            r0 = tr.begin + (wid & 7) * 32;
            r1 = min(r0 + 32, tr.end);
        }
        if (r1 <= r0) { mbar_arrive(bar); return; }
        const uint32_t bytes = (uint32_t)(r1 - r0) * 16u;
        const uint32_t dst = stage_sa + kStageBytes * st;
        mbar_arrive_expect_tx(bar, bytes * kArr);
        tma_load_1d(dst, reinterpret_cast<const float4 *>(p.pts) + r0, bytes, bar);
        if constexpr (PERTURB) tma_load_1d(dst + 512, reinterpret_cast<const float4 *>(p.raw) + r0, bytes, bar);
    };
    if constexpr (kTma) {
        if (wlane == 0) {
            mbar_init(bar_sa, 1);
            mbar_init(bar_sa + 8u * (kPT / 32), 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            if (iters > 0) issue_rows(0);
        }
        __syncwarp();
    }
    // pull this CTA's point / perturbation tiles into L2 while the table is being cleared
    for (int it = 0; it < iters; ++it) {
        TileRange tr;
        const int tile = my_range(it, tr);
        if (tile < 0) continue;
        const size_t b0 = (size_t)tr.begin * C * 4, b1 = (size_t)tr.end * C * 4;
        for (size_t o = (b0 & ~size_t(127)) + (size_t)(tid & (kTile - 1)) * 128; o < b1; o += (size_t)kTile * 128) {
            asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const char *>(p.pts) + o));
            if constexpr (PERTURB)
                asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const char *>(p.raw) + o));
        }
    }
    {
        uint4 *t4 = reinterpret_cast<uint4 *>(p.tab);
        const long long n4 = p.tab_total >> 1;
        for (long long i = (long long)blockIdx.x * kPT + tid; i < n4; i += (long long)nct * kPT)
            t4[i] = make_uint4(~0u, ~0u, ~0u, ~0u);
    }
    if (p.stamps) __syncthreads();
    stamp(p, 8);
    grid.sync();
    stamp(p, 1);

    // ------------------------------------------------------------------------------- P1 insert
    // coordinate loads -> cell; 32-byte bucket load -> local scan; CAS only where a slot looked empty,
    // atomicMin only where the holder has a larger index.
    // C == 4: every WARP runs its own two-stage TMA pipeline -- lane 0 fetches the warp's 32 rows (512 bytes
    // per array, contiguous and 16-byte aligned) of iteration it + 1 into the other shared-memory stage and
    // arms that stage's mbarrier with the byte count while iteration it probes the table.  Per warp, not per
    // CTA: a CTA-wide stage (one barrier per iteration) made every warp wait for the slowest probe chain.
    for (int it = 0; it < iters; ++it) {
        TileRange tr;
        const int tile = my_range(it, tr);
        int idx = -1, cell = -1;
        uint32_t off = 0, nb = 1;
        if (tile >= 0) {
            if (tr.begin + (tid & (kTile - 1)) < tr.end) {
                idx = tr.begin + (tid & (kTile - 1));
                off = p.bd.tab_off[tr.b]; nb = p.bd.tab_size[tr.b] >> 2;
            }
        }
        if constexpr (kTma) {
            // every lane is past its reads of the other stage (__syncwarp at the end of the iteration): refill it
            if (wlane == 0 && it + 1 < iters) issue_rows(it + 1);
            mbar_wait(bar_sa + 8u * (kPT / 32) * (it & 1), (uint32_t)(it >> 1) & 1u);
        }
        if (idx >= 0) {
            float x, y, z;
            if constexpr (kTma) {
                const float4 *sp = reinterpret_cast<const float4 *>(dyn_sm) +
                                   ((it & 1) * (kPT / 32) + wid) * kArr * 32 + wlane;
                const float4 v = sp[0];
                x = v.x; y = v.y; z = v.z;
                if constexpr (PERTURB) {
                    const float4 r = sp[32];
                    x = __fadd_rn(x, perturb_delta(r.x, p.pd.eb[0], p.pd.fb[0]));
                    y = __fadd_rn(y, perturb_delta(r.y, p.pd.eb[1], p.pd.fb[1]));
                    z = __fadd_rn(z, perturb_delta(r.z, p.pd.eb[2], p.pd.fb[2]));
                }
            } else {
                load_xyz<C, PERTURB>(p.pts, p.raw, p.pd, idx, x, y, z);
            }
            cell = point_cell(x, y, z, p.gd);
            if (p.p2v) p.p2v[idx] = -1;
            if (p.p2s) p.p2s[idx] = -1;
            if (cell < 0) p.hslot[idx] = kDropped;
        }
        if constexpr (kTma) __syncwarp();                            // the stage may be refilled from here on
        if (p.stamps && it == 0) { __syncthreads(); stamp(p, 13); }
        // Probe order of a cell: linear over the entries from its home entry.  The first kOptProbes probes are
        // blind CASes (claim-or-find in ONE L2 round trip: at LiDAR load factors ~0.3 the home entry is free or
        // already the cell's own for most points); whoever is still pending goes on in the same order with
        // 32-byte bucket loads, which look at up to four entries per round trip.
        uint32_t pos = cell >= 0 ? probe_start((uint32_t)cell, nb << 2) : 0u;
        bool pend = cell >= 0;
        const unsigned long long mine = ((unsigned long long)(uint32_t)cell << 32) | (uint32_t)idx;
#pragma unroll 1
        for (int k = 0; k < kOptProbes; ++k) {
            if (pend) {
                uint32_t h = off + pos;
                const unsigned long long old = atomicCAS(&p.tab[h], kEmpty64, mine);
                if (old == kEmpty64) {
                    p.hslot[idx] = h | kCandBit;                     // claimed: first so far
                    pend = false;
                } else if ((uint32_t)(old >> 32) == (uint32_t)cell) {
                    if ((uint32_t)old > (uint32_t)idx) {
                        atomicMin(&p.tab[h], mine);
                        h |= kCandBit;                               // displaced a later point
                    }
                    p.hslot[idx] = h;
                    pend = false;
                } else {
                    pos = pos + 1 == (nb << 2) ? 0u : pos + 1;
                }
            }
        }
        uint32_t bkt = pos >> 2;
        int j0 = (int)(pos & 3u);                                    // first slot of the bucket still to look at
        while (pend) {
#ifndef RPC_NO_LD256
            unsigned long long e[4];                                 // the whole 32-byte bucket with one 256-bit load
            asm volatile("ld.global.cg.v4.u64 {%0,%1,%2,%3}, [%4];"
                         : "=l"(e[0]), "=l"(e[1]), "=l"(e[2]), "=l"(e[3]) : "l"(p.tab + off + 4 * bkt) : "memory");
#else
            const ulonglong2 *bp = reinterpret_cast<const ulonglong2 *>(p.tab + off + 4 * bkt);
            const ulonglong2 e01 = __ldcg(bp), e23 = __ldcg(bp + 1);
            const unsigned long long e[4] = {e01.x, e01.y, e23.x, e23.y};
#endif
            int j = 4;
#pragma unroll
            for (int q = 3; q >= 0; --q)
                if (q >= j0 && (e[q] == kEmpty64 || (uint32_t)(e[q] >> 32) == (uint32_t)cell)) j = q;
            if (j == 4) { bkt = bkt + 1 == nb ? 0 : bkt + 1; j0 = 0; continue; }
            unsigned long long old = j == 0 ? e[0] : j == 1 ? e[1] : j == 2 ? e[2] : e[3];
            j0 = j;
            if (old == kEmpty64) {
                old = atomicCAS(&p.tab[off + 4 * bkt + j], kEmpty64, mine);
                if (old != kEmpty64 && (uint32_t)(old >> 32) != (uint32_t)cell) {
                    j0 += 1;                                         // lost the slot to another cell
                    if (j0 == 4) { bkt = bkt + 1 == nb ? 0 : bkt + 1; j0 = 0; }
                    continue;
                }
            }
            uint32_t h = off + 4 * bkt + j0;
            if (old == kEmpty64) {
                h |= kCandBit;                                       // claimed: first so far
            } else if ((uint32_t)old > (uint32_t)idx) {
                atomicMin(&p.tab[h], mine);
                h |= kCandBit;                                       // displaced a later point
            }
            p.hslot[idx] = h;
            pend = false;
        }
    }
    if (p.stamps) __syncthreads();
    stamp(p, 9);
    grid.sync();
    stamp(p, 2);

    // ------------------------------------------------------------------------------- P2 flag
    // thread order = point order inside the tile, which the scan of P3 needs
    for (int it = 0; it < iters; ++it) {
        TileRange tr;
        const int tile = my_range(it, tr);
        uint32_t hs = kDropped;
        int idx = -1;
        if (tile >= 0) {
            if (tr.begin + (tid & (kTile - 1)) < tr.end) { idx = tr.begin + (tid & (kTile - 1)); hs = __ldcg(p.hslot + idx); }
        }
        const uint32_t lo = (hs != kDropped && (hs & kCandBit)) ? (uint32_t)__ldcg(p.tab + (hs & kEntryMask)) : 0xFFFFFFFEu;
        const bool f = lo == (uint32_t)idx;
        if (f) p.hslot[idx] = hs | kFirstBit;
        // first points per warp, one byte per warp of the tile: no CTA barrier in this phase, and P3 gets both
        // the tile total and the warp's offset inside the tile from one 8-byte load
        const uint32_t bal = __ballot_sync(0xFFFFFFFFu, f);
        if (wlane == 0 && tile >= 0)
            reinterpret_cast<unsigned char *>(p.wcnt)[(size_t)tile * 8 + (wid & 7)] = (unsigned char)__popc(bal);
    }
    if (p.stamps) __syncthreads();
    stamp(p, 10);
    grid.sync();
    stamp(p, 3);

    // ------------------------------------------------------------------------------- P3 assign
    // every CTA scans all tile counts into shared memory (ntiles <= kMaxTiles): prefix[j] = first points
    // in tiles < j; a sweep's voxel index base is prefix[its first tile]
    int *prefix = reinterpret_cast<int *>(dyn_sm);
    // (batches of more than kMaxTiles tiles: the scan runs over GROUPS of 2^gs <= 32 tiles -- big_* helpers, kept out
    // of line so that the usual path pays no registers for them)
    const int gs = p.gs;
    if (gs == 0) {
        const int per = (p.ntiles + kPT - 1) / kPT;
        const int j0 = tid * per, j1 = min(j0 + per, p.ntiles);
        constexpr int kKeep = 6;                                     // counts kept in registers between the two passes
        int s = 0, c[kKeep];
#pragma unroll
        for (int q = 0; q < kKeep; ++q) {
            c[q] = j0 + q < j1 ? tile_first_points(__ldcg(p.wcnt + j0 + q)) : 0;
            s += c[q];
        }
        for (int j = j0 + kKeep; j < j1; ++j) s += tile_first_points(__ldcg(p.wcnt + j));
        int tot;
        int run = block_excl_scan<kPT>(s, sm, &tot);
#pragma unroll
        for (int q = 0; q < kKeep; ++q)
            if (j0 + q < j1) { prefix[j0 + q] = run; run += c[q]; }
        for (int j = j0 + kKeep; j < j1; ++j) { prefix[j] = run; run += tile_first_points(__ldcg(p.wcnt + j)); }
        if (tid == kPT - 1) prefix[p.ntiles] = tot;
        __syncthreads();
        if (tid <= p.bd.B) s_pb[tid] = prefix[p.bd.tile_begin[tid]];
    } else {
        big_scan(p.wcnt, p.ntiles, gs, prefix, sm);
        if (wid == 0)
            for (int b = 0; b <= p.bd.B; ++b) {
                const int v = big_prefix_at(p.wcnt, prefix, p.bd.tile_begin[b], gs);
                if (wlane == 0) s_pb[b] = v;
            }
    }
    __syncthreads();
    if (tid == 0) {
        int acc = 0;
        for (int b = 0; b < p.bd.B; ++b) {
            const int m = min(s_pb[b + 1] - s_pb[b], V);
            s_voff[b] = acc;
            if (blockIdx.x == 0) p.voxel_num[b] = m;
            acc += m;
        }
        s_voff[p.bd.B] = acc;
        if (blockIdx.x == 0) p.voxel_num[p.bd.B] = acc;
    }
    __syncthreads();
    const int lvl_step = kTile % P;
    for (int it = 0; it < iters; ++it) {
        TileRange tr;
        const int tile = my_range(it, tr);
        uint32_t hs = kDropped;
        int idx = -1, base = 0, voff = 0;
        if (tile >= 0) {
            base = (gs == 0 ? prefix[tile] : big_prefix_at(p.wcnt, prefix, tile, gs)) - s_pb[tr.b];
            voff = s_voff[tr.b];
            if (tr.begin + (tid & (kTile - 1)) < tr.end) { idx = tr.begin + (tid & (kTile - 1)); hs = __ldcg(p.hslot + idx); }
        }
        const bool f = hs != kDropped && (hs & kFirstBit);
        // rank of a first point inside its tile = first points of the tile's earlier warps (the bytes P2 wrote)
        // + earlier lanes of this warp; no CTA barrier
        const unsigned long long w8 = tile >= 0 ? __ldcg(p.wcnt + tile) : 0ull;
        const int total = tile_first_points(w8);
        const int rank = tile_first_points(w8 & ((1ull << (8 * (wid & 7))) - 1ull)) +
                         __popc(__ballot_sync(0xFFFFFFFFu, f) & ((1u << wlane) - 1u));
        // slot lists of the tile's admitted voxels: rows [voff+base, voff+min(base+total, V)), levels 1.. := kInf
        // (level 0 is written by the first point's own thread below, so the two need no ordering)
        {
            const int r0 = min(base, V), r1 = min(base + total, V);
            int *sl = p.slots + (int64_t)(voff + r0) * P;
            int lvl = (tid & (kTile - 1)) % P;                       // level of element t, kept without a division
            for (int t = tid & (kTile - 1); t < (r1 - r0) * P; t += kTile) {
                if (lvl) sl[t] = kInf;
                lvl += lvl_step;
                if (lvl >= P) lvl -= P;
            }
        }
        if (f) {
            const int r = base + rank;
            uint32_t *lo = reinterpret_cast<uint32_t *>(p.tab + (hs & kEntryMask));   // low word
            if (r >= V) {
                *lo = kRefused;
            } else {
                *lo = (uint32_t)(voff + r);
                p.slots[(int64_t)(voff + r) * P] = idx;
            }
        }
    }
    if (p.stamps) __syncthreads();
    stamp(p, 11);
    grid.sync();
    stamp(p, 4);

    // ------------------------------------------------------------------------------- P4 rank
    // Level q of a voxel's list only ever decreases (atomicMin); whatever loses at level q -- the
    // arriving index or the one it displaces -- carries to level q+1.  Every index but the final
    // minimum of a level is carried exactly once, so level q ends up holding the (q+1)-th smallest
    // index regardless of interleaving.  Slot 0 is the first point (already known).  An index
    // larger than the current last level can never enter the list, which prunes crowded voxels.
    // kIlp iterations of a thread are taken together, stage by stage (entry words, voxel rows, last levels, then
    // the chains level by level), so that their L2 round trips overlap instead of adding up.
    if (P > 1 && P < kIlpMinP) {
        // short lists (KITTI: P = 5, most cells refused by the cap): one point at a time measured faster
        for (int it = 0; it < iters; ++it) {
            TileRange tr;
            const int tile = my_range(it, tr);
            if (tile < 0) continue;
            const int idx = tr.begin + (tid & (kTile - 1));
            if (idx >= tr.end) continue;
            const uint32_t hs = __ldcg(p.hslot + idx);
            if (hs & kFirstBit) continue;                                // first point or dropped (kDropped has the bit)
            const uint32_t v = (uint32_t)__ldcg(p.tab + (hs & kEntryMask));
            if (v == kRefused) continue;
            int *sl = p.slots + (int64_t)v * P;
            if (__ldcg(sl + P - 1) <= idx) continue;
            int carry = idx;
            for (int q = 1; q < P && carry != kInf; ++q) {
                const int old = atomicMin(sl + q, carry);
                if (old > carry) carry = old;                           // displaced (kInf: was free)
            }
        }
    } else if (P > 1) {
        for (int it0 = 0; it0 < iters; it0 += kIlp) {
            int carry[kIlp];
            uint32_t v[kIlp];
#pragma unroll
            for (int u = 0; u < kIlp; ++u) {
                carry[u] = kInf;
                v[u] = kDropped;                                       // here: the point's hslot word
                if (it0 + u < iters) {
                    TileRange tr;
                    const int tile = my_range(it0 + u, tr);
                    const int idx = tr.begin + (tid & (kTile - 1));
                    if (tile >= 0 && idx < tr.end) { carry[u] = idx; v[u] = __ldcg(p.hslot + idx); }
                }
            }
#pragma unroll
            for (int u = 0; u < kIlp; ++u)                             // voxel row of the point's cell
                v[u] = (v[u] & kFirstBit) ? kRefused : (uint32_t)__ldcg(p.tab + (v[u] & kEntryMask));   // kDropped has the bit
            {
                int last[kIlp];
#pragma unroll
                for (int u = 0; u < kIlp; ++u) last[u] = v[u] != kRefused ? __ldcg(p.slots + (int64_t)v[u] * P + P - 1) : 0;
#pragma unroll
                for (int u = 0; u < kIlp; ++u)
                    if (last[u] <= carry[u]) carry[u] = kInf;           // cannot enter the list (or nothing to do)
            }
            for (int q = 1; q < P; ++q) {
                bool any = false;
#pragma unroll
                for (int u = 0; u < kIlp; ++u) {
                    if (carry[u] != kInf) {
                        const int old = atomicMin(p.slots + (int64_t)v[u] * P + q, carry[u]);
                        if (old > carry[u]) carry[u] = old;             // displaced (kInf: was free)
                        any |= carry[u] != kInf;
                    }
                }
                if (!any) break;
            }
        }
    }
    if (P > 1) {
        if (p.stamps) __syncthreads();
        stamp(p, 12);
        grid.sync();
    }
    stamp(p, 5);

    // ------------------------------------------------------------------------------- P5 gather
    // Warp-granular, no CTA barrier, one LANE per voxel; tasks of VT = 32 consecutive voxel rows are dealt
    // round-robin over the CTAs first, then over their warps.  The padded [VT,P,C] block of the output is one
    // contiguous 16-byte-aligned range: the warp zero-fills it with coalesced float4 stores, then (after
    // __syncwarp, which orders the stores of the participating lanes) every lane walks the kept slots of
    // its voxel in order: gather (and perturb) the point row, store it over the padding, add it to the
    // running per-channel sum -- the HardSimpleVFE order; adding the zero padding would not change it --
    // and record point2voxel / point2slot.
    {
        const int F = p.gd.F;
        const int M = s_voff[p.bd.B];
        const int lane = tid & 31, warp = tid >> 5;
        const int VT = p.VT;                                               // 32 (fewer when P is large), % 4 == 0
        int *s_idx = reinterpret_cast<int *>(dyn_sm) + (size_t)warp * p.gather_stride;   // [VT * P] slot lists
        const long long ntask = ((long long)M + VT - 1) / VT;
        for (long long t = (long long)warp * nct + blockIdx.x; t < ntask; t += (long long)nct * (kPT / 32)) {
            const long long v0 = t * VT;
            const int nv = (int)min((long long)VT, M - v0);
            const int ns = nv * P;
            __syncwarp();
            for (int j = lane; j < ns; j += 32) s_idx[j] = __ldcg(p.slots + v0 * P + j);
            {
                float *out = p.voxels + v0 * P * C;
                const int nf = ns * C;
                const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
                for (int q = lane; q < (nf >> 2); q += 32) reinterpret_cast<float4 *>(out)[q] = z4;
                for (int q = (nf & ~3) + lane; q < nf; q += 32) out[q] = 0.0f;
            }
            __syncwarp();
            if (lane < nv) {
                const long long v = v0 + lane;
                float *row = p.voxels + v * P * C;
                float sum[C];
#pragma unroll
                for (int c = 0; c < C; ++c) sum[c] = 0.0f;
                int n = 0;
                int cx = 0, cy = 0, cz = 0;
                for (; n < P; ++n) {
                    const int idx = s_idx[lane * P + n];
                    if (idx == kInf) break;                                    // levels fill from slot 0 up
                    float x[C];
                    if constexpr (C == 4) {
                        const float4 t4 = __ldg(reinterpret_cast<const float4 *>(p.pts) + idx);
                        x[0] = t4.x; x[1] = t4.y; x[2] = t4.z; x[3] = t4.w;
                        if constexpr (PERTURB) {
                            const float4 d = __ldg(reinterpret_cast<const float4 *>(p.raw) + idx);
                            const float dv[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
                            for (int c = 0; c < 4; ++c) x[c] = __fadd_rn(x[c], perturb_delta(dv[c], p.pd.eb[c], p.pd.fb[c]));
                        }
                        reinterpret_cast<float4 *>(row)[n] = make_float4(x[0], x[1], x[2], x[3]);
                    } else {
#pragma unroll
                        for (int c = 0; c < C; ++c) x[c] = __ldg(p.pts + (int64_t)idx * C + c);
                        if constexpr (PERTURB) {
#pragma unroll
                            for (int c = 0; c < C; ++c)
                                x[c] = __fadd_rn(x[c], perturb_delta(__ldg(p.raw + (int64_t)idx * C + c), p.pd.eb[c], p.pd.fb[c]));
                        }
#pragma unroll
                        for (int c = 0; c < C; ++c) row[n * C + c] = x[c];
                    }
#pragma unroll
                    for (int c = 0; c < C; ++c) sum[c] = __fadd_rn(sum[c], x[c]);
                    if (n == 0) {                                              // coors from the first point
                        cx = axis_cell(x[0], p.gd, 0);
                        cy = axis_cell(x[1], p.gd, 1);
                        cz = axis_cell(x[2], p.gd, 2);
                    }
                    if (p.p2v) p.p2v[idx] = (int)v;
                    if (p.p2s) p.p2s[idx] = n;
                }
                p.num_points[v] = n;
                int b = 0;
                while (b + 1 < p.bd.B && v >= s_voff[b + 1]) ++b;
                reinterpret_cast<int4 *>(p.coors)[v] = make_int4(b, cz, cy, cx);
                if (p.feats) {
                    const float fn = (float)n;
#pragma unroll
                    for (int c = 0; c < C; ++c)
                        if (c < F) p.feats[v * F + c] = __fdiv_rn(sum[c], fn);
                }
            }
        }
    }
    if (p.stamps) __syncthreads();
    stamp(p, 6);
}

}  // namespace rpc

// hard_fwd.cuh -- deterministic hard voxelization of a batch of sweeps (+ optional perturbation and
// HardSimpleVFE mean) as ONE persistent cooperative kernel, sm_100a.
//
// Replaces mmcv `hard_voxelize_forward` with a parallel formulation that gives the SAME
// result as the serial CPU walk.  At the named sizes the path moves ~30-200 MB, i.e. 5-30 us at HBM
// speed, so the forward is one launch of one 1024-thread CTA per SM that walks through six phases
// separated by grid-wide barriers; all intermediate state (hash table, per-point entry, slot lists)
// stays L2-resident between phases.  Measured on B200 the scatter/gather phases are bound by the
// per-SM L1/LSU rate of one 128-byte line per cycle, not by HBM, so every phase is laid out to touch as
// few distinct lines per warp instruction as possible and points are dealt to CTAs in interleaved
// 1024-point tiles (a contiguous split leaves the in-range work on a few SMs).
//
//   P0 clear     64-bit hash entries := EMPTY (also makes the table L2-resident for the atomics of P1)
//   P1 insert    per point: [perturb] -> cell id -> claim / find the cell's entry with ONE 64-bit CAS
//                ((cell << 32) | point index); an entry already held by the cell with a larger index
//                gets an atomicMin -> entry.lo = index of the FIRST point of the cell.  Points that
//                saw a smaller index can never be first and are not looked at again in P2.
//   P2 flag      candidates: first-of-cell flag (entry.lo == index); first points counted per tile
//   P3 assign    exclusive scan over tiles in point order = mmcv's voxel index; rank >= V -> cell
//                refused for good; entry.lo := voxel row; slot lists initialised (coalesced fill)
//   P4 rank      per kept cell the P smallest point indices, by an atomicMin insertion chain (level k
//                keeps the k-th smallest; losers carry to level k+1)
//   P5 gather    per voxel: gather the <= P points in slot order into shared memory and to the
//                zero-padded [P,C] rows (coalesced), coors from the first point, num_points,
//                point2voxel/slot and the mean
#pragma once
#include <cooperative_groups.h>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace rpc {

constexpr int kPT = 1024;                       // threads per CTA of the persistent kernel = points per tile
constexpr int kKB = 1;                          // tiles a CTA keeps in flight in the unordered phases
constexpr int kCtasPerSM = 2;                   // resident CTAs per SM (register cap 32 at 1024 threads)
constexpr int kMaxTiles = 32768;                // tile counts are scanned in shared memory (128 KB)
constexpr int kGatherSlots = 2048;              // (voxel, slot) pairs per gather iteration
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
    int *tile_cnt;                              // [ntiles]   first points per tile
    int *slots;                                 // [cap * P]  per voxel: point index per slot, kInf empty
    float *voxels; int *coors; int *num_points; int *voxel_num; int *p2v; int *p2s; float *feats;
    unsigned long long *stamps;                 // [gridDim][16] %globaltimer per CTA and phase, or NULL
    long long tab_total;
    int ntiles, VB;
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
    r.begin = p.bd.pt_begin[r.b] + (tile - p.bd.tile_begin[r.b]) * kPT;
    r.end = min(r.begin + kPT, p.bd.pt_begin[r.b + 1]);
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

// For kKB tiles at once: exclusive rank of every flagged thread within its tile (thread order = point
// order) and the tile totals.  sm: kKB * 33 ints.
__device__ __forceinline__ void block_rank_flags(const bool f[kKB], int rank[kKB], int total[kKB], int *sm)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t bal[kKB];
#pragma unroll
    for (int k = 0; k < kKB; ++k) {
        bal[k] = __ballot_sync(0xFFFFFFFFu, f[k]);
        if (lane == 0) sm[k * 33 + warp] = __popc(bal[k]);
    }
    __syncthreads();
    if (warp < kKB) {                            // kKB <= 32 warps
        const int w = sm[warp * 33 + lane];
        int incl = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int t = __shfl_up_sync(0xFFFFFFFFu, incl, d);
            if (lane >= d) incl += t;
        }
        sm[warp * 33 + lane] = incl - w;
        if (lane == 31) sm[warp * 33 + 32] = incl;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < kKB; ++k) {
        rank[k] = sm[k * 33 + warp] + __popc(bal[k] & ((1u << lane) - 1u));
        total[k] = sm[k * 33 + 32];
    }
    __syncthreads();
}

template <int C, bool PERTURB>
__global__ void __launch_bounds__(kPT, kCtasPerSM)
k_hard_fwd(const __grid_constant__ FwdParams p)
{
    cg::grid_group grid = cg::this_grid();
    extern __shared__ __align__(16) unsigned char dyn_sm[];     // P3: int prefix[ntiles]; P5: gather tile
    __shared__ int sm[kKB * 33];
    __shared__ int s_voff[RPC_MAX_BATCH + 1];                   // first voxel row of each sweep, then M
    const int tid = threadIdx.x;
    const int P = p.gd.P, V = p.gd.V;
    const int nct = gridDim.x;

    // ------------------------------------------------------------------------------- P0 clear
    stamp(p, 0);
    // pull this CTA's point / perturbation tiles into L2 while the table is being cleared
    for (int tile = blockIdx.x; tile < p.ntiles; tile += nct) {
        const TileRange tr = tile_range(p, tile);
        const size_t b0 = (size_t)tr.begin * C * 4, b1 = (size_t)tr.end * C * 4;
        for (size_t o = (b0 & ~size_t(127)) + (size_t)tid * 128; o < b1; o += (size_t)kPT * 128) {
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
    // thread t owns point t of kKB tiles (blockIdx.x + k*gridDim.x ...) at once.  The SM issues in
    // order, so every step is written as "issue kKB independent memory operations, then consume them":
    // coordinate loads -> cells; 32-byte bucket loads -> local scan; CAS only where a slot looked empty.
    for (int t0 = blockIdx.x; t0 < p.ntiles; t0 += kKB * nct) {
        int cell[kKB], idx[kKB];
        uint32_t bkt[kKB], off[kKB], nb[kKB];
#pragma unroll
        for (int k = 0; k < kKB; ++k) {
            const int tile = t0 + k * nct;
            idx[k] = -1; off[k] = 0; nb[k] = 1;
            if (tile < p.ntiles) {
                const TileRange tr = tile_range(p, tile);
                if (tr.begin + tid < tr.end) {
                    idx[k] = tr.begin + tid;
                    off[k] = p.bd.tab_off[tr.b]; nb[k] = p.bd.tab_size[tr.b] >> 2;
                }
            }
        }
        {
            float xyz[kKB][3], d[kKB][3];
#pragma unroll
            for (int k = 0; k < kKB; ++k) {
                const int i = max(idx[k], 0);                       // tail threads re-read row 0 (unused)
                if constexpr (C == 4) {
                    const float4 v = __ldg(reinterpret_cast<const float4 *>(p.pts) + i);
                    xyz[k][0] = v.x; xyz[k][1] = v.y; xyz[k][2] = v.z;
                    if constexpr (PERTURB) {
                        const float4 r = __ldg(reinterpret_cast<const float4 *>(p.raw) + i);
                        d[k][0] = r.x; d[k][1] = r.y; d[k][2] = r.z;
                    }
                } else {
#pragma unroll
                    for (int c = 0; c < 3; ++c) {
                        xyz[k][c] = __ldg(p.pts + (int64_t)i * C + c);
                        if constexpr (PERTURB) d[k][c] = __ldg(p.raw + (int64_t)i * C + c);
                    }
                }
            }
#pragma unroll
            for (int k = 0; k < kKB; ++k) {
                if constexpr (PERTURB) {
#pragma unroll
                    for (int c = 0; c < 3; ++c)
                        xyz[k][c] = __fadd_rn(xyz[k][c], perturb_delta(d[k][c], p.pd.eb[c], p.pd.fb[c]));
                }
                cell[k] = idx[k] >= 0 ? point_cell(xyz[k][0], xyz[k][1], xyz[k][2], p.gd) : -1;
            }
        }
#pragma unroll
        for (int k = 0; k < kKB; ++k) {
            if (idx[k] < 0) continue;
            if (p.p2v) p.p2v[idx[k]] = -1;
            if (p.p2s) p.p2s[idx[k]] = -1;
            if (cell[k] < 0) p.hslot[idx[k]] = kDropped;
            else bkt[k] = probe_start((uint32_t)cell[k], nb[k]);
        }
        if (p.stamps && t0 == blockIdx.x) { __syncthreads(); stamp(p, 13); }
        int j0[kKB];                                                // first slot of the bucket still to look at
        bool pend[kKB];
#pragma unroll
        for (int k = 0; k < kKB; ++k) { pend[k] = cell[k] >= 0; j0[k] = 0; }
        for (;;) {
            bool any = false;
#pragma unroll
            for (int k = 0; k < kKB; ++k) any |= pend[k];
            if (!any) break;
            ulonglong2 e01[kKB], e23[kKB];
#pragma unroll
            for (int k = 0; k < kKB; ++k) {
                if (!pend[k]) continue;
                const ulonglong2 *bp = reinterpret_cast<const ulonglong2 *>(p.tab + off[k] + 4 * bkt[k]);
                e01[k] = __ldcg(bp); e23[k] = __ldcg(bp + 1);
            }
            int jc[kKB];                                            // slot to CAS, or -1
            unsigned long long old[kKB];
#pragma unroll
            for (int k = 0; k < kKB; ++k) {
                jc[k] = -1;
                if (!pend[k]) continue;
                const unsigned long long e[4] = {e01[k].x, e01[k].y, e23[k].x, e23[k].y};
                int j = 4;
#pragma unroll
                for (int q = 3; q >= 0; --q)
                    if (q >= j0[k] && (e[q] == kEmpty64 || (uint32_t)(e[q] >> 32) == (uint32_t)cell[k])) j = q;
                if (j == 4) { bkt[k] = bkt[k] + 1 == nb[k] ? 0 : bkt[k] + 1; j0[k] = 0; continue; }
                const unsigned long long ej = j == 0 ? e[0] : j == 1 ? e[1] : j == 2 ? e[2] : e[3];
                j0[k] = j;
                if (ej == kEmpty64) { jc[k] = j; continue; }
                old[k] = ej;                                         // the cell's entry, no atomic needed to find it
                jc[k] = -2;
            }
#pragma unroll
            for (int k = 0; k < kKB; ++k)
                if (jc[k] >= 0)
                    old[k] = atomicCAS(&p.tab[off[k] + 4 * bkt[k] + jc[k]], kEmpty64,
                                       ((unsigned long long)(uint32_t)cell[k] << 32) | (uint32_t)idx[k]);
#pragma unroll
            for (int k = 0; k < kKB; ++k) {
                if (jc[k] == -1) continue;
                const uint32_t i = (uint32_t)idx[k];
                if (old[k] != kEmpty64 && (uint32_t)(old[k] >> 32) != (uint32_t)cell[k]) {
                    j0[k] += 1;                                      // lost the slot to another cell
                    if (j0[k] == 4) { bkt[k] = bkt[k] + 1 == nb[k] ? 0 : bkt[k] + 1; j0[k] = 0; }
                    continue;
                }
                uint32_t h = off[k] + 4 * bkt[k] + j0[k];
                if (old[k] == kEmpty64) {
                    h |= kCandBit;                                   // claimed: first so far
                } else if ((uint32_t)old[k] > i) {
                    atomicMin(&p.tab[h], ((unsigned long long)(uint32_t)cell[k] << 32) | i);
                    h |= kCandBit;                                   // displaced a later point
                }
                p.hslot[i] = h;
                pend[k] = false;
            }
        }
    }
    if (p.stamps) __syncthreads();
    stamp(p, 9);
    grid.sync();
    stamp(p, 2);

    // ------------------------------------------------------------------------------- P2 flag
    // thread t owns point t of the tile (thread order = point order, which the scan of P3 needs)
    for (int t0 = blockIdx.x; t0 < p.ntiles; t0 += kKB * nct) {
        uint32_t hs[kKB];
        int idx[kKB];
#pragma unroll
        for (int k = 0; k < kKB; ++k) {
            const int tile = t0 + k * nct;
            hs[k] = kDropped; idx[k] = -1;
            if (tile < p.ntiles) {
                const TileRange tr = tile_range(p, tile);
                if (tr.begin + tid < tr.end) { idx[k] = tr.begin + tid; hs[k] = __ldcg(p.hslot + idx[k]); }
            }
        }
        uint32_t lo[kKB];
#pragma unroll
        for (int k = 0; k < kKB; ++k)
            lo[k] = (hs[k] != kDropped && (hs[k] & kCandBit)) ? (uint32_t)__ldcg(p.tab + (hs[k] & kEntryMask))
                                                              : 0xFFFFFFFEu;
        bool f[kKB];
        int rank[kKB], total[kKB];
#pragma unroll
        for (int k = 0; k < kKB; ++k) {
            f[k] = lo[k] == (uint32_t)idx[k];
            if (f[k]) p.hslot[idx[k]] = hs[k] | kFirstBit;
        }
        block_rank_flags(f, rank, total, sm);
#pragma unroll
        for (int k = 0; k < kKB; ++k)
            if (tid == k && t0 + k * nct < p.ntiles) p.tile_cnt[t0 + k * nct] = total[k];
    }
    if (p.stamps) __syncthreads();
    stamp(p, 10);
    grid.sync();
    stamp(p, 3);

    // ------------------------------------------------------------------------------- P3 assign
    // every CTA scans all tile counts into shared memory (ntiles <= kMaxTiles): prefix[j] = first points
    // in tiles < j; a sweep's voxel index base is prefix[its first tile]
    int *prefix = reinterpret_cast<int *>(dyn_sm);
    {
        const int per = (p.ntiles + kPT - 1) / kPT;
        const int j0 = tid * per, j1 = min(j0 + per, p.ntiles);
        int s = 0;
        for (int j = j0; j < j1; ++j) s += __ldcg(p.tile_cnt + j);
        int tot;
        int run = block_excl_scan<kPT>(s, sm, &tot);
        for (int j = j0; j < j1; ++j) { prefix[j] = run; run += __ldcg(p.tile_cnt + j); }
        if (tid == kPT - 1) prefix[p.ntiles] = tot;
        __syncthreads();
        if (tid == 0) {
            int acc = 0;
            for (int b = 0; b < p.bd.B; ++b) {
                const int m = min(prefix[p.bd.tile_begin[b + 1]] - prefix[p.bd.tile_begin[b]], V);
                s_voff[b] = acc;
                if (blockIdx.x == 0) p.voxel_num[b] = m;
                acc += m;
            }
            s_voff[p.bd.B] = acc;
            if (blockIdx.x == 0) p.voxel_num[p.bd.B] = acc;
        }
        __syncthreads();
    }
    for (int t0 = blockIdx.x; t0 < p.ntiles; t0 += kKB * nct) {
        uint32_t hs[kKB];
        int idx[kKB], base[kKB], voff[kKB];
        bool f[kKB];
#pragma unroll
        for (int k = 0; k < kKB; ++k) {
            const int tile = t0 + k * nct;
            hs[k] = kDropped; idx[k] = -1; base[k] = 0; voff[k] = 0;
            if (tile < p.ntiles) {
                const TileRange tr = tile_range(p, tile);
                base[k] = prefix[tile] - prefix[p.bd.tile_begin[tr.b]];
                voff[k] = s_voff[tr.b];
                if (tr.begin + tid < tr.end) { idx[k] = tr.begin + tid; hs[k] = __ldcg(p.hslot + idx[k]); }
            }
            f[k] = hs[k] != kDropped && (hs[k] & kFirstBit);
        }
        int rank[kKB], total[kKB];
        block_rank_flags(f, rank, total, sm);
        // slot lists of the tile's admitted voxels: rows [voff+base, voff+min(base+total, V)) := kInf
#pragma unroll
        for (int k = 0; k < kKB; ++k) {
            const int r0 = min(base[k], V), r1 = min(base[k] + total[k], V);
            int *sl = p.slots + (int64_t)(voff[k] + r0) * P;
            for (int t = tid; t < (r1 - r0) * P; t += kPT) sl[t] = kInf;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < kKB; ++k) {
            if (!f[k]) continue;
            const int r = base[k] + rank[k];
            uint32_t *lo = reinterpret_cast<uint32_t *>(p.tab + (hs[k] & kEntryMask));   // low word
            if (r >= V) { *lo = kRefused; continue; }
            *lo = (uint32_t)(voff[k] + r);
            p.slots[(int64_t)(voff[k] + r) * P] = idx[k];
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
    if (P > 1) {
        for (int t0 = blockIdx.x; t0 < p.ntiles; t0 += kKB * nct) {
            uint32_t hs[kKB];
            int idx[kKB];
#pragma unroll
            for (int k = 0; k < kKB; ++k) {
                const int tile = t0 + k * nct;
                hs[k] = kDropped; idx[k] = -1;
                if (tile < p.ntiles) {
                    const TileRange tr = tile_range(p, tile);
                    if (tr.begin + tid < tr.end) { idx[k] = tr.begin + tid; hs[k] = __ldcg(p.hslot + idx[k]); }
                }
            }
            uint32_t v[kKB];
#pragma unroll
            for (int k = 0; k < kKB; ++k)
                v[k] = (hs[k] != kDropped && !(hs[k] & kFirstBit)) ? (uint32_t)__ldcg(p.tab + (hs[k] & kEntryMask))
                                                                   : kRefused;
            int *sl[kKB];
            int last[kKB], carry[kKB];
#pragma unroll
            for (int k = 0; k < kKB; ++k) {
                sl[k] = p.slots + (int64_t)(v[k] == kRefused ? 0 : v[k]) * P;
                last[k] = v[k] != kRefused ? __ldcg(sl[k] + P - 1) : 0;
            }
#pragma unroll
            for (int k = 0; k < kKB; ++k) carry[k] = (v[k] != kRefused && last[k] > idx[k]) ? idx[k] : kInf;
            for (int q = 1; q < P; ++q) {
                bool any = false;
#pragma unroll
                for (int k = 0; k < kKB; ++k) any |= carry[k] != kInf;
                if (!any) break;
                int old[kKB];
#pragma unroll
                for (int k = 0; k < kKB; ++k)
                    if (carry[k] != kInf) old[k] = atomicMin(sl[k] + q, carry[k]);
#pragma unroll
                for (int k = 0; k < kKB; ++k)
                    if (carry[k] != kInf && old[k] > carry[k]) carry[k] = old[k];   // displaced (kInf: was free)
            }
        }
        if (p.stamps) __syncthreads();
        stamp(p, 12);
        grid.sync();
    }
    stamp(p, 5);

    // ------------------------------------------------------------------------------- P5 gather
    // One CTA handles VB consecutive voxel rows per iteration (VB * P <= kGatherSlots).  Most slots of
    // the padded output are empty (84 % on a nuScenes sweep), so the padding and the points are handled
    // separately: the [VB,P,C] tile is zeroed in shared memory with 16-byte stores, only the KEPT
    // (voxel, slot) rows -- compacted through a block scan of the counts -- gather their point, and the
    // finished tile, which is one contiguous block of the output, leaves with a single bulk (TMA)
    // shared->global copy that overlaps the next iteration.  Thread per (voxel, channel) sums the kept
    // slots in order and divides by the count (HardSimpleVFE; adding the zero padding would not change
    // the sum); thread per voxel re-derives the cell of the first point for coors.
    {
        const int VB = p.VB, F = p.gd.F;
        const int M = s_voff[p.bd.B];
        int *s_idx = reinterpret_cast<int *>(dyn_sm);                       // [VB*P] point index per slot
        float *tile = reinterpret_cast<float *>(s_idx + VB * P);            // [VB*P*C]
        int *s_cnt = reinterpret_cast<int *>(tile + VB * P * C);            // [VB]
        int *s_pre = s_cnt + VB;                                            // [VB] kept rows before the voxel
        unsigned short *s_list = reinterpret_cast<unsigned short *>(s_pre + VB);   // [VB*P] kept rows
        const uint32_t tile_sa = (uint32_t)__cvta_generic_to_shared(tile);
        for (int64_t v0 = (int64_t)blockIdx.x * VB; v0 < M; v0 += (int64_t)nct * VB) {
            const int nv = (int)min((int64_t)VB, M - v0);
            const int ns = nv * P;
            const int nf = ns * C;
            if (tid == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // tile free again
            __syncthreads();
            for (int t = tid; t < (nf + 3) >> 2; t += kPT)
                reinterpret_cast<float4 *>(tile)[t] = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int t = tid; t < ns; t += kPT) s_idx[t] = __ldcg(p.slots + v0 * P + t);
            __syncthreads();
            int kept_total;
            {
                int c = 0;
                if (tid < nv)
                    for (int q = 0; q < P; ++q) c += s_idx[tid * P + q] != kInf ? 1 : 0;
                const int pre = block_excl_scan<kPT>(c, sm, &kept_total);
                if (tid < nv) {
                    s_cnt[tid] = c;
                    s_pre[tid] = pre;
                    p.num_points[v0 + tid] = c;
                    for (int q = 0; q < c; ++q) s_list[pre + q] = (unsigned short)(tid * P + q);
                }
            }
            __syncthreads();
            if constexpr (C == 4) {
                for (int j = tid; j < kept_total; j += kPT) {
                    const int r = s_list[j];
                    const int idx = s_idx[r];
                    float4 x = __ldg(reinterpret_cast<const float4 *>(p.pts) + idx);
                    if constexpr (PERTURB) {
                        const float4 d = __ldg(reinterpret_cast<const float4 *>(p.raw) + idx);
                        x.x = __fadd_rn(x.x, perturb_delta(d.x, p.pd.eb[0], p.pd.fb[0]));
                        x.y = __fadd_rn(x.y, perturb_delta(d.y, p.pd.eb[1], p.pd.fb[1]));
                        x.z = __fadd_rn(x.z, perturb_delta(d.z, p.pd.eb[2], p.pd.fb[2]));
                        x.w = __fadd_rn(x.w, perturb_delta(d.w, p.pd.eb[3], p.pd.fb[3]));
                    }
                    reinterpret_cast<float4 *>(tile)[r] = x;
                    if (p.p2v) p.p2v[idx] = (int)(v0 + r / P);
                    if (p.p2s) p.p2s[idx] = r % P;
                }
            } else {
                // one ELEMENT per lane: a warp instruction touches ~32/C rows (lines) instead of 32
                for (int j = tid; j < kept_total * C; j += kPT) {
                    const int jr = j / C, c = j - jr * C;
                    const int r = s_list[jr];
                    const int idx = s_idx[r];
                    float x = __ldg(p.pts + (int64_t)idx * C + c);
                    if constexpr (PERTURB)
                        x = __fadd_rn(x, perturb_delta(__ldg(p.raw + (int64_t)idx * C + c), p.pd.eb[c], p.pd.fb[c]));
                    tile[r * C + c] = x;
                    if (c == 0) {
                        if (p.p2v) p.p2v[idx] = (int)(v0 + r / P);
                        if (p.p2s) p.p2s[idx] = r % P;
                    }
                }
            }
            __syncthreads();
            if (p.feats) {
                for (int t = tid; t < nv * F; t += kPT) {
                    const int v = t / F, c = t - v * F;
                    const int n = s_cnt[v];
                    float s = 0.0f;
                    for (int q = 0; q < n; ++q) s = __fadd_rn(s, tile[(v * P + q) * C + c]);
                    p.feats[(v0 + v) * F + c] = __fdiv_rn(s, (float)n);
                }
            }
            if (tid < nv) {
                const float *f0 = tile + (int64_t)tid * P * C;            // the voxel's first point
                const int cx = axis_cell(f0[0], p.gd, 0);
                const int cy = axis_cell(f0[1], p.gd, 1);
                const int cz = axis_cell(f0[2], p.gd, 2);
                int b = 0;
                while (b + 1 < p.bd.B && v0 + tid >= s_voff[b + 1]) ++b;
                reinterpret_cast<int4 *>(p.coors)[v0 + tid] = make_int4(b, cz, cy, cx);
            }
            float *out = p.voxels + v0 * P * C;
            if ((nf & 3) == 0) {
                // generic-proxy writes to the tile -> visible to the async proxy, then one bulk store
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                __syncthreads();
                if (tid == 0) {
                    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                                 ::"l"(out), "r"(tile_sa), "r"(nf * 4) : "memory");
                    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                }
            } else {                                                     // ragged tail tile
                for (int t = tid; t < nf; t += kPT) out[t] = tile[t];
            }
        }
        if (tid == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
    if (p.stamps) __syncthreads();
    stamp(p, 6);
}

}  // namespace rpc

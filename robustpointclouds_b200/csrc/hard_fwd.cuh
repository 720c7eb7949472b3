// hard_fwd.cuh -- deterministic hard voxelization of a batch of sweeps (+ optional perturbation and
// HardSimpleVFE mean) as ONE persistent cooperative kernel, sm_100a.
//
// Replaces mmcv `hard_voxelize_forward` with a parallel formulation that gives the SAME
// result as the serial CPU walk.  The path is latency-bound at the named sizes (a KITTI batch is
// ~30 MB of compulsory traffic, ~5 us at HBM speed), so the whole forward is one launch of one CTA
// per SM that walks through six phases separated by grid-wide barriers; all intermediate state
// (hash table, per-point entry, slot lists) stays L2-resident between phases.
//
//   P0 clear     64-bit hash entries := EMPTY (also makes the table L2-resident for the atomics of P1)
//   P1 insert    per point: [perturb] -> cell id -> claim / find the cell's entry with ONE 64-bit CAS
//                ((cell << 32) | point index); an entry already held by the cell with a larger index
//                gets an atomicMin -> entry.lo = index of the FIRST point of the cell
//   P2 flag      per point: first-of-cell flag (entry.lo == index); first points counted per tile
//   P3 assign    exclusive scan over tiles in point order = mmcv's voxel index; rank >= V -> cell
//                refused for good; entry.lo := voxel row; coors; slot list initialised
//   P4 rank      per kept cell the P smallest point indices, by an atomicMin insertion chain (level k
//                keeps the k-th smallest; losers carry to level k+1)
//   P5 gather    per voxel: gather the <= P points in slot order into shared memory, write the
//                zero-padded [P,C] rows coalesced, num_points, point2voxel/slot and the mean
#pragma once
#include <cooperative_groups.h>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace rpc {

constexpr int kPT = 1024;                       // threads per CTA of the persistent kernel
constexpr int kUMax = 8;                        // points per thread per tile (tile = U * kPT points)
constexpr uint32_t kDropped = 0xFFFFFFFFu;      // hslot: point outside the range
constexpr uint32_t kFirstBit = 0x80000000u;     // hslot: point is the first of its cell
constexpr uint32_t kRefused = 0xFFFFFFFFu;      // entry.lo after P3: cell came after the V-th voxel
constexpr unsigned long long kEmpty64 = ~0ull;

struct FwdParams {
    const float *pts, *raw;                     // [ntotal, C]
    unsigned long long *tab;                    // [tab_total] (cell << 32) | first index, later | voxel row
    uint32_t *hslot;                            // [ntotal]   table entry of each point (+ kFirstBit)
    int *tile_cnt;                              // [ntiles]   first points per tile
    int *slots;                                 // [cap * P]  per voxel: point index per slot, kInf empty
    float *voxels; int *coors; int *num_points; int *voxel_num; int *p2v; int *p2s; float *feats;
    unsigned long long *stamps;                 // [8] %globaltimer at the phase boundaries (CTA 0), or NULL
    long long tab_total;
    int ntiles, T, U, VB;
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

__device__ __forceinline__ int block_sum(int v, int *sm)       // sm: 33 ints; all threads get the sum
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, d);
    if (lane == 0) sm[warp] = v;
    __syncthreads();
    if (warp == 0) {
        int w = lane < kPT / 32 ? sm[lane] : 0;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) w += __shfl_xor_sync(0xFFFFFFFFu, w, d);
        if (lane == 0) sm[32] = w;
    }
    __syncthreads();
    const int r = sm[32];
    __syncthreads();
    return r;
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
    r.begin = p.bd.pt_begin[r.b] + (tile - p.bd.tile_begin[r.b]) * p.T;
    r.end = min(r.begin + p.T, p.bd.pt_begin[r.b + 1]);
    return r;
}

template <int C, bool PERTURB>
__global__ void __launch_bounds__(kPT, 1)
k_hard_fwd(const __grid_constant__ FwdParams p)
{
    cg::grid_group grid = cg::this_grid();
    extern __shared__ __align__(16) float tile_sm[];              // P5: [VB*P*C] floats + int cnt[VB]
    __shared__ int sm[33];
    __shared__ int s_total[RPC_MAX_BATCH];
    __shared__ int s_voff[RPC_MAX_BATCH + 1];
    const int tid = threadIdx.x;
    const int P = p.gd.P, V = p.gd.V, U = p.U;

    // ------------------------------------------------------------------------------- P0 clear
    stamp(p, 0);
    {
        uint4 *t4 = reinterpret_cast<uint4 *>(p.tab);
        const long long n4 = p.tab_total >> 1;
        for (long long i = (long long)blockIdx.x * kPT + tid; i < n4; i += (long long)gridDim.x * kPT)
            t4[i] = make_uint4(~0u, ~0u, ~0u, ~0u);
    }
    stamp(p, 8);
    grid.sync();
    stamp(p, 1);

    // ------------------------------------------------------------------------------- P1 insert
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
        const TileRange tr = tile_range(p, tile);
        const uint32_t off = p.bd.tab_off[tr.b], mask = p.bd.tab_mask[tr.b];
        const int shift = p.bd.tab_shift[tr.b];
        for (int k0 = 0; k0 < U; k0 += 4) {
            int cell[4];
            uint32_t s[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int i = tr.begin + (k0 + k) * kPT + tid;
                cell[k] = -1;
                if (k0 + k < U && i < tr.end) {
                    float x, y, z;
                    load_xyz<C, PERTURB>(p.pts, p.raw, p.pd, i, x, y, z);
                    cell[k] = point_cell(x, y, z, p.gd);
                    if (p.p2v) p.p2v[i] = -1;
                    if (p.p2s) p.p2s[i] = -1;
                    if (cell[k] < 0) p.hslot[i] = kDropped;
                }
                s[k] = hash_cell((uint32_t)cell[k], shift);
            }
            bool pend[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) pend[k] = cell[k] >= 0;
            while (pend[0] | pend[1] | pend[2] | pend[3]) {
                unsigned long long old[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const uint32_t i = (uint32_t)(tr.begin + (k0 + k) * kPT + tid);
                    if (pend[k])
                        old[k] = atomicCAS(&p.tab[off + s[k]], kEmpty64,
                                           ((unsigned long long)(uint32_t)cell[k] << 32) | i);
                }
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    if (!pend[k]) continue;
                    const uint32_t i = (uint32_t)(tr.begin + (k0 + k) * kPT + tid);
                    if (old[k] != kEmpty64 && (uint32_t)(old[k] >> 32) != (uint32_t)cell[k]) {
                        s[k] = (s[k] + 1) & mask;                           // collision: next entry
                        continue;
                    }
                    if (old[k] != kEmpty64 && (uint32_t)old[k] > i)
                        atomicMin(&p.tab[off + s[k]], ((unsigned long long)(uint32_t)cell[k] << 32) | i);
                    p.hslot[i] = off + s[k];
                    pend[k] = false;
                }
            }
        }
    }
    stamp(p, 9);
    grid.sync();
    stamp(p, 2);

    // ------------------------------------------------------------------------------- P2 flag
    // thread t owns points begin + t*U .. begin + t*U + U-1 (point order is what the scan needs)
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
        const TileRange tr = tile_range(p, tile);
        const int i0 = tr.begin + tid * U;
        uint32_t hs[kUMax];
#pragma unroll
        for (int k = 0; k < kUMax; ++k)
            hs[k] = (k < U && i0 + k < tr.end) ? __ldcg(p.hslot + i0 + k) : kDropped;
        uint32_t lo[kUMax];
#pragma unroll
        for (int k = 0; k < kUMax; ++k)
            lo[k] = hs[k] != kDropped ? (uint32_t)__ldcg(p.tab + hs[k]) : 0u;
        int c = 0;
#pragma unroll
        for (int k = 0; k < kUMax; ++k)
            if (hs[k] != kDropped && lo[k] == (uint32_t)(i0 + k)) {
                p.hslot[i0 + k] = hs[k] | kFirstBit;
                ++c;
            }
        c = block_sum(c, sm);
        if (tid == 0) p.tile_cnt[tile] = c;
    }
    stamp(p, 10);
    grid.sync();
    stamp(p, 3);

    // ------------------------------------------------------------------------------- P3 assign
    if (tid < RPC_MAX_BATCH) s_total[tid] = 0;
    __syncthreads();
    for (int j = tid; j < p.ntiles; j += kPT) {
        const int c = __ldcg(p.tile_cnt + j);
        if (c) atomicAdd(&s_total[tile_to_sweep(p.bd, j)], c);
    }
    __syncthreads();
    if (tid == 0) {
        int run = 0;
        for (int b = 0; b < p.bd.B; ++b) {
            s_voff[b] = run;
            const int m = min(s_total[b], V);
            if (blockIdx.x == 0) p.voxel_num[b] = m;
            run += m;
        }
        s_voff[p.bd.B] = run;
        if (blockIdx.x == 0) p.voxel_num[p.bd.B] = run;
    }
    __syncthreads();
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
        const TileRange tr = tile_range(p, tile);
        int part = 0;
        for (int j = p.bd.tile_begin[tr.b] + tid; j < tile; j += kPT) part += __ldcg(p.tile_cnt + j);
        const int prefix = block_sum(part, sm);
        const int i0 = tr.begin + tid * U;
        uint32_t hs[kUMax];
        int c = 0;
#pragma unroll
        for (int k = 0; k < kUMax; ++k) {
            hs[k] = (k < U && i0 + k < tr.end) ? __ldcg(p.hslot + i0 + k) : kDropped;
            c += (hs[k] != kDropped && (hs[k] & kFirstBit)) ? 1 : 0;
        }
        int total;
        int rank = prefix + block_excl_scan<kPT>(c, sm, &total);
        if (!c) continue;
        const int voff = s_voff[tr.b];
#pragma unroll
        for (int k = 0; k < kUMax; ++k) {
            if (hs[k] == kDropped || !(hs[k] & kFirstBit)) continue;
            const uint32_t e = hs[k] & ~kFirstBit;
            uint32_t *lo = reinterpret_cast<uint32_t *>(p.tab + e);         // little endian: low word
            const int r = rank++;
            if (r >= V) { *lo = kRefused; continue; }
            const int v = voff + r;
            const int cell = (int)lo[1];
            *lo = (uint32_t)v;
            const int x = cell % p.gd.g[0];
            const int zy = cell / p.gd.g[0];
            reinterpret_cast<int4 *>(p.coors)[v] = make_int4(tr.b, zy / p.gd.g[1], zy % p.gd.g[1], x);
            int *sl = p.slots + (int64_t)v * P;
            sl[0] = i0 + k;
            for (int q = 1; q < P; ++q) sl[q] = kInf;
        }
    }
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
        for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
            const TileRange tr = tile_range(p, tile);
            for (int k0 = 0; k0 < U; k0 += 4) {
                int carry[4];
                int *sl[4];
                uint32_t hs[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const int i = tr.begin + (k0 + k) * kPT + tid;
                    hs[k] = (k0 + k < U && i < tr.end) ? __ldcg(p.hslot + i) : kDropped;
                }
                uint32_t v[4];
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    v[k] = (hs[k] != kDropped && !(hs[k] & kFirstBit)) ? (uint32_t)__ldcg(p.tab + hs[k]) : kRefused;
                int last[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    sl[k] = p.slots + (int64_t)(v[k] == kRefused ? 0 : v[k]) * P;
                    last[k] = v[k] != kRefused ? __ldcg(sl[k] + P - 1) : 0;
                }
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const int i = tr.begin + (k0 + k) * kPT + tid;
                    carry[k] = (v[k] != kRefused && last[k] > i) ? i : kInf;
                }
                for (int q = 1; q < P; ++q) {
                    if (carry[0] == kInf && carry[1] == kInf && carry[2] == kInf && carry[3] == kInf) break;
                    int old[4];
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        if (carry[k] != kInf) old[k] = atomicMin(sl[k] + q, carry[k]);
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        if (carry[k] != kInf && old[k] > carry[k]) carry[k] = old[k];   // displaced (kInf: was free)
                        else if (carry[k] != kInf) { /* carry stays: it lost at this level */ }
                }
            }
        }
        stamp(p, 12);
        grid.sync();
    }
    stamp(p, 5);

    // ------------------------------------------------------------------------------- P5 gather
    // One CTA handles VB consecutive voxel rows per iteration.  (a) thread per (voxel, slot) gathers the
    // point ([perturbed] features) into shared memory, zero for empty slots, and records
    // point2voxel/slot.  (b) the [VB,P,C] tile is contiguous in the output -> coalesced 16-byte stores.
    // (c) thread per (voxel, channel) sums the slots in order and divides by the count (HardSimpleVFE).
    {
        const int VB = p.VB, F = p.gd.F;
        int *cnt = reinterpret_cast<int *>(tile_sm + VB * P * C);
        const int M = s_voff[p.bd.B];
        for (int64_t v0 = (int64_t)blockIdx.x * VB; v0 < M; v0 += (int64_t)gridDim.x * VB) {
            const int nv = (int)min((int64_t)VB, M - v0);
            for (int t = tid; t < VB; t += kPT) cnt[t] = 0;
            __syncthreads();
            for (int t = tid; t < nv * P; t += kPT) {
                const int idx = __ldcg(p.slots + v0 * P + t);
                float *dst = tile_sm + t * C;
                if (idx != kInf) {
                    if constexpr (C == 4) {
                        float4 x = __ldg(reinterpret_cast<const float4 *>(p.pts) + idx);
                        if constexpr (PERTURB) {
                            float4 r = __ldg(reinterpret_cast<const float4 *>(p.raw) + idx);
                            x.x = __fadd_rn(x.x, perturb_delta(r.x, p.pd.eb[0], p.pd.fb[0]));
                            x.y = __fadd_rn(x.y, perturb_delta(r.y, p.pd.eb[1], p.pd.fb[1]));
                            x.z = __fadd_rn(x.z, perturb_delta(r.z, p.pd.eb[2], p.pd.fb[2]));
                            x.w = __fadd_rn(x.w, perturb_delta(r.w, p.pd.eb[3], p.pd.fb[3]));
                        }
                        *reinterpret_cast<float4 *>(dst) = x;
                    } else {
#pragma unroll
                        for (int c = 0; c < C; ++c) {
                            float x = __ldg(p.pts + (int64_t)idx * C + c);
                            if constexpr (PERTURB)
                                x = __fadd_rn(x, perturb_delta(__ldg(p.raw + (int64_t)idx * C + c), p.pd.eb[c], p.pd.fb[c]));
                            dst[c] = x;
                        }
                    }
                    atomicAdd(&cnt[t / P], 1);
                    if (p.p2v) p.p2v[idx] = (int)(v0 + t / P);
                    if (p.p2s) p.p2s[idx] = t % P;
                } else {
#pragma unroll
                    for (int c = 0; c < C; ++c) dst[c] = 0.0f;
                }
            }
            __syncthreads();
            {   // coalesced store of the tile; v0*P*C*4 bytes is 16-byte aligned because VB % 4 == 0
                const int nf = nv * P * C;
                float *out = p.voxels + v0 * P * C;
                const int nf4 = nf >> 2;
                for (int t = tid; t < nf4; t += kPT)
                    reinterpret_cast<float4 *>(out)[t] = reinterpret_cast<const float4 *>(tile_sm)[t];
                for (int t = (nf4 << 2) + tid; t < nf; t += kPT) out[t] = tile_sm[t];
            }
            for (int t = tid; t < nv; t += kPT) p.num_points[v0 + t] = cnt[t];
            if (p.feats) {
                for (int t = tid; t < nv * F; t += kPT) {
                    const int v = t / F, c = t % F;
                    float s = 0.0f;
                    for (int k = 0; k < P; ++k) s = __fadd_rn(s, tile_sm[(v * P + k) * C + c]);
                    p.feats[(v0 + v) * F + c] = __fdiv_rn(s, (float)cnt[v]);
                }
            }
            __syncthreads();
        }
    }
    stamp(p, 6);                                 // CTA 0 only: its own end, not the grid's
}

}  // namespace rpc

// voxelize_hard.cuh -- deterministic hard voxelization of a batch of sweeps, sm_100a.
//
// Replaces mmcv `hard_voxelize_forward` with a parallel formulation that gives the
// SAME result as the serial CPU walk:
//   K1 k_cell_hash      per point: [perturb] -> cell id -> insert into the sweep's open-addressing
//                       table, atomicMin(first[slot], point index)          (first point per cell)
//   K2 k_count_firsts / k_scan_tiles / k_assign_voxels
//                       flag points that are the first of their cell, exclusive scan in point
//                       order = mmcv's voxel index; rank >= V -> cell refused for good
//   K3 k_rank_in_voxel  per kept cell the P smallest point indices, by an atomicMin insertion
//                       chain (level k keeps the k-th smallest; losers carry to level k+1)
//   K4 k_gather_mean    per voxel: gather the <= P points in slot order into shared memory, write
//                       the zero-padded [P,C] rows coalesced, num_points, point2voxel/slot and
//                       the HardSimpleVFE mean in the same pass
#pragma once
#include "common.cuh"

namespace rpc {

struct HardWs {                 // views into the caller's workspace
    uint32_t *keys;             // [tab_total]  cell id or kEmpty
    uint32_t *first;            // [tab_total]  smallest point index of the cell (0xFFFFFFFF none)
    int *vid;                   // [tab_total]  global voxel row of the cell, -1 refused
    int *hslot;                 // [ntotal]     table entry of each point, -1 dropped
    int *tile_count;            // [ntiles]     firsts per tile, then exclusive prefix per sweep
    int *slots;                 // [cap * P]    per voxel: point index per slot, kInf empty
    int *voff;                  // [B + 1]      first voxel row of each sweep
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

// ------------------------------------------------------------------------------------------ K1
template <int C, bool PERTURB>
__global__ void __launch_bounds__(kThreads)
k_cell_hash(const float *__restrict__ pts, const float *__restrict__ raw,
            const __grid_constant__ GridDesc gd, const __grid_constant__ BatchDesc bd,
            const __grid_constant__ PerturbDesc pd, HardWs ws,
            int *__restrict__ p2v, int *__restrict__ p2s)
{
    const int tile = blockIdx.x;
    const int b = tile_to_sweep(bd, tile);
    const int begin = bd.pt_begin[b] + (tile - bd.tile_begin[b]) * kTile;
    const int end = min(begin + kTile, bd.pt_begin[b + 1]);
    const uint32_t off = bd.tab_off[b], mask = bd.tab_mask[b];
    const int shift = bd.tab_shift[b];
#pragma unroll
    for (int k = 0; k < kTile / kThreads; ++k) {
        const int i = begin + k * kThreads + threadIdx.x;
        if (i >= end) break;
        float x, y, z;
        load_xyz<C, PERTURB>(pts, raw, pd, i, x, y, z);
        const int cell = point_cell(x, y, z, gd);
        int entry = -1;
        if (cell >= 0) {
            uint32_t s = hash_cell((uint32_t)cell, shift);
            for (;;) {
                uint32_t old = ws.keys[off + s];
                if (old == kEmpty) old = atomicCAS(&ws.keys[off + s], kEmpty, (uint32_t)cell);
                if (old == kEmpty || old == (uint32_t)cell) break;
                s = (s + 1) & mask;
            }
            entry = (int)(off + s);
            if (ws.first[entry] > (uint32_t)i) atomicMin(&ws.first[entry], (uint32_t)i);
        }
        ws.hslot[i] = entry;
        if (p2v) p2v[i] = -1;
        if (p2s) p2s[i] = -1;
    }
}

// ------------------------------------------------------------------------------------------ K2a
// thread t owns points begin + 4t .. begin + 4t + 3 (point order is what the scan needs)
__device__ __forceinline__ int first_flags(const HardWs &ws, int begin, int end, int e[4])
{
    const int i0 = begin + 4 * threadIdx.x;
    int flags = 0;
    if (i0 + 3 < end && ((i0 & 3) == 0)) {
        int4 v = *reinterpret_cast<const int4 *>(ws.hslot + i0);
        e[0] = v.x; e[1] = v.y; e[2] = v.z; e[3] = v.w;
    } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) e[k] = (i0 + k < end) ? ws.hslot[i0 + k] : -1;
    }
#pragma unroll
    for (int k = 0; k < 4; ++k)
        if (e[k] >= 0 && ws.first[e[k]] == (uint32_t)(i0 + k)) flags |= 1 << k;
    return flags;
}

__global__ void __launch_bounds__(kThreads)
k_count_firsts(const __grid_constant__ BatchDesc bd, HardWs ws)
{
    __shared__ int sm[33];
    const int tile = blockIdx.x;
    const int b = tile_to_sweep(bd, tile);
    const int begin = bd.pt_begin[b] + (tile - bd.tile_begin[b]) * kTile;
    const int end = min(begin + kTile, bd.pt_begin[b + 1]);
    int e[4];
    int c = __popc(first_flags(ws, begin, end, e));
    c = warp_incl_scan(c);
    if ((threadIdx.x & 31) == 31) sm[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
#pragma unroll
        for (int w = 0; w < kThreads / 32; ++w) t += sm[w];
        ws.tile_count[tile] = t;
    }
}

// ------------------------------------------------------------------------------------------ K2b
// one CTA: per sweep, exclusive scan of the tile counts (in place); voxel_num[b] = min(firsts, V),
// voff = exclusive scan of voxel_num, voxel_num[B] = total.
constexpr int kScanThreads = 1024;
__global__ void __launch_bounds__(kScanThreads)
k_scan_tiles(const __grid_constant__ BatchDesc bd, HardWs ws, int V, int *__restrict__ voxel_num)
{
    __shared__ int sm[33];
    __shared__ int s_run;
    int run_vox = 0;
    for (int b = 0; b < bd.B; ++b) {
        const int t0 = bd.tile_begin[b], t1 = bd.tile_begin[b + 1];
        int carry = 0;
        for (int base = t0; base < t1; base += kScanThreads) {
            const int t = base + threadIdx.x;
            const int v = t < t1 ? ws.tile_count[t] : 0;
            int total;
            const int ex = block_excl_scan<kScanThreads>(v, sm, &total);
            if (t < t1) ws.tile_count[t] = carry + ex;
            carry += total;
        }
        if (threadIdx.x == 0) {
            const int m = min(carry, V);
            ws.voff[b] = run_vox;
            voxel_num[b] = m;
            s_run = run_vox + m;
        }
        __syncthreads();
        run_vox = s_run;
        __syncthreads();
    }
    if (threadIdx.x == 0) { ws.voff[bd.B] = run_vox; voxel_num[bd.B] = run_vox; }
}

// ------------------------------------------------------------------------------------------ K2c
// rank of each first point = mmcv's voxel index; publish the voxel row in the table, write coors
// (batch, z, y, x) and initialise the voxel's slot list (slot 0 = the first point).
__global__ void __launch_bounds__(kThreads)
k_assign_voxels(const __grid_constant__ GridDesc gd, const __grid_constant__ BatchDesc bd, HardWs ws,
                int *__restrict__ coors)
{
    __shared__ int sm[33];
    const int tile = blockIdx.x;
    const int b = tile_to_sweep(bd, tile);
    const int begin = bd.pt_begin[b] + (tile - bd.tile_begin[b]) * kTile;
    const int end = min(begin + kTile, bd.pt_begin[b + 1]);
    int e[4];
    const int flags = first_flags(ws, begin, end, e);
    int total;
    int rank = ws.tile_count[tile] + block_excl_scan<kThreads>(__popc(flags), sm, &total);
    if (!flags) return;
    const int voff = ws.voff[b];
    const int P = gd.P;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        if (!(flags & (1 << k))) continue;
        const int r = rank++;
        if (r >= gd.V) { ws.vid[e[k]] = -1; continue; }
        const int v = voff + r;
        ws.vid[e[k]] = v;
        const int cell = (int)ws.keys[e[k]];
        const int x = cell % gd.g[0];
        const int zy = cell / gd.g[0];
        reinterpret_cast<int4 *>(coors)[v] = make_int4(b, zy / gd.g[1], zy % gd.g[1], x);
        int *sl = ws.slots + (int64_t)v * P;
        sl[0] = begin + 4 * threadIdx.x + k;
        for (int s = 1; s < P; ++s) sl[s] = kInf;
    }
}

// ------------------------------------------------------------------------------------------ K3
// Per kept cell, the P smallest point indices in increasing order.  Level s of the voxel's list
// only ever decreases (atomicMin); whatever loses at level s -- the arriving index or the one it
// displaces -- carries to level s+1.  Every index but the final minimum of a level is carried
// exactly once, so level s ends up holding the (s+1)-th smallest index regardless of interleaving.
// Slot 0 is the first point (already known).  An index larger than the current last level can
// never enter the list (levels only decrease), which prunes crowded voxels.
__global__ void __launch_bounds__(kThreads)
k_rank_in_voxel(const __grid_constant__ BatchDesc bd, HardWs ws, int P)
{
    const int tile = blockIdx.x;
    const int b = tile_to_sweep(bd, tile);
    const int begin = bd.pt_begin[b] + (tile - bd.tile_begin[b]) * kTile;
    const int end = min(begin + kTile, bd.pt_begin[b + 1]);
#pragma unroll
    for (int k = 0; k < kTile / kThreads; ++k) {
        const int i = begin + k * kThreads + threadIdx.x;
        if (i >= end) break;
        const int entry = ws.hslot[i];
        if (entry < 0) continue;
        const int v = ws.vid[entry];
        if (v < 0) continue;
        int *sl = ws.slots + (int64_t)v * P;
        if (ws.first[entry] == (uint32_t)i) continue;                       // slot 0
        if (*(volatile int *)(sl + P - 1) < i) continue;          // list already full of smaller
        int carry = i;
        for (int s = 1; s < P && carry != kInf; ++s) {
            const int old = atomicMin(sl + s, carry);
            if (old > carry) carry = old;                         // displaced (kInf: slot was free)
        }
    }
}

// ------------------------------------------------------------------------------------------ K4
// One CTA handles VB consecutive voxel rows.  Phase 1: thread per (voxel, slot) gathers the point
// ([perturbed] features) into shared memory, zero for empty slots, and records point2voxel/slot.
// Phase 2: the [VB, P, C] tile is contiguous in the output -> coalesced copy.  Phase 3: thread per
// (voxel, channel) sums the slots in order and divides by the count (HardSimpleVFE).
template <int C, bool PERTURB>
__global__ void __launch_bounds__(kThreads)
k_gather_mean(const float *__restrict__ pts, const float *__restrict__ raw,
              const __grid_constant__ PerturbDesc pd, HardWs ws, int nB, int P, int F, int VB,
              float *__restrict__ voxels, int *__restrict__ num_points, float *__restrict__ feats,
              int *__restrict__ p2v, int *__restrict__ p2s)
{
    extern __shared__ __align__(16) float tile[];                 // [VB*P*C] + int cnt[VB]
    int *cnt = reinterpret_cast<int *>(tile + VB * P * C);
    const int M = ws.voff[nB];
    for (int64_t v0 = (int64_t)blockIdx.x * VB; v0 < M; v0 += (int64_t)gridDim.x * VB) {
        const int nv = (int)min((int64_t)VB, M - v0);
        for (int t = threadIdx.x; t < VB; t += kThreads) cnt[t] = 0;
        __syncthreads();
        for (int t = threadIdx.x; t < nv * P; t += kThreads) {
            const int idx = ws.slots[v0 * P + t];
            float *dst = tile + t * C;
            if (idx != kInf) {
                if constexpr (C == 4) {
                    float4 x = __ldg(reinterpret_cast<const float4 *>(pts) + idx);
                    if constexpr (PERTURB) {
                        float4 r = __ldg(reinterpret_cast<const float4 *>(raw) + idx);
                        x.x = __fadd_rn(x.x, perturb_delta(r.x, pd.eb[0], pd.fb[0]));
                        x.y = __fadd_rn(x.y, perturb_delta(r.y, pd.eb[1], pd.fb[1]));
                        x.z = __fadd_rn(x.z, perturb_delta(r.z, pd.eb[2], pd.fb[2]));
                        x.w = __fadd_rn(x.w, perturb_delta(r.w, pd.eb[3], pd.fb[3]));
                    }
                    *reinterpret_cast<float4 *>(dst) = x;
                } else {
#pragma unroll
                    for (int c = 0; c < C; ++c) {
                        float x = __ldg(pts + (int64_t)idx * C + c);
                        if constexpr (PERTURB)
                            x = __fadd_rn(x, perturb_delta(__ldg(raw + (int64_t)idx * C + c), pd.eb[c], pd.fb[c]));
                        dst[c] = x;
                    }
                }
                atomicAdd(&cnt[t / P], 1);
                if (p2v) p2v[idx] = (int)(v0 + t / P);
                if (p2s) p2s[idx] = t % P;
            } else {
#pragma unroll
                for (int c = 0; c < C; ++c) dst[c] = 0.0f;
            }
        }
        __syncthreads();
        {   // coalesced store of the tile; v0*P*C*4 bytes is 16-byte aligned when VB % 4 == 0
            const int nf = nv * P * C;
            float *out = voxels + v0 * P * C;
            const int nf4 = nf >> 2;
            for (int t = threadIdx.x; t < nf4; t += kThreads)
                reinterpret_cast<float4 *>(out)[t] = reinterpret_cast<const float4 *>(tile)[t];
            for (int t = (nf4 << 2) + threadIdx.x; t < nf; t += kThreads) out[t] = tile[t];
        }
        for (int t = threadIdx.x; t < nv; t += kThreads) num_points[v0 + t] = cnt[t];
        if (feats) {
            for (int t = threadIdx.x; t < nv * F; t += kThreads) {
                const int v = t / F, c = t % F;
                float s = 0.0f;
                for (int k = 0; k < P; ++k) s = __fadd_rn(s, tile[(v * P + k) * C + c]);
                feats[(v0 + v) * F + c] = __fdiv_rn(s, (float)cnt[v]);
            }
        }
        __syncthreads();
    }
}

}  // namespace rpc

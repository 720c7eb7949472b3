// common.cuh -- shared device helpers for the sm_100a hot path.
//
// Arithmetic rules: every fp32 operation that decides a voxel cell is a single
// IEEE round-to-nearest operation -- __fsub_rn / __fdiv_rn / __fmul_rn / __fadd_rn -- so that no
// FMA contraction or reciprocal-multiply can move a point across a cell boundary.  The file is
// compiled WITHOUT -use_fast_math.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/rpc_b200.h"

namespace rpc {

constexpr int kThreads = 256;        // threads per CTA of the streaming kernels
constexpr uint32_t kEmpty = 0xFFFFFFFFu;
constexpr int kInf = 0x7FFFFFFF;     // "no point" in the per-voxel slot lists

// voxel grid, passed by value in kernel parameter space
struct GridDesc {
    float lo[3];      // x, y, z lower bounds
    float vs[3];      // voxel sizes
    float rvs[3];     // RN(1 / vs): quotient estimate of axis_cell
    float gf[3];      // grid extents as float (comparison is done in fp32 like mmcv's int compare)
    int g[3];         // grid extents gx, gy, gz
    int P, V, C, F;
};

// batch of sweeps: points of sweep b are rows [pt_begin[b], pt_begin[b+1]) of the concatenated
// array; tiles never straddle two sweeps.
struct BatchDesc {
    int B;
    int pt_begin[RPC_MAX_BATCH + 1];
    int tile_begin[RPC_MAX_BATCH + 1];
    uint32_t tab_off[RPC_MAX_BATCH];   // first entry of the sweep's hash sub-table
    uint32_t tab_size[RPC_MAX_BATCH];  // its entries (even, >= 1.5 x points of the sweep)
};

struct PerturbDesc {
    float eb[RPC_MAX_CHANNELS];
    float fb[RPC_MAX_CHANNELS];
};

// which sweep a tile belongs to (B <= 32: a short uniform loop over parameter space)
__device__ __forceinline__ int tile_to_sweep(const BatchDesc &bd, int tile)
{
    int b = 0;
#pragma unroll 1
    while (b + 1 < bd.B && tile >= bd.tile_begin[b + 1]) ++b;
    return b;
}

// VoxelPerturber tail, one channel: p = clamp(raw*eb, -fb, +fb), NaN -> 0   (voxel_perturber.py:256,
// 357-365).  torch.clamp == min(max(x, lo), hi) with NaN propagated, then nan_to_num.
__device__ __forceinline__ float perturb_delta(float raw, float eb, float fb)
{
    float p = __fmul_rn(raw, eb);
    if (p != p) return 0.0f;
    p = p < -fb ? -fb : p;
    p = p > fb ? fb : p;
    return p;
}
// clamp passes the gradient where -fb <= raw*eb <= fb, inclusive (SURVEY a-3)
__device__ __forceinline__ bool perturb_pass(float raw, float eb, float fb)
{
    float p = __fmul_rn(raw, eb);
    return (p >= -fb) && (p <= fb);
}

// one coordinate -> cell index along axis d, or -1 (mmcv dynamic_voxelize_forward_cpu_kernel:
// c = floor((p - lo) / vs); fail if c < 0 or c >= grid; NaN fails).
//
// The IEEE division is a ~40-instruction routine and three of them per point are a visible share of the
// point-parallel phases, so the quotient is first estimated as q = (p - lo) * RN(1 / vs).  q differs from
// RN((p - lo) / vs) by at most 3 * 2^-24 |q|; whenever q is farther than 4 * 2^-23 |q| from an integer the
// two floors are equal and the estimate is used, otherwise (and for NaN / inf / |q| >= 2^22) the true
// division decides.  Results are bit-identical to the division-only form.
__device__ __forceinline__ int axis_cell(float p, const GridDesc &gd, int d)
{
    const float a = __fsub_rn(p, gd.lo[d]);
    const float q = __fmul_rn(a, gd.rvs[d]);
    float f = floorf(q);
    const float frac = __fsub_rn(q, f);                         // exact, in [0, 1)
    const float tol = __fmul_rn(fabsf(q), 4.8e-7f);
    if (!(frac > tol && __fsub_rn(1.0f, frac) > tol) || !(fabsf(q) < 4194304.0f))
        f = floorf(__fdiv_rn(a, gd.vs[d]));
    if (!(f >= 0.0f) || !(f < gd.gf[d])) return -1;
    return (int)f;
}

// (x,y,z) -> linear cell id (z*gy + y)*gx + x, or -1 when the point is dropped
__device__ __forceinline__ int point_cell(float x, float y, float z, const GridDesc &gd)
{
    int cx = axis_cell(x, gd, 0);
    int cy = axis_cell(y, gd, 1);
    int cz = axis_cell(z, gd, 2);
    if ((cx | cy | cz) < 0) return -1;
    return (cz * gd.g[1] + cy) * gd.g[0] + cx;
}

// warp / block exclusive scan of small integers (kThreads threads)
__device__ __forceinline__ int warp_incl_scan(int v)
{
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int t = __shfl_up_sync(0xFFFFFFFFu, v, d);
        if (lane >= d) v += t;
    }
    return v;
}

// returns the exclusive prefix of v over the CTA; *total receives the CTA sum.  smem: 33 ints.
template <int NT>
__device__ __forceinline__ int block_excl_scan(int v, int *smem, int *total)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int incl = warp_incl_scan(v);
    if (lane == 31) smem[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        int w = lane < NT / 32 ? smem[lane] : 0;
        int wi = warp_incl_scan(w);
        smem[lane] = wi - w;
        if (lane == 31) smem[32] = wi;
    }
    __syncthreads();
    int r = smem[warp] + incl - v;
    *total = smem[32];
    __syncthreads();
    return r;
}

}  // namespace rpc

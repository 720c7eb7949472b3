// rows.cuh -- the reference-order composition on the padded voxel tensor:
//   k_valid_rows      extract_feat's padding mask + boolean gather
//                     (/root/reference/models/detectors/adversarial_voxelnet.py:86-90,
//                      adversarial_centerpoint.py:73-77): valid = row_sum != 0 ('rowsum', the reference
//                     quirk) or slot < num_points ('count'); valid rows compacted in row order
//   k_scatter_rows    the gradient-connected copy + masked scatter-back (adversarial_voxelnet.py:113-117)
//   k_gather_rows     its autograd (rows of the gradient picked by the same index)
//   k_perturb_apply_* VoxelPerturber's elementwise tail and its autograd
//                     (/root/reference/models/adversarial/voxel_perturber.py:256, 357-365, 285)
#pragma once
#include <cooperative_groups.h>

#include "common.cuh"

namespace rpc {

constexpr int kRowTile = 1024;                  // rows per tile of the compaction

struct RangeDesc { float lo[3], hi[3]; };       // mode 2: open box, mmdet3d BasePoints.in_range_3d

// row r of the [M*P, C] view is valid?  rowsum: fp32 sum left to right != 0 (torch sum over 4-5 floats)
__device__ __forceinline__ bool row_valid(const float *__restrict__ vox, const int *__restrict__ num,
                                          int64_t r, int P, int C, int mode, const RangeDesc &rd)
{
    if (mode == 1) return (int)(r % P) < __ldg(num + r / P);
    const float *x = vox + r * C;
    if (mode == 2) {                                            // NaN compares false: dropped, like the ATen mask
        const float a = __ldg(x), b = __ldg(x + 1), c = __ldg(x + 2);
        return a > rd.lo[0] && b > rd.lo[1] && c > rd.lo[2] && a < rd.hi[0] && b < rd.hi[1] && c < rd.hi[2];
    }
    float s = __ldg(x);
    for (int c = 1; c < C; ++c) s = __fadd_rn(s, __ldg(x + c));
    return s != 0.0f;
}

// One cooperative launch: (1) per-tile valid counts, grid barrier, (2) every CTA scans the tile counts
// in shared memory and writes its tiles' rows.  row_idx[k] = k-th valid row, row2k[r] = k or -1,
// rows_out[k, :] = the row itself, count[0] = K'.
__global__ void __launch_bounds__(kRowTile)
k_valid_rows(const float *__restrict__ vox, const int *__restrict__ num, int64_t M, const int *__restrict__ m_dev,
             int P, int C, int mode, int ntiles, int *__restrict__ tile_cnt, int *__restrict__ row_idx,
             int *__restrict__ row2k, float *__restrict__ rows_out, int *__restrict__ count,
             const __grid_constant__ RangeDesc rd)
{
    namespace cg = cooperative_groups;
    extern __shared__ int prefix[];                               // [ntiles + 1]
    __shared__ int sm[33];
    if (m_dev) M = min(M, (int64_t)*m_dev);
    const int64_t R = M * P;
    const int tid = threadIdx.x;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int64_t r = (int64_t)tile * kRowTile + tid;
        const bool v = r < R && row_valid(vox, num, r, P, C, mode, rd);
        const int c = __syncthreads_count(v);
        if (tid == 0) tile_cnt[tile] = c;
    }
    cg::this_grid().sync();
    {
        const int per = (ntiles + kRowTile - 1) / kRowTile;
        const int j0 = tid * per, j1 = min(j0 + per, ntiles);
        int s = 0;
        for (int j = j0; j < j1; ++j) s += __ldcg(tile_cnt + j);
        int tot;
        int run = block_excl_scan<kRowTile>(s, sm, &tot);
        for (int j = j0; j < j1; ++j) { prefix[j] = run; run += __ldcg(tile_cnt + j); }
        if (tid == 0 && blockIdx.x == 0) count[0] = tot;
        __syncthreads();
    }
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int64_t r = (int64_t)tile * kRowTile + tid;
        const bool v = r < R && row_valid(vox, num, r, P, C, mode, rd);
        int tot;
        const int k = prefix[tile] + block_excl_scan<kRowTile>(v ? 1 : 0, sm, &tot);
        if (r < R && row2k) row2k[r] = v ? k : -1;
        if (v) {
            row_idx[k] = (int)r;
            if (rows_out)
                for (int c = 0; c < C; ++c) rows_out[(int64_t)k * C + c] = __ldg(vox + r * C + c);
        }
    }
}

// out = vox; out[row_idx[k], :] = rows[k, :]      (element per thread, k looked up through row2k)
__global__ void __launch_bounds__(kThreads)
k_scatter_rows(const float *__restrict__ vox, const int *__restrict__ row2k, const float *__restrict__ rows,
               int64_t R, int C, float *__restrict__ out)
{
    const int64_t total = R * C;
    for (int64_t e = (int64_t)blockIdx.x * kThreads + threadIdx.x; e < total; e += (int64_t)gridDim.x * kThreads) {
        const int64_t r = e / C;
        const int k = __ldg(row2k + r);
        out[e] = k >= 0 ? __ldg(rows + (int64_t)k * C + (e - r * C)) : __ldg(vox + e);
    }
}

// out[k, :] = src[row_idx[k], :]
__global__ void __launch_bounds__(kThreads)
k_gather_rows(const float *__restrict__ src, const int *__restrict__ row_idx, int64_t K, const int *__restrict__ k_dev,
              int C, float *__restrict__ out)
{
    if (k_dev) K = min(K, (int64_t)*k_dev);
    const int64_t total = K * C;
    for (int64_t e = (int64_t)blockIdx.x * kThreads + threadIdx.x; e < total; e += (int64_t)gridDim.x * kThreads) {
        const int64_t k = e / C;
        out[e] = __ldg(src + (int64_t)__ldg(row_idx + k) * C + (e - k * C));
    }
}

// p = nan_to_num(clamp(raw * eb, -fb, fb)); y = x + p          (p_out optional)
__global__ void __launch_bounds__(kThreads)
k_perturb_apply_fwd(const float *__restrict__ x, const float *__restrict__ raw,
                    const __grid_constant__ PerturbDesc pd, int64_t K, const int *__restrict__ k_dev, int C,
                    float *__restrict__ y, float *__restrict__ p_out)
{
    if (k_dev) K = min(K, (int64_t)*k_dev);
    const int64_t total = K * C;
    for (int64_t e = (int64_t)blockIdx.x * kThreads + threadIdx.x; e < total; e += (int64_t)gridDim.x * kThreads) {
        const int c = (int)(e % C);
        const float p = perturb_delta(__ldg(raw + e), pd.eb[c], pd.fb[c]);
        if (p_out) p_out[e] = p;
        y[e] = __fadd_rn(__ldg(x + e), p);
    }
}

// grad_raw = (-fb <= raw*eb <= fb) ? grad_p * eb : 0
__global__ void __launch_bounds__(kThreads)
k_perturb_apply_bwd(const float *__restrict__ grad_p, const float *__restrict__ raw,
                    const __grid_constant__ PerturbDesc pd, int64_t K, const int *__restrict__ k_dev, int C,
                    float *__restrict__ grad_raw)
{
    if (k_dev) K = min(K, (int64_t)*k_dev);
    const int64_t total = K * C;
    for (int64_t e = (int64_t)blockIdx.x * kThreads + threadIdx.x; e < total; e += (int64_t)gridDim.x * kThreads) {
        const int c = (int)(e % C);
        grad_raw[e] = perturb_pass(__ldg(raw + e), pd.eb[c], pd.fb[c]) ? __fmul_rn(__ldg(grad_p + e), pd.eb[c]) : 0.0f;
    }
}

// voxel_centers of Det3DDataPreprocessor.voxelize: (coors[:, [x, y, z]] + 0.5) * voxel_size + range_min, the same
// fp32 operations in the same order as the ATen expression (separate multiply and add)
struct CenterDesc { float vs[3], lo[3]; };
__global__ void __launch_bounds__(kThreads)
k_voxel_centers(const int *__restrict__ coors, int ncol, int64_t M, const int *__restrict__ m_dev,
                const __grid_constant__ CenterDesc cd, float *__restrict__ out)
{
    if (m_dev) M = min(M, (int64_t)*m_dev);
    for (int64_t e = (int64_t)blockIdx.x * kThreads + threadIdx.x; e < M * 3; e += (int64_t)gridDim.x * kThreads) {
        const int64_t v = e / 3;
        const int d = (int)(e - v * 3);                              // 0: x, 1: y, 2: z; coors rows end in (z, y, x)
        const float c = (float)__ldg(coors + v * ncol + (ncol - 1 - d));
        out[e] = __fadd_rn(__fmul_rn(__fadd_rn(c, 0.5f), cd.vs[d]), cd.lo[d]);
    }
}

}  // namespace rpc

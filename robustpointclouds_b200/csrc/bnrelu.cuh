// bnrelu.cuh -- training-mode BatchNorm1d + ReLU of the perturber's MLP (third slice of SURVEY 8f row 2):
//   z = relu((y - mean_k y) / sqrt(var_k y + eps) * gamma + beta)          nn.BatchNorm1d(h) -> nn.ReLU of
//                                                                         /root/reference/models/adversarial/voxel_perturber.py:84-98
// ATen runs this as statistics + transform + ReLU forward and threshold + reduce + elementwise backward: six passes that
// move 12 [K, H] tensors through HBM.  Here: moments (read y) and transform (read y, write z) forward; reduce (read g, y)
// and elementwise (read g, y, write grad_y) backward -- the ReLU mask is recomputed from y, not stored -- 8 tensors.
// Thread = (row group, column) over the row-major [K, H] matrix, H <= 128: a warp reads 128 contiguous bytes, per-column
// values sit in registers, sums are fp64 (per thread, block tree, one atomicAdd per block and column).
#pragma once
#include "common.cuh"

namespace rpc {

constexpr int kBnThreads = 256;
constexpr int kBnMaxH = 128;
constexpr int kBnU = 4;                         // rows in flight per thread

// sum over the row groups of a block, then one atomicAdd(double) per column and value
__device__ __forceinline__ void bn_block_add2(double a, double b, int H, int groups, double *acc)
{
    __shared__ double sh[kBnThreads][2];
    const int tid = threadIdx.x;
    sh[tid][0] = a; sh[tid][1] = b;
    __syncthreads();
    if (tid < H) {
        double s0 = 0.0, s1 = 0.0;
        for (int q = 0; q < groups; ++q) { s0 += sh[q * H + tid][0]; s1 += sh[q * H + tid][1]; }
        atomicAdd(acc + tid, s0);
        atomicAdd(acc + H + tid, s1);
    }
}

// acc[0..H) += sum_k y, acc[H..2H) += sum_k y^2
__global__ void __launch_bounds__(kBnThreads)
k_bn_moments(const float *__restrict__ y, int64_t K, int H, double *__restrict__ acc)
{
    const int groups = kBnThreads / H, tid = threadIdx.x, r = tid / H, j = tid - r * H;
    double s = 0.0, q = 0.0;
    if (r < groups) {
        const int64_t stride = (int64_t)gridDim.x * groups;
        for (int64_t k0 = (int64_t)blockIdx.x * groups + r; k0 < K; k0 += kBnU * stride) {
            float v[kBnU];
#pragma unroll
            for (int u = 0; u < kBnU; ++u) {
                const int64_t k = k0 + u * stride;
                v[u] = k < K ? __ldg(y + k * H + j) : 0.0f;
            }
#pragma unroll
            for (int u = 0; u < kBnU; ++u) { s += (double)v[u]; q += (double)v[u] * (double)v[u]; }
        }
    }
    bn_block_add2(s, q, H, groups, acc);
}

// mean / invstd from the moments (every block, for its own use); block 0 also saves them and updates the running
// statistics like torch (biased variance for the normalisation, unbiased for running_var); then the transform + ReLU
__global__ void __launch_bounds__(kBnThreads)
k_bn_relu_fwd(const float *__restrict__ y, const double *__restrict__ acc, int64_t K, int H,
              const float *__restrict__ gamma, const float *__restrict__ beta, float eps, float momentum,
              float *running_mean, float *running_var, float *save_mean, float *save_invstd, float *__restrict__ z)
{
    const int groups = kBnThreads / H, tid = threadIdx.x, r = tid / H, j = tid - r * H;
    if (r >= groups) return;
    const double mean_d = acc[j] / (double)K;
    double var_d = acc[H + j] / (double)K - mean_d * mean_d;
    if (var_d < 0.0) var_d = 0.0;
    const float mean = (float)mean_d, invstd = (float)(1.0 / sqrt(var_d + (double)eps));
    if (blockIdx.x == 0 && r == 0) {
        save_mean[j] = mean;
        save_invstd[j] = invstd;
        if (running_mean) running_mean[j] = (1.0f - momentum) * running_mean[j] + momentum * mean;
        if (running_var) {
            const double unb = K > 1 ? var_d * (double)K / (double)(K - 1) : var_d;
            running_var[j] = (1.0f - momentum) * running_var[j] + momentum * (float)unb;
        }
    }
    const float ga = __ldg(gamma + j), be = __ldg(beta + j);
    const int64_t stride = (int64_t)gridDim.x * groups;
    for (int64_t k0 = (int64_t)blockIdx.x * groups + r; k0 < K; k0 += kBnU * stride) {
        float v[kBnU];
#pragma unroll
        for (int u = 0; u < kBnU; ++u) {
            const int64_t k = k0 + u * stride;
            v[u] = k < K ? __ldg(y + k * H + j) : 0.0f;
        }
#pragma unroll
        for (int u = 0; u < kBnU; ++u) {
            const int64_t k = k0 + u * stride;
            if (k < K) {
                const float t = __fadd_rn(__fmul_rn(__fmul_rn(__fsub_rn(v[u], mean), invstd), ga), be);
                z[k * H + j] = t < 0.0f ? 0.0f : t;                  // NaN goes through, like ATen's relu
            }
        }
    }
}

// acc[0..H) += sum_k g', acc[H..2H) += sum_k g' * xhat, with g' = g where the ReLU was open (recomputed from y)
__global__ void __launch_bounds__(kBnThreads)
k_bn_relu_bwd_reduce(const float *__restrict__ g, const float *__restrict__ y, const float *__restrict__ save_mean,
                     const float *__restrict__ save_invstd, const float *__restrict__ gamma,
                     const float *__restrict__ beta, int64_t K, int H, double *__restrict__ acc)
{
    const int groups = kBnThreads / H, tid = threadIdx.x, r = tid / H, j = tid - r * H;
    double s1 = 0.0, s2 = 0.0;
    if (r < groups) {
        const float mean = __ldg(save_mean + j), invstd = __ldg(save_invstd + j), ga = __ldg(gamma + j), be = __ldg(beta + j);
        const int64_t stride = (int64_t)gridDim.x * groups;
        for (int64_t k0 = (int64_t)blockIdx.x * groups + r; k0 < K; k0 += kBnU * stride) {
            float v[kBnU], gv[kBnU];
#pragma unroll
            for (int u = 0; u < kBnU; ++u) {
                const int64_t k = k0 + u * stride;
                v[u] = k < K ? __ldg(y + k * H + j) : 0.0f;
                gv[u] = k < K ? __ldg(g + k * H + j) : 0.0f;
            }
#pragma unroll
            for (int u = 0; u < kBnU; ++u) {
                const float xh = __fmul_rn(__fsub_rn(v[u], mean), invstd);
                const float t = __fadd_rn(__fmul_rn(xh, ga), be);
                const float gp = t <= 0.0f ? 0.0f : gv[u];            // threshold_backward's rule (NaN passes)
                s1 += (double)gp;
                s2 += (double)gp * (double)xh;
            }
        }
    }
    bn_block_add2(s1, s2, H, groups, acc);
}

// grad_y = gamma * invstd * (g' - mean_k g' - xhat * mean_k(g' xhat)); block 0 writes grad_gamma = sum g' xhat,
// grad_beta = sum g'
__global__ void __launch_bounds__(kBnThreads)
k_bn_relu_bwd_elemt(const float *__restrict__ g, const float *__restrict__ y, const float *__restrict__ save_mean,
                    const float *__restrict__ save_invstd, const float *__restrict__ gamma,
                    const float *__restrict__ beta, const double *__restrict__ acc, int64_t K, int H,
                    float *__restrict__ grad_y, float *grad_gamma, float *grad_beta)
{
    const int groups = kBnThreads / H, tid = threadIdx.x, r = tid / H, j = tid - r * H;
    if (r >= groups) return;
    const float mean = __ldg(save_mean + j), invstd = __ldg(save_invstd + j), ga = __ldg(gamma + j), be = __ldg(beta + j);
    const float m1 = (float)(acc[j] / (double)K), m2 = (float)(acc[H + j] / (double)K);
    if (blockIdx.x == 0 && r == 0) {
        if (grad_beta) grad_beta[j] = (float)acc[j];
        if (grad_gamma) grad_gamma[j] = (float)acc[H + j];
    }
    const float sc = ga * invstd;
    const int64_t stride = (int64_t)gridDim.x * groups;
    for (int64_t k0 = (int64_t)blockIdx.x * groups + r; k0 < K; k0 += kBnU * stride) {
        float v[kBnU], gv[kBnU];
#pragma unroll
        for (int u = 0; u < kBnU; ++u) {
            const int64_t k = k0 + u * stride;
            v[u] = k < K ? __ldg(y + k * H + j) : 0.0f;
            gv[u] = k < K ? __ldg(g + k * H + j) : 0.0f;
        }
#pragma unroll
        for (int u = 0; u < kBnU; ++u) {
            const int64_t k = k0 + u * stride;
            if (k < K) {
                const float xh = __fmul_rn(__fsub_rn(v[u], mean), invstd);
                const float t = __fadd_rn(__fmul_rn(xh, ga), be);
                const float gp = t <= 0.0f ? 0.0f : gv[u];            // threshold_backward's rule (NaN passes)
                grad_y[k * H + j] = sc * (gp - m1 - xh * m2);
            }
        }
    }
}

}  // namespace rpc

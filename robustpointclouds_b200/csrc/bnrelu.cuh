// bnrelu.cuh -- training-mode BatchNorm1d + ReLU of the perturber's MLP (third slice of SURVEY 8f row 2):
//   z = relu((y - mean_k y) / sqrt(var_k y + eps) * gamma + beta)          nn.BatchNorm1d(h) -> nn.ReLU of
//                                                                         /root/reference/models/adversarial/voxel_perturber.py:84-98
// ATen runs this as statistics + transform + ReLU forward and threshold + reduce + elementwise backward: six passes that
// move 12 [K, H] tensors through HBM.  Here: moments (read y) and transform (read y, write z) forward; reduce (read g, y)
// and elementwise (read g, y, write grad_y) backward -- the ReLU mask is recomputed from y, not stored -- 8 tensors.
// Thread = (row group, V columns) over the row-major [K, H] matrix, H <= 128 (V = 4: 16-byte loads when H % 4 == 0),
// per-column values sit in registers, sums are fp64 (per thread, block tree, one atomicAdd per block and column).
#pragma once
#include "common.cuh"

namespace rpc {

constexpr int kBnThreads = 256;
constexpr int kBnMaxH = 128;
constexpr int kBnU = 4;                         // rows in flight per thread

// V consecutive columns of row k starting at column j (V = 4: one 16-byte load; rows beyond K read as 0)
template <int V>
__device__ __forceinline__ void bn_load(const float *__restrict__ m, int64_t k, int64_t K, int H, int j, float (&v)[V])
{
    if (k < K) {
        if constexpr (V == 4) {
            const float4 t = __ldg(reinterpret_cast<const float4 *>(m + k * H + j));
            v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
        } else {
#pragma unroll
            for (int c = 0; c < V; ++c) v[c] = __ldg(m + k * H + j + c);
        }
    } else {
#pragma unroll
        for (int c = 0; c < V; ++c) v[c] = 0.0f;
    }
}

template <int V>
__device__ __forceinline__ void bn_store(float *__restrict__ m, int64_t k, int H, int j, const float (&v)[V])
{
    if constexpr (V == 4) {
        *reinterpret_cast<float4 *>(m + k * H + j) = make_float4(v[0], v[1], v[2], v[3]);
    } else {
#pragma unroll
        for (int c = 0; c < V; ++c) m[k * H + j + c] = v[c];
    }
}

// sum over the row groups of a block, then one atomicAdd(double) per column and value; HC = H / V column groups
template <int V>
__device__ __forceinline__ void bn_block_add2(const double (&a)[V], const double (&b)[V], int H, int HC, int groups,
                                              double *acc)
{
    __shared__ double sh[kBnThreads][2 * V];
    const int tid = threadIdx.x;
#pragma unroll
    for (int c = 0; c < V; ++c) { sh[tid][c] = a[c]; sh[tid][V + c] = b[c]; }
    __syncthreads();
    if (tid < HC) {
#pragma unroll
        for (int c = 0; c < V; ++c) {
            double s0 = 0.0, s1 = 0.0;
            for (int q = 0; q < groups; ++q) { s0 += sh[q * HC + tid][c]; s1 += sh[q * HC + tid][V + c]; }
            atomicAdd(acc + tid * V + c, s0);
            atomicAdd(acc + H + tid * V + c, s1);
        }
    }
}

// acc[0..H) += sum_k y, acc[H..2H) += sum_k y^2
template <int V>
__global__ void __launch_bounds__(kBnThreads)
k_bn_moments(const float *__restrict__ y, int64_t K, int H, double *__restrict__ acc)
{
    const int HC = H / V, groups = kBnThreads / HC, tid = threadIdx.x, r = tid / HC, j = (tid - r * HC) * V;
    double s[V], q[V];
#pragma unroll
    for (int c = 0; c < V; ++c) { s[c] = 0.0; q[c] = 0.0; }
    if (r < groups) {
        const int64_t stride = (int64_t)gridDim.x * groups;
        for (int64_t k0 = (int64_t)blockIdx.x * groups + r; k0 < K; k0 += kBnU * stride) {
            float v[kBnU][V];
#pragma unroll
            for (int u = 0; u < kBnU; ++u) bn_load<V>(y, k0 + u * stride, K, H, j, v[u]);
#pragma unroll
            for (int u = 0; u < kBnU; ++u)
#pragma unroll
                for (int c = 0; c < V; ++c) { s[c] += (double)v[u][c]; q[c] += (double)v[u][c] * (double)v[u][c]; }
        }
    }
    bn_block_add2<V>(s, q, H, HC, groups, acc);
}

// mean / invstd from the moments (every block, for its own use); block 0 also saves them and updates the running
// statistics like torch (biased variance for the normalisation, unbiased for running_var); then the transform + ReLU
template <int V>
__global__ void __launch_bounds__(kBnThreads)
k_bn_relu_fwd(const float *__restrict__ y, const double *__restrict__ acc, int64_t K, int H,
              const float *__restrict__ gamma, const float *__restrict__ beta, float eps, float momentum,
              float *running_mean, float *running_var, float *save_mean, float *save_invstd, float *__restrict__ z)
{
    const int HC = H / V, groups = kBnThreads / HC, tid = threadIdx.x, r = tid / HC, j = (tid - r * HC) * V;
    if (r >= groups) return;
    float mean[V], invstd[V], ga[V], be[V];
#pragma unroll
    for (int c = 0; c < V; ++c) {
        const double mean_d = acc[j + c] / (double)K;
        double var_d = acc[H + j + c] / (double)K - mean_d * mean_d;
        if (var_d < 0.0) var_d = 0.0;
        mean[c] = (float)mean_d;
        invstd[c] = (float)(1.0 / sqrt(var_d + (double)eps));
        ga[c] = __ldg(gamma + j + c);
        be[c] = __ldg(beta + j + c);
        if (blockIdx.x == 0 && r == 0) {
            save_mean[j + c] = mean[c];
            save_invstd[j + c] = invstd[c];
            if (running_mean) running_mean[j + c] = (1.0f - momentum) * running_mean[j + c] + momentum * mean[c];
            if (running_var) {
                const double unb = K > 1 ? var_d * (double)K / (double)(K - 1) : var_d;
                running_var[j + c] = (1.0f - momentum) * running_var[j + c] + momentum * (float)unb;
            }
        }
    }
    const int64_t stride = (int64_t)gridDim.x * groups;
    for (int64_t k0 = (int64_t)blockIdx.x * groups + r; k0 < K; k0 += kBnU * stride) {
        float v[kBnU][V];
#pragma unroll
        for (int u = 0; u < kBnU; ++u) bn_load<V>(y, k0 + u * stride, K, H, j, v[u]);
#pragma unroll
        for (int u = 0; u < kBnU; ++u) {
            const int64_t k = k0 + u * stride;
            if (k < K) {
                float o[V];
#pragma unroll
                for (int c = 0; c < V; ++c) {
                    const float t = __fadd_rn(__fmul_rn(__fmul_rn(__fsub_rn(v[u][c], mean[c]), invstd[c]), ga[c]), be[c]);
                    o[c] = t < 0.0f ? 0.0f : t;                      // NaN goes through, like ATen's relu
                }
                bn_store<V>(z, k, H, j, o);
            }
        }
    }
}

// acc[0..H) += sum_k g', acc[H..2H) += sum_k g' * xhat, with g' = g where the ReLU was open (recomputed from y)
template <int V>
__global__ void __launch_bounds__(kBnThreads)
k_bn_relu_bwd_reduce(const float *__restrict__ g, const float *__restrict__ y, const float *__restrict__ save_mean,
                     const float *__restrict__ save_invstd, const float *__restrict__ gamma,
                     const float *__restrict__ beta, int64_t K, int H, double *__restrict__ acc)
{
    const int HC = H / V, groups = kBnThreads / HC, tid = threadIdx.x, r = tid / HC, j = (tid - r * HC) * V;
    double s1[V], s2[V];
#pragma unroll
    for (int c = 0; c < V; ++c) { s1[c] = 0.0; s2[c] = 0.0; }
    if (r < groups) {
        float mean[V], invstd[V], ga[V], be[V];
#pragma unroll
        for (int c = 0; c < V; ++c) {
            mean[c] = __ldg(save_mean + j + c); invstd[c] = __ldg(save_invstd + j + c);
            ga[c] = __ldg(gamma + j + c); be[c] = __ldg(beta + j + c);
        }
        const int64_t stride = (int64_t)gridDim.x * groups;
        for (int64_t k0 = (int64_t)blockIdx.x * groups + r; k0 < K; k0 += kBnU * stride) {
            float v[kBnU][V], gv[kBnU][V];
#pragma unroll
            for (int u = 0; u < kBnU; ++u) {
                bn_load<V>(y, k0 + u * stride, K, H, j, v[u]);
                bn_load<V>(g, k0 + u * stride, K, H, j, gv[u]);
            }
#pragma unroll
            for (int u = 0; u < kBnU; ++u)
#pragma unroll
                for (int c = 0; c < V; ++c) {
                    const float xh = __fmul_rn(__fsub_rn(v[u][c], mean[c]), invstd[c]);
                    const float t = __fadd_rn(__fmul_rn(xh, ga[c]), be[c]);
                    const float gp = t <= 0.0f ? 0.0f : gv[u][c];    // threshold_backward's rule (NaN passes)
                    s1[c] += (double)gp;
                    s2[c] += (double)gp * (double)xh;
                }
        }
    }
    bn_block_add2<V>(s1, s2, H, HC, groups, acc);
}

// grad_y = gamma * invstd * (g' - mean_k g' - xhat * mean_k(g' xhat)); block 0 writes grad_gamma = sum g' xhat,
// grad_beta = sum g'
template <int V>
__global__ void __launch_bounds__(kBnThreads)
k_bn_relu_bwd_elemt(const float *__restrict__ g, const float *__restrict__ y, const float *__restrict__ save_mean,
                    const float *__restrict__ save_invstd, const float *__restrict__ gamma,
                    const float *__restrict__ beta, const double *__restrict__ acc, int64_t K, int H,
                    float *__restrict__ grad_y, float *grad_gamma, float *grad_beta)
{
    const int HC = H / V, groups = kBnThreads / HC, tid = threadIdx.x, r = tid / HC, j = (tid - r * HC) * V;
    if (r >= groups) return;
    float mean[V], invstd[V], ga[V], be[V], m1[V], m2[V], sc[V];
#pragma unroll
    for (int c = 0; c < V; ++c) {
        mean[c] = __ldg(save_mean + j + c); invstd[c] = __ldg(save_invstd + j + c);
        ga[c] = __ldg(gamma + j + c); be[c] = __ldg(beta + j + c);
        m1[c] = (float)(acc[j + c] / (double)K);
        m2[c] = (float)(acc[H + j + c] / (double)K);
        sc[c] = ga[c] * invstd[c];
        if (blockIdx.x == 0 && r == 0) {
            if (grad_beta) grad_beta[j + c] = (float)acc[j + c];
            if (grad_gamma) grad_gamma[j + c] = (float)acc[H + j + c];
        }
    }
    const int64_t stride = (int64_t)gridDim.x * groups;
    for (int64_t k0 = (int64_t)blockIdx.x * groups + r; k0 < K; k0 += kBnU * stride) {
        float v[kBnU][V], gv[kBnU][V];
#pragma unroll
        for (int u = 0; u < kBnU; ++u) {
            bn_load<V>(y, k0 + u * stride, K, H, j, v[u]);
            bn_load<V>(g, k0 + u * stride, K, H, j, gv[u]);
        }
#pragma unroll
        for (int u = 0; u < kBnU; ++u) {
            const int64_t k = k0 + u * stride;
            if (k < K) {
                float o[V];
#pragma unroll
                for (int c = 0; c < V; ++c) {
                    const float xh = __fmul_rn(__fsub_rn(v[u][c], mean[c]), invstd[c]);
                    const float t = __fadd_rn(__fmul_rn(xh, ga[c]), be[c]);
                    const float gp = t <= 0.0f ? 0.0f : gv[u][c];
                    o[c] = sc[c] * (gp - m1[c] - xh * m2[c]);
                }
                bn_store<V>(grad_y, k, H, j, o);
            }
        }
    }
}

}  // namespace rpc

// perturber.cuh -- the elementwise / reduction parts of VoxelPerturber.forward around the MLP:
//   k_col_moments       per-channel sum and sum of squares (+ sum of row norms, sum |channel 3|) of a [K, C] matrix
//   k_normalize_fwd     xn = clamp(x / (std(x, 0) + 1e-6), -10, 10)      (/root/reference/models/adversarial/
//                       voxel_perturber.py:157-168; NaN / Inf scale -> unit scaling, :161-163)
//   k_losses_finalize   l2_norm = mean_k ||p_k||, intensity = mean_k |p_k3|, bias = mean_c |mean_k p_kc|,
//                       imbalance = std_c(std_k p_kc)                     (:268-299) + the per-channel constants
//                       their gradient needs
//   k_losses_bwd        grad_p from the four upstream scalars (torch's rules: norm and std backward are 0 where the
//                       norm / std is 0)
// The reference runs ~25 small ATen kernels and 4 host syncs for this; here it is 5 launches and no sync.
// Moments are accumulated in fp64 (block tree, then one atomicAdd per block and value), which is at least as
// accurate as torch's fp32 Welford; results agree with the reference fixtures to 1e-6 relative.
#pragma once
#include "common.cuh"

namespace rpc {

constexpr int kMomVals = 2 * RPC_MAX_CHANNELS + 2;      // S_c, Q_c, sum of row norms, sum |x_k3|

// acc[0..C) += sum_k x_kc, acc[8..8+C) += sum_k x_kc^2, acc[16] += sum_k ||x_k||, acc[17] += sum_k |x_k3|
template <int C>
__global__ void __launch_bounds__(kThreads)
k_col_moments(const float *__restrict__ x, int64_t K, const int *__restrict__ k_dev, double *__restrict__ acc)
{
    if (k_dev) K = min(K, (int64_t)*k_dev);
    double s[C], q[C], nrm = 0.0, a3 = 0.0;
#pragma unroll
    for (int c = 0; c < C; ++c) { s[c] = 0.0; q[c] = 0.0; }
    for (int64_t k = (int64_t)blockIdx.x * kThreads + threadIdx.x; k < K; k += (int64_t)gridDim.x * kThreads) {
        float v[C];
        if constexpr (C == 4) {
            const float4 t = __ldg(reinterpret_cast<const float4 *>(x) + k);
            v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
        } else {
#pragma unroll
            for (int c = 0; c < C; ++c) v[c] = __ldg(x + k * C + c);
        }
        float n2 = 0.0f;
#pragma unroll
        for (int c = 0; c < C; ++c) {
            s[c] += (double)v[c];
            q[c] += (double)v[c] * (double)v[c];
            n2 = __fadd_rn(n2, __fmul_rn(v[c], v[c]));
        }
        nrm += (double)sqrtf(n2);
        if (C >= 4) a3 += (double)fabsf(v[3 % C]);
    }
    __shared__ double sh[kThreads / 32][kMomVals];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    auto wsum = [](double v) {
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) v += __shfl_down_sync(0xFFFFFFFFu, v, d);
        return v;
    };
#pragma unroll
    for (int c = 0; c < C; ++c) {
        const double a = wsum(s[c]), b = wsum(q[c]);
        if (lane == 0) { sh[warp][c] = a; sh[warp][RPC_MAX_CHANNELS + c] = b; }
    }
    {
        const double a = wsum(nrm), b = wsum(a3);
        if (lane == 0) { sh[warp][2 * RPC_MAX_CHANNELS] = a; sh[warp][2 * RPC_MAX_CHANNELS + 1] = b; }
    }
    __syncthreads();
    if (threadIdx.x < kMomVals) {
        const int j = threadIdx.x;
        const bool used = j >= 2 * RPC_MAX_CHANNELS || (j % RPC_MAX_CHANNELS) < C;
        if (used) {
            double t = 0.0;
            for (int w = 0; w < kThreads / 32; ++w) t += sh[w][j];
            atomicAdd(acc + j, t);
        }
    }
}

// unbiased std over rows of channel c from the moments, as fp32 (NaN when K < 2, like torch)
__device__ __forceinline__ float col_std(const double *acc, int c, int64_t K)
{
    const double S = acc[c], Q = acc[RPC_MAX_CHANNELS + c];
    double var = (Q - S * S / (double)K) / (double)(K - 1);
    if (var < 0.0) var = 0.0;                                       // rounding of an exactly constant column
    return (float)sqrt(var);
}

__global__ void __launch_bounds__(kThreads)
k_normalize_fwd(const float *__restrict__ x, const double *__restrict__ acc, int64_t K, const int *__restrict__ k_dev,
                int C, float *__restrict__ xn)
{
    if (k_dev) K = min(K, (int64_t)*k_dev);
    __shared__ float scale[RPC_MAX_CHANNELS];
    if (threadIdx.x == 0) {
        bool bad = false;
        for (int c = 0; c < C; ++c) {
            scale[c] = __fadd_rn(col_std(acc, c, K), 1e-6f);
            bad |= isnan(scale[c]) || isinf(scale[c]);
        }
        if (bad)                                                     // voxel_perturber.py:161-163
            for (int c = 0; c < C; ++c) scale[c] = 1.0f;
    }
    __syncthreads();
    const int64_t total = K * C;
    for (int64_t e = (int64_t)blockIdx.x * kThreads + threadIdx.x; e < total; e += (int64_t)gridDim.x * kThreads) {
        float v = __fdiv_rn(__ldg(x + e), scale[e % C]);
        v = v < -10.0f ? -10.0f : v;                                 // torch.clamp propagates NaN
        v = v > 10.0f ? 10.0f : v;
        xn[e] = v;
    }
}

// losses[4] = {l2_norm, intensity, bias, imbalance}; consts: [0..C) column means m_c, [8..8+C) weights w_c of
// (p_kc - m_c) in d imbalance / d p_kc, [16..16+C) d bias / d p_kc, [24] 1/K
__global__ void k_losses_finalize(const double *__restrict__ acc, int64_t K, const int *__restrict__ k_dev, int C,
                                  float *__restrict__ losses, float *__restrict__ consts)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    if (k_dev) K = min(K, (int64_t)*k_dev);
    const double invK = 1.0 / (double)K;
    double m[RPC_MAX_CHANNELS], sd[RPC_MAX_CHANNELS], bias = 0.0, smean = 0.0;
    for (int c = 0; c < C; ++c) {
        m[c] = acc[c] * invK;
        sd[c] = (double)col_std(acc, c, K);
        bias += fabs(m[c]);
        smean += sd[c];
    }
    smean /= C;
    double v = 0.0;
    for (int c = 0; c < C; ++c) v += (sd[c] - smean) * (sd[c] - smean);
    const double imb = sqrt(v / (double)(C - 1));
    losses[0] = (float)(acc[2 * RPC_MAX_CHANNELS] * invK);
    losses[1] = C >= 4 ? (float)(acc[2 * RPC_MAX_CHANNELS + 1] * invK) : 0.0f;
    losses[2] = (float)(bias / C);
    losses[3] = (float)imb;
    for (int c = 0; c < C; ++c) {
        consts[c] = (float)m[c];
        // torch: std_backward is 0 where the std is 0 (both the per-column and the across-column one)
        const double w = (imb > 0.0 && sd[c] > 0.0) ? (sd[c] - smean) / ((double)(C - 1) * imb) / ((double)(K - 1) * sd[c]) : 0.0;
        consts[RPC_MAX_CHANNELS + c] = (float)w;
        consts[2 * RPC_MAX_CHANNELS + c] = (float)((m[c] > 0.0 ? 1.0 : m[c] < 0.0 ? -1.0 : 0.0) * invK / C);
    }
    consts[3 * RPC_MAX_CHANNELS] = (float)invK;
}

// grad_p[k,c] = g0 * p_kc / (||p_k|| K) + [c == 3] g1 * sign(p_k3) / K + g2 * sign(m_c) / (C K) + g3 * w_c * (p_kc - m_c)
__global__ void __launch_bounds__(kThreads)
k_losses_bwd(const float *__restrict__ p, const float *__restrict__ consts, const float *__restrict__ g, int64_t K,
             const int *__restrict__ k_dev, int C, float *__restrict__ grad_p)
{
    if (k_dev) K = min(K, (int64_t)*k_dev);
    const float g0 = __ldg(g), g1 = __ldg(g + 1), g2 = __ldg(g + 2), g3 = __ldg(g + 3);
    const float invK = __ldg(consts + 3 * RPC_MAX_CHANNELS);
    for (int64_t k = (int64_t)blockIdx.x * kThreads + threadIdx.x; k < K; k += (int64_t)gridDim.x * kThreads) {
        float v[RPC_MAX_CHANNELS], n2 = 0.0f;
        for (int c = 0; c < C; ++c) { v[c] = __ldg(p + k * C + c); n2 = __fadd_rn(n2, __fmul_rn(v[c], v[c])); }
        const float nrm = sqrtf(n2);
        const float l2w = nrm > 0.0f ? g0 * invK / nrm : 0.0f;      // torch: norm backward is 0 at the zero vector
        for (int c = 0; c < C; ++c) {
            float d = l2w * v[c] + g2 * __ldg(consts + 2 * RPC_MAX_CHANNELS + c) +
                      g3 * __ldg(consts + RPC_MAX_CHANNELS + c) * (v[c] - __ldg(consts + c));
            if (c == 3) d += g1 * invK * (v[3] > 0.0f ? 1.0f : v[3] < 0.0f ? -1.0f : 0.0f);
            grad_p[k * C + c] = d;
        }
    }
}

// ---- post-VFE variant: StrongAdversarialVoxelNet.apply_enhanced_perturbations
// (/root/reference/models/detectors/strong_adversarial_voxelnet.py:156-180)
//   sp = (y - x) * scaling [+ alpha * last]   (y_is_delta: the adversary returned the perturbation itself, sp = y * scaling ...)
//    out = x + sp;  last' = sp;  l2 = ||sp||_2 over all elements
// One pass instead of five ATen kernels; the elementwise results are the same fp32 operations in the same order.
__global__ void __launch_bounds__(kThreads)
k_momentum_apply_fwd(const float *__restrict__ x, const float *__restrict__ y, const float *__restrict__ last,
                     int y_is_delta, float scaling, float alpha, int64_t n, float *__restrict__ out,
                     float *__restrict__ sp_out, double *__restrict__ sumsq)
{
    double acc = 0.0;
    for (int64_t e = (int64_t)blockIdx.x * kThreads + threadIdx.x; e < n; e += (int64_t)gridDim.x * kThreads) {
        const float xv = __ldg(x + e);
        const float yv = __ldg(y + e);
        float sp = __fmul_rn(y_is_delta ? yv : __fsub_rn(yv, xv), scaling);
        if (last) sp = __fadd_rn(sp, __fmul_rn(alpha, __ldg(last + e)));
        out[e] = __fadd_rn(xv, sp);
        sp_out[e] = sp;
        acc += (double)sp * (double)sp;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) acc += __shfl_down_sync(0xFFFFFFFFu, acc, d);
    __shared__ double sh[kThreads / 32];
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < kThreads / 32; ++w) t += sh[w];
        atomicAdd(sumsq, t);
    }
}

__global__ void k_momentum_l2(const double *__restrict__ sumsq, float *__restrict__ l2)
{
    if (threadIdx.x == 0 && blockIdx.x == 0) *l2 = (float)sqrt(*sumsq);
}

// g_sp = g_out + g_l2 * sp / l2 (second term 0 when l2 == 0, torch's norm backward);  grad_y = scaling * g_sp;
// grad_x = g_out - scaling * g_sp.  g_out or g_l2 may be NULL (no gradient arrived on that output).
__global__ void __launch_bounds__(kThreads)
k_momentum_apply_bwd(const float *__restrict__ g_out, const float *__restrict__ g_l2, const float *__restrict__ sp,
                     const float *__restrict__ l2, int y_is_delta, float scaling, int64_t n,
                     float *__restrict__ grad_x, float *__restrict__ grad_y)
{
    const float nrm = __ldg(l2);
    const float w = (g_l2 && nrm > 0.0f) ? __ldg(g_l2) / nrm : 0.0f;
    for (int64_t e = (int64_t)blockIdx.x * kThreads + threadIdx.x; e < n; e += (int64_t)gridDim.x * kThreads) {
        const float go = g_out ? __ldg(g_out + e) : 0.0f;
        const float gsp = go + w * __ldg(sp + e);
        if (grad_y) grad_y[e] = scaling * gsp;
        if (grad_x) grad_x[e] = y_is_delta ? go : go - scaling * gsp;
    }
}

}  // namespace rpc

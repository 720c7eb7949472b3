// vfe.cuh -- HardSimpleVFE mean forward / backward on the padded [M,P,C] tensor, and the backward
// of the fused north-star step (gradient of the voxel means back to the raw per-point perturbation).
#pragma once
#include "common.cuh"

namespace rpc {

// out[m,c] = sum_k voxels[m,k,c] / num_points[m]   (mmdet3d HardSimpleVFE.forward; all P slots are
// summed, in slot order).  One CTA stages VB voxel rows ([VB*P*C] contiguous floats) through shared
// memory with 16-byte loads, then one thread per (voxel, channel) reduces.
__global__ void __launch_bounds__(kThreads)
k_vfe_mean_fwd(const float *__restrict__ voxels, const int *__restrict__ num_points, int64_t M,
               const int *__restrict__ m_dev, int P, int C, int F, int VB, float *__restrict__ out)
{
    extern __shared__ __align__(16) float tile[];
    if (m_dev) M = min(M, (int64_t)*m_dev);
    for (int64_t v0 = (int64_t)blockIdx.x * VB; v0 < M; v0 += (int64_t)gridDim.x * VB) {
        const int nv = (int)min((int64_t)VB, M - v0);
        const int nf = nv * P * C;
        const float *src = voxels + v0 * P * C;
        const int nf4 = nf >> 2;
        for (int t = threadIdx.x; t < nf4; t += kThreads)
            reinterpret_cast<float4 *>(tile)[t] = __ldg(reinterpret_cast<const float4 *>(src) + t);
        for (int t = (nf4 << 2) + threadIdx.x; t < nf; t += kThreads) tile[t] = __ldg(src + t);
        __syncthreads();
        for (int t = threadIdx.x; t < nv * F; t += kThreads) {
            const int v = t / F, c = t % F;
            float s = 0.0f;
            for (int k = 0; k < P; ++k) s = __fadd_rn(s, tile[(v * P + k) * C + c]);
            out[(v0 + v) * F + c] = __fdiv_rn(s, (float)num_points[v0 + v]);
        }
        __syncthreads();
    }
}

// grad_voxels[m,k,c] = grad_out[m,c] / num_points[m] for every slot, 0 for c >= F
__global__ void __launch_bounds__(kThreads)
k_vfe_mean_bwd(const float *__restrict__ grad_out, const int *__restrict__ num_points, int64_t M,
               const int *__restrict__ m_dev, int P, int C, int F, float *__restrict__ grad_voxels)
{
    if (m_dev) M = min(M, (int64_t)*m_dev);
    const int64_t total = M * P * C;
    const int pc = P * C;
    for (int64_t t = (int64_t)blockIdx.x * kThreads + threadIdx.x; t < total;
         t += (int64_t)gridDim.x * kThreads) {
        const int64_t m = t / pc;
        const int c = (int)(t % C);
        grad_voxels[t] = c < F ? __fdiv_rn(__ldg(grad_out + m * F + c), (float)__ldg(num_points + m)) : 0.0f;
    }
}

// Backward of the fused step, per input point i (SURVEY a-3):
//   g = grad_feats[v,c] / num_points[v] if the point was stored in voxel v (c < F), else 0
//   grad_raw[i,c] = (-fb <= raw*eb <= fb) ? g * eb[c] : 0
// point2voxel comes from the forward.  With raw == NULL (no perturbation) the gradient w.r.t. the
// point features themselves is written (g).
template <int C>
__global__ void __launch_bounds__(kThreads)
k_step_bwd(const float *__restrict__ grad_feats, const int *__restrict__ num_points,
           const int *__restrict__ p2v, const float *__restrict__ raw,
           const __grid_constant__ PerturbDesc pd, int64_t ntotal, int F, float *__restrict__ grad_raw)
{
    for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < ntotal;
         i += (int64_t)gridDim.x * kThreads) {
        const int v = __ldg(p2v + i);
        float g[C];
#pragma unroll
        for (int c = 0; c < C; ++c) g[c] = 0.0f;
        if (v >= 0) {
            const float n = (float)__ldg(num_points + v);
            if (C == 4 && F == 4) {
                float4 t = __ldg(reinterpret_cast<const float4 *>(grad_feats) + v);
                g[0] = __fdiv_rn(t.x, n); g[1] = __fdiv_rn(t.y, n);
                g[2] = __fdiv_rn(t.z, n); g[3 % C] = __fdiv_rn(t.w, n);
            } else {
#pragma unroll
                for (int c = 0; c < C; ++c)
                    if (c < F) g[c] = __fdiv_rn(__ldg(grad_feats + (int64_t)v * F + c), n);
            }
        }
        if (raw) {
            if constexpr (C == 4) {
                float4 r = __ldg(reinterpret_cast<const float4 *>(raw) + i);
                const float rr[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    g[c] = perturb_pass(rr[c], pd.eb[c], pd.fb[c]) ? __fmul_rn(g[c], pd.eb[c]) : 0.0f;
            } else {
#pragma unroll
                for (int c = 0; c < C; ++c)
                    g[c] = perturb_pass(__ldg(raw + i * C + c), pd.eb[c], pd.fb[c]) ? __fmul_rn(g[c], pd.eb[c]) : 0.0f;
            }
        }
        if constexpr (C == 4) {
            reinterpret_cast<float4 *>(grad_raw)[i] = make_float4(g[0], g[1], g[2], g[3]);
        } else {
#pragma unroll
            for (int c = 0; c < C; ++c) grad_raw[i * C + c] = g[c];
        }
    }
}

}  // namespace rpc

// attention.cuh -- the spatial-attention gate of VoxelPerturber (first slice of SURVEY 8f row 2):
//   a_k   = sigmoid(w2 . relu(W1 xn_k + b1) + b2)          attention = Linear(C, C/2) -> ReLU -> Linear(C/2, 1) -> Sigmoid
//                                                          (/root/reference/models/adversarial/voxel_perturber.py:104-110)
//   out_k = t_k * a_k                                      raw_perturbations * attention_weights (:203-205)
// The reference runs this as two addmm, a ReLU, a sigmoid and a broadcast multiply (5 launches, 4 [K, <=C] round trips
// through HBM) and about twice that in the backward.  Here the forward is ONE streaming pass (xn and t in, out
// written) and the backward ONE pass that recomputes the gate instead of saving it: it writes grad_t (and grad_xn
// when asked for) and reduces the (C/2)(C + 2) + 1 parameter gradients in fp64 (registers -> shuffles -> shared
// memory -> one atomicAdd per block and value), so nothing of size K is kept between the two.
// HBM-bound: 12 C bytes per row forward, 16-20 C bytes backward; the weights (<= 41 floats) live in shared memory.
#pragma once
#include "common.cuh"

namespace rpc {

constexpr int kAttMaxMid = RPC_MAX_CHANNELS / 2;
constexpr int kAttMaxParams = kAttMaxMid * (RPC_MAX_CHANNELS + 2) + 1;      // W1, b1, w2, b2

template <int C> struct AttDims {
    static constexpr int mid = C / 2 > 0 ? C / 2 : 1;
    static constexpr int np = mid * (C + 2) + 1;
    // layout of the packed parameter (and gradient) vector
    static constexpr int w1 = 0, b1 = mid * C, w2 = mid * C + mid, b2 = mid * C + 2 * mid;
};

template <int C>
__device__ __forceinline__ void att_load_params(const float *w1, const float *b1, const float *w2, const float *b2,
                                                float *s)
{
    using D = AttDims<C>;
    for (int j = threadIdx.x; j < D::np; j += blockDim.x) {
        float v;
        if (j < D::b1) v = __ldg(w1 + j);
        else if (j < D::w2) v = __ldg(b1 + (j - D::b1));
        else if (j < D::b2) v = __ldg(w2 + (j - D::w2));
        else v = __ldg(b2);
        s[j] = v;
    }
    __syncthreads();
}

template <int C>
__device__ __forceinline__ void att_load_row(const float *__restrict__ m, int64_t k, float (&v)[C])
{
    if constexpr (C == 4) {
        const float4 t = __ldg(reinterpret_cast<const float4 *>(m) + k);
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    } else {
#pragma unroll
        for (int c = 0; c < C; ++c) v[c] = __ldg(m + k * C + c);
    }
}

template <int C>
__device__ __forceinline__ void att_store_row(float *__restrict__ m, int64_t k, const float (&v)[C])
{
    if constexpr (C == 4) {
        reinterpret_cast<float4 *>(m)[k] = make_float4(v[0], v[1], v[2], v[3]);
    } else {
#pragma unroll
        for (int c = 0; c < C; ++c) m[k * C + c] = v[c];
    }
}

// hidden activations (after the ReLU) and the gate of one row
template <int C>
__device__ __forceinline__ float att_gate(const float *s, const float (&xn)[C], float (&h)[AttDims<C>::mid])
{
    using D = AttDims<C>;
    float z = s[D::b2];
#pragma unroll
    for (int j = 0; j < D::mid; ++j) {
        float a = s[D::b1 + j];
#pragma unroll
        for (int c = 0; c < C; ++c) a += s[D::w1 + j * C + c] * xn[c];
        h[j] = a > 0.0f ? a : 0.0f;
        z += s[D::w2 + j] * h[j];
    }
    return 1.0f / (1.0f + expf(-z));                               // ATen's sigmoid formula
}

template <int C>
__global__ void __launch_bounds__(kThreads)
k_attention_gate_fwd(const float *__restrict__ xn, const float *__restrict__ t, const float *w1, const float *b1,
                     const float *w2, const float *b2, int64_t K, float *__restrict__ out, float *__restrict__ gate)
{
    __shared__ float s[kAttMaxParams];
    att_load_params<C>(w1, b1, w2, b2, s);
    for (int64_t k = (int64_t)blockIdx.x * kThreads + threadIdx.x; k < K; k += (int64_t)gridDim.x * kThreads) {
        float x[C], tv[C], h[AttDims<C>::mid];
        att_load_row<C>(xn, k, x);
        att_load_row<C>(t, k, tv);
        const float a = att_gate<C>(s, x, h);
#pragma unroll
        for (int c = 0; c < C; ++c) tv[c] *= a;
        att_store_row<C>(out, k, tv);
        if (gate) gate[k] = a;
    }
}

// acc[np] (fp64, zeroed by the caller) += parameter gradients; grad_t = g * a; grad_xn = W1^T (relu' . w2 gz)
template <int C>
__global__ void __launch_bounds__(kThreads)
k_attention_gate_bwd(const float *__restrict__ g, const float *__restrict__ xn, const float *__restrict__ t,
                     const float *w1, const float *b1, const float *w2, const float *b2, int64_t K,
                     float *__restrict__ grad_t, float *__restrict__ grad_xn, double *__restrict__ acc)
{
    using D = AttDims<C>;
    __shared__ float s[kAttMaxParams];
    __shared__ double sh[kThreads / 32][D::np];
    att_load_params<C>(w1, b1, w2, b2, s);
    float pg[D::np];
#pragma unroll
    for (int j = 0; j < D::np; ++j) pg[j] = 0.0f;
    for (int64_t k = (int64_t)blockIdx.x * kThreads + threadIdx.x; k < K; k += (int64_t)gridDim.x * kThreads) {
        float x[C], tv[C], gv[C], h[D::mid];
        att_load_row<C>(xn, k, x);
        att_load_row<C>(t, k, tv);
        att_load_row<C>(g, k, gv);
        const float a = att_gate<C>(s, x, h);
        float ga = 0.0f;
#pragma unroll
        for (int c = 0; c < C; ++c) { ga += gv[c] * tv[c]; gv[c] *= a; }
        att_store_row<C>(grad_t, k, gv);
        const float gz = ga * a * (1.0f - a);                       // sigmoid backward
        pg[D::b2] += gz;
        float gx[C];
#pragma unroll
        for (int c = 0; c < C; ++c) gx[c] = 0.0f;
#pragma unroll
        for (int j = 0; j < D::mid; ++j) {
            pg[D::w2 + j] += gz * h[j];
            const float gh = h[j] > 0.0f ? gz * s[D::w2 + j] : 0.0f;   // ReLU backward (0 at 0, like ATen's threshold)
            pg[D::b1 + j] += gh;
#pragma unroll
            for (int c = 0; c < C; ++c) {
                pg[D::w1 + j * C + c] += gh * x[c];
                gx[c] += gh * s[D::w1 + j * C + c];
            }
        }
        if (grad_xn) att_store_row<C>(grad_xn, k, gx);
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int j = 0; j < D::np; ++j) {
        double v = (double)pg[j];
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) v += __shfl_down_sync(0xFFFFFFFFu, v, d);
        if (lane == 0) sh[warp][j] = v;
    }
    __syncthreads();
    if (threadIdx.x < D::np) {
        double v = 0.0;
        for (int w = 0; w < kThreads / 32; ++w) v += sh[w][threadIdx.x];
        atomicAdd(acc + threadIdx.x, v);
    }
}

// acc -> the four fp32 gradient tensors of the attention parameters
__global__ void k_attention_gate_finalize(const double *__restrict__ acc, int C, float *gw1, float *gb1, float *gw2, float *gb2)
{
    const int mid = C / 2 > 0 ? C / 2 : 1;
    const int np = mid * (C + 2) + 1;
    for (int j = threadIdx.x; j < np; j += blockDim.x) {
        const float v = (float)acc[j];
        if (j < mid * C) gw1[j] = v;
        else if (j < mid * C + mid) gb1[j - mid * C] = v;
        else if (j < mid * C + 2 * mid) gw2[j - mid * C - mid] = v;
        else gb2[0] = v;
    }
}

}  // namespace rpc

// linear.cuh -- weight and bias gradient of the perturber's Linear layers (second slice of SURVEY 8f row 2):
//   grad_w[m, n] = sum_k g[k, m] * a[k, n],   grad_b[m] = sum_k g[k, m]         (g = dL/d(out) [K, M], a = input [K, N])
// for the MLP of /root/reference/models/adversarial/voxel_perturber.py:82-101 (M, N <= 128, K = 0.1-1 M rows).
// Measured on B200 (profiles/r01_mlp_in_loop_launches.txt): cuBLAS runs these K-long reductions as SIMT split-K
// GEMMs on 50-72 CTAs (185-200 us each at K = 131 k) plus a split-K epilogue and a 65-block column-sum for the bias
// -- about a quarter of the whole training step with the real perturber in the loop.  Here: every CTA of a
// 2-per-SM grid takes an equal chunk of rows, streams it through shared memory in tiles of kLinKT rows (both
// operands are K-major, so tiles are contiguous float4 copies and the inner product needs no transpose), keeps the
// whole M x N result in registers (16 x 16 threads, (TM/16) x (TN/16) each, explicit FMAs -- the library is built with
// --fmad=false) and adds it to the zeroed output with one fp32 `red` per element.  fp32 throughout, like cuBLAS'
// SIMT path; summation order differs (as between any two split-K factors), tests bound it at 1e-5 of the scale.
#pragma once
#include "common.cuh"

namespace rpc {

constexpr int kLinKT = 16;                      // rows per shared-memory tile
constexpr int kLinThreads = 256;

template <int TM, int TN>
__global__ void __launch_bounds__(kLinThreads, 2)
k_linear_wgrad(const float *__restrict__ g, const float *__restrict__ a, int64_t K, int M, int N, int64_t chunk,
               float *__restrict__ gw, float *__restrict__ gb)
{
    constexpr int RM = TM / 16, RN = TN / 16;
    __shared__ __align__(16) float sg[kLinKT][TM];
    __shared__ __align__(16) float sa[kLinKT][TN];
    const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
    const int64_t r0 = (int64_t)blockIdx.x * chunk, r1 = min(r0 + chunk, K);
    float acc[RM][RN], bs[RM];
#pragma unroll
    for (int i = 0; i < RM; ++i) {
        bs[i] = 0.0f;
#pragma unroll
        for (int j = 0; j < RN; ++j) acc[i][j] = 0.0f;
    }
    const bool vec_g = (M == TM) && ((reinterpret_cast<uintptr_t>(g) & 15u) == 0);
    const bool vec_a = (N == TN) && ((reinterpret_cast<uintptr_t>(a) & 15u) == 0);
    for (int64_t row = r0; row < r1; row += kLinKT) {
        const int rows = (int)min((int64_t)kLinKT, r1 - row);
        // ---- stage the two tiles (rows beyond the chunk and columns beyond M / N are zero)
        if (vec_g) {
            const float4 *src = reinterpret_cast<const float4 *>(g + row * M);
            for (int e = tid; e < kLinKT * TM / 4; e += kLinThreads)
                reinterpret_cast<float4 *>(&sg[0][0])[e] = e < rows * TM / 4 ? __ldg(src + e) : make_float4(0.f, 0.f, 0.f, 0.f);
        } else {
            for (int e = tid; e < kLinKT * TM; e += kLinThreads) {
                const int kk = e / TM, m = e % TM;
                sg[kk][m] = (kk < rows && m < M) ? __ldg(g + (row + kk) * M + m) : 0.0f;
            }
        }
        if (vec_a) {
            const float4 *src = reinterpret_cast<const float4 *>(a + row * N);
            for (int e = tid; e < kLinKT * TN / 4; e += kLinThreads)
                reinterpret_cast<float4 *>(&sa[0][0])[e] = e < rows * TN / 4 ? __ldg(src + e) : make_float4(0.f, 0.f, 0.f, 0.f);
        } else {
            for (int e = tid; e < kLinKT * TN; e += kLinThreads) {
                const int kk = e / TN, n = e % TN;
                sa[kk][n] = (kk < rows && n < N) ? __ldg(a + (row + kk) * N + n) : 0.0f;
            }
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < kLinKT; ++kk) {
            float gv[RM], av[RN];
#pragma unroll
            for (int i = 0; i < RM; ++i) gv[i] = sg[kk][ty * RM + i];
#pragma unroll
            for (int j = 0; j < RN; ++j) av[j] = sa[kk][tx * RN + j];
#pragma unroll
            for (int i = 0; i < RM; ++i) {
#pragma unroll
                for (int j = 0; j < RN; ++j) acc[i][j] = fmaf(gv[i], av[j], acc[i][j]);
                if (tx == 0) bs[i] += gv[i];
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < RM; ++i) {
        const int m = ty * RM + i;
        if (m >= M) continue;
#pragma unroll
        for (int j = 0; j < RN; ++j) {
            const int n = tx * RN + j;
            if (n < N) atomicAdd(gw + (int64_t)m * N + n, acc[i][j]);
        }
        if (tx == 0 && gb) atomicAdd(gb + m, bs[i]);
    }
}

}  // namespace rpc

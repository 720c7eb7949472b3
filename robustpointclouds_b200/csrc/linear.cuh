// linear.cuh -- weight and bias gradient of the perturber's Linear layers (second slice of SURVEY 8f row 2):
//   grad_w[m, n] = sum_k g[k, m] * a[k, n],   grad_b[m] = sum_k g[k, m]         (g = dL/d(out) [K, M], a = input [K, N])
// for the MLP of /root/reference/models/adversarial/voxel_perturber.py:82-101 (M, N <= 128, K = 0.1-1 M rows).
// Measured on B200 (profiles/r01_mlp_in_loop_launches.txt): cuBLAS runs these K-long reductions as SIMT split-K
// GEMMs on 50-72 CTAs (185-200 us each at K = 131 k) plus a split-K epilogue and a 65-block column-sum for the bias
// -- about a quarter of the whole training step with the real perturber in the loop.  Here: every CTA of a
// 2-per-SM grid takes an equal chunk of rows, streams it through shared memory in tiles of kLinKT rows (both
// operands are K-major, so tiles are contiguous float4 copies and the inner product needs no transpose), keeps the
// whole M x N result in registers (8 x 8 accumulators per thread, 64-256 threads per CTA, the next tile prefetched
// into registers while the current one is multiplied; explicit FMAs -- the library is built with --fmad=false) and adds it to the zeroed output with one fp32 `red` per element.  fp32 throughout, like cuBLAS'
// SIMT path; summation order differs (as between any two split-K factors), tests bound it at 1e-5 of the scale.
#pragma once
#include "common.cuh"

namespace rpc {

constexpr int kLinKT = 16;                      // rows per shared-memory tile

// optional prologue on the layer input: a = relu(batch_norm(a_pre)) with saved statistics, applied while the operand is
// staged (the trunk keeps only pre-activations between forward and backward; operation order of k_bn_relu_fwd)
struct BnIn { const float *mean, *invstd, *gamma, *beta; };
__device__ __forceinline__ float lin_bn_relu(float x, const float4 &b)
{
    const float t = __fadd_rn(__fmul_rn(__fmul_rn(__fsub_rn(x, b.x), b.y), b.z), b.w);
    return t < 0.0f ? 0.0f : t;
}

// thread tiling: 8 x 8 accumulators per thread where the tile is at least 64 wide (4 LDS.128 per 64 FMAs), 2 where it is 16
template <int TM, int TN> struct LinTile {
    static constexpr int RM = TM >= 64 ? 8 : 2, RN = TN >= 64 ? 8 : 2;
    static constexpr int TYD = TM / RM, TXD = TN / RN, NT = TYD * TXD;      // 64, 128 or 256 threads
    static constexpr int G4 = kLinKT * TM / 4, A4 = kLinKT * TN / 4;         // float4 per staged tile
    static constexpr int GV = (G4 + NT - 1) / NT, AV = (A4 + NT - 1) / NT;   // float4 per thread and tile
};

// column c (0 .. R-1) of a thread's micro-tile: R = 8 is two groups of four, TDIM * 4 apart, so that the float4
// reads of neighbouring threads are contiguous in shared memory (no bank conflicts)
template <int R, int TDIM>
__device__ __forceinline__ int lin_col(int t, int c)
{
    if constexpr (R == 8) return (c >> 2) * (TDIM * 4) + t * 4 + (c & 3);
    else return t * R + c;
}

template <int TM, int TN>
__global__ void __launch_bounds__(LinTile<TM, TN>::NT)
k_linear_wgrad(const float *__restrict__ g, const float *__restrict__ a, int64_t K, int M, int N, int64_t chunk,
               float *__restrict__ gw, float *__restrict__ gb, const BnIn bn)
{
    using T = LinTile<TM, TN>;
    constexpr int RM = T::RM, RN = T::RN, NT = T::NT;
    __shared__ __align__(16) float sg[kLinKT][TM];
    __shared__ __align__(16) float sa[kLinKT][TN];
    __shared__ float4 s_bn[TN];
    const int tid = threadIdx.x, ty = tid / T::TXD, tx = tid % T::TXD;
    const bool has_bn = bn.mean != nullptr;
    if (has_bn) {
        for (int n = tid; n < TN; n += NT)
            s_bn[n] = n < N ? make_float4(__ldg(bn.mean + n), __ldg(bn.invstd + n), __ldg(bn.gamma + n), __ldg(bn.beta + n))
                            : make_float4(0.f, 0.f, 0.f, 0.f);
        __syncthreads();
    }
    const int64_t r0 = (int64_t)blockIdx.x * chunk, r1 = min(r0 + chunk, K);
    float acc[RM][RN], bs[RM];
#pragma unroll
    for (int i = 0; i < RM; ++i) {
        bs[i] = 0.0f;
#pragma unroll
        for (int j = 0; j < RN; ++j) acc[i][j] = 0.0f;
    }
    const bool vec = (M == TM) && (N == TN) && ((reinterpret_cast<uintptr_t>(g) & 15u) == 0) &&
                     ((reinterpret_cast<uintptr_t>(a) & 15u) == 0);
    const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
    float4 pg[T::GV], pa[T::AV];                  // the next tile, on its way from global memory (vec path)
    auto fetch = [&](int64_t row) {
        const int rows = (int)min((int64_t)kLinKT, r1 - row);
        const float4 *src_g = reinterpret_cast<const float4 *>(g + row * M);
        const float4 *src_a = reinterpret_cast<const float4 *>(a + row * N);
#pragma unroll
        for (int q = 0; q < T::GV; ++q) {
            const int e = tid + q * NT;
            pg[q] = (e < rows * TM / 4) ? __ldg(src_g + e) : z4;
        }
#pragma unroll
        for (int q = 0; q < T::AV; ++q) {
            const int e = tid + q * NT;
            pa[q] = (e < rows * TN / 4) ? __ldg(src_a + e) : z4;
        }
    };
    if (vec && r0 < r1) fetch(r0);
    for (int64_t row = r0; row < r1; row += kLinKT) {
        // ---- stage the two tiles (rows beyond the chunk and columns beyond M / N are zero)
        if (vec) {
#pragma unroll
            for (int q = 0; q < T::GV; ++q) {
                const int e = tid + q * NT;
                if (e < T::G4) reinterpret_cast<float4 *>(&sg[0][0])[e] = pg[q];
            }
            const int rows_v = (int)min((int64_t)kLinKT, r1 - row);
#pragma unroll
            for (int q = 0; q < T::AV; ++q) {
                const int e = tid + q * NT;
                if (e < T::A4) {
                    float4 v = pa[q];
                    if (has_bn && e < rows_v * TN / 4) {                 // (rows beyond the chunk stay zero)
                        const int n = (e * 4) % TN;
                        v.x = lin_bn_relu(v.x, s_bn[n]); v.y = lin_bn_relu(v.y, s_bn[n + 1]);
                        v.z = lin_bn_relu(v.z, s_bn[n + 2]); v.w = lin_bn_relu(v.w, s_bn[n + 3]);
                    }
                    reinterpret_cast<float4 *>(&sa[0][0])[e] = v;
                }
            }
        } else {
            const int rows = (int)min((int64_t)kLinKT, r1 - row);
            for (int e = tid; e < kLinKT * TM; e += NT) {
                const int kk = e / TM, m = e % TM;
                sg[kk][m] = (kk < rows && m < M) ? __ldg(g + (row + kk) * M + m) : 0.0f;
            }
            for (int e = tid; e < kLinKT * TN; e += NT) {
                const int kk = e / TN, n = e % TN;
                float v = (kk < rows && n < N) ? __ldg(a + (row + kk) * N + n) : 0.0f;
                if (has_bn && kk < rows && n < N) v = lin_bn_relu(v, s_bn[n]);
                sa[kk][n] = v;
            }
        }
        __syncthreads();
        if (vec && row + kLinKT < r1) fetch(row + kLinKT);          // travels while this tile is multiplied
#pragma unroll
        for (int kk = 0; kk < kLinKT; ++kk) {
            float gv[RM], av[RN];
            if constexpr (RM == 8) {
                const float4 u = *reinterpret_cast<const float4 *>(&sg[kk][ty * 4]);
                const float4 v = *reinterpret_cast<const float4 *>(&sg[kk][T::TYD * 4 + ty * 4]);
                gv[0] = u.x; gv[1] = u.y; gv[2] = u.z; gv[3] = u.w; gv[4] = v.x; gv[5] = v.y; gv[6] = v.z; gv[7] = v.w;
            } else {
#pragma unroll
                for (int i = 0; i < RM; ++i) gv[i] = sg[kk][ty * RM + i];
            }
            if constexpr (RN == 8) {
                const float4 u = *reinterpret_cast<const float4 *>(&sa[kk][tx * 4]);
                const float4 v = *reinterpret_cast<const float4 *>(&sa[kk][T::TXD * 4 + tx * 4]);
                av[0] = u.x; av[1] = u.y; av[2] = u.z; av[3] = u.w; av[4] = v.x; av[5] = v.y; av[6] = v.z; av[7] = v.w;
            } else {
#pragma unroll
                for (int j = 0; j < RN; ++j) av[j] = sa[kk][tx * RN + j];
            }
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
        const int m = lin_col<RM, T::TYD>(ty, i);
        if (m >= M) continue;
#pragma unroll
        for (int j = 0; j < RN; ++j) {
            const int n = lin_col<RN, T::TXD>(tx, j);
            if (n < N) atomicAdd(gw + (int64_t)m * N + n, acc[i][j]);
        }
        if (tx == 0 && gb) atomicAdd(gb + m, bs[i]);
    }
}

// ---- skinny layers (first: C -> h0, last: h0 -> C): one operand is at most 8 wide, so the reduction is a stream over
// the wide operand [K, W] at HBM speed: thread = (row group, wide column), S accumulators each, the narrow row read
// as a broadcast; block tree over the row groups, then one fp32 `red` per CTA and element.
//   WIDE_IS_G: wide = g [K, M = W], narrow = a [K, N = S]  -> gw[j, s], gb[j] = column sums of the wide operand
//   else     : wide = a [K, N = W], narrow = g [K, M = S]  -> gw[s, j], gb[s] = column sums of the narrow operand
constexpr int kSkinnyThreads = 256;

template <int S, bool WIDE_IS_G>
__global__ void __launch_bounds__(kSkinnyThreads)
k_linear_wgrad_skinny(const float *__restrict__ wide, const float *__restrict__ narrow, int64_t K, int W,
                      float *__restrict__ gw, float *__restrict__ gb, const BnIn bn)
{
    const int groups = kSkinnyThreads / W;                          // rows taken per CTA iteration
    const int tid = threadIdx.x, r = tid / W, j = tid - r * W;
    const bool has_bn = !WIDE_IS_G && bn.mean != nullptr;           // the layer input is the wide operand
    float4 bj = make_float4(0.f, 0.f, 0.f, 0.f);
    if (has_bn && r < groups) bj = make_float4(__ldg(bn.mean + j), __ldg(bn.invstd + j), __ldg(bn.gamma + j), __ldg(bn.beta + j));
    float acc[S], colsum = 0.0f, nsum[S];
#pragma unroll
    for (int s = 0; s < S; ++s) { acc[s] = 0.0f; nsum[s] = 0.0f; }
    if (r < groups) {
        constexpr int U = 4;                                        // rows in flight per thread
        const int64_t stride = (int64_t)gridDim.x * groups;
        for (int64_t k0 = (int64_t)blockIdx.x * groups + r; k0 < K; k0 += U * stride) {
            float w[U], nv[U][S];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int64_t k = k0 + u * stride;
                const bool ok = k < K;
                w[u] = ok ? __ldg(wide + k * W + j) : 0.0f;
                if (has_bn && ok) w[u] = lin_bn_relu(w[u], bj);
                if constexpr (S == 4) {
                    const float4 t = ok ? __ldg(reinterpret_cast<const float4 *>(narrow) + k) : make_float4(0.f, 0.f, 0.f, 0.f);
                    nv[u][0] = t.x; nv[u][1] = t.y; nv[u][2] = t.z; nv[u][3] = t.w;
                } else {
#pragma unroll
                    for (int s = 0; s < S; ++s) nv[u][s] = ok ? __ldg(narrow + k * S + s) : 0.0f;
                }
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
#pragma unroll
                for (int s = 0; s < S; ++s) acc[s] = fmaf(w[u], nv[u][s], acc[s]);
                if constexpr (WIDE_IS_G) colsum += w[u];
                else if (j == 0) {
#pragma unroll
                    for (int s = 0; s < S; ++s) nsum[s] += nv[u][s];
                }
            }
        }
    }
    __shared__ float sh[kSkinnyThreads][S + 1];
#pragma unroll
    for (int s = 0; s < S; ++s) sh[tid][s] = acc[s];
    sh[tid][S] = colsum;
    __syncthreads();
    if (tid < W) {                                                  // row group 0 adds up its column over the groups
        float t[S + 1];
#pragma unroll
        for (int s = 0; s <= S; ++s) t[s] = 0.0f;
        for (int q = 0; q < groups; ++q)
#pragma unroll
            for (int s = 0; s <= S; ++s) t[s] += sh[q * W + tid][s];
#pragma unroll
        for (int s = 0; s < S; ++s) atomicAdd(WIDE_IS_G ? gw + (int64_t)tid * S + s : gw + (int64_t)s * W + tid, t[s]);
        if (WIDE_IS_G && gb) atomicAdd(gb + tid, t[S]);
    }
    if constexpr (!WIDE_IS_G) {
        __syncthreads();
        if (j == 0 && r < groups)
#pragma unroll
            for (int s = 0; s < S; ++s) sh[r][s] = nsum[s];
        __syncthreads();
        if (gb && tid < S) {
            float t = 0.0f;
            for (int q = 0; q < groups; ++q) t += sh[q][tid];
            atomicAdd(gb + tid, t);
        }
    }
}

}  // namespace rpc

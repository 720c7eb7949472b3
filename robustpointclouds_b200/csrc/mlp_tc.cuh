// mlp_tc.cuh -- the Linear layers of VoxelPerturber's MLP trunk on the 5th-generation tensor cores.
//
// The reference builds the trunk at /root/reference/models/adversarial/voxel_perturber.py:82-103 (six nn.Linear with
// BatchNorm1d + ReLU between them, Tanh at the end) and runs it at :176; autograd's addmm backward gives the input
// gradients.  Both products of a layer are the same shape of problem,
//     Y[K, NO] = X[K, NI] * B^T,   B[NO, NI],   K = 0.1 .. 1.2 M rows,  NI, NO <= 128,
// (forward: B = W; input gradient: X = dL/dY, B = W^T), i.e. a stream of 128-row tiles against a weight matrix that
// fits shared memory.  k_linear_tc runs it as fp32-accurate 3 x TF32 on tcgen05:
//     x = x_hi + x_lo, w = w_hi + w_lo (each part rounded to TF32; x_hi + x_lo carries 22 mantissa bits)
//     D = x_lo * w_hi + x_hi * w_lo + x_hi * w_hi        (x_lo * w_lo < 2^-22 relative: dropped)
// accumulated in fp32 in tensor memory: relative error ~1e-6, inside the 1e-5 parity bar, where single-pass TF32
// (1e-3) is not.
//
// One persistent CTA per SM, 256 threads:
//   * the weight matrix is split once into hi / lo and kept in shared memory in the tcgen05 K-major canonical layout
//     without swizzle (8 x 16-byte core matrices; SBO = 128 B between 8-row groups, LBO between 16-byte K chunks);
//   * per tile, all warps bring 128 rows x 64 columns of X through registers (optional fused prologue: the previous
//     layer's BatchNorm + ReLU, x' = relu((x - mean) * invstd * gamma + beta), so the normalised activation never exists in HBM),
//     split them and store hi / lo into one of two shared-memory stages, then one thread issues the 3 x (NI / 8)
//     tcgen05.mma (M = 128, N = NO, K = 8) into one of two TMEM accumulators and commits to mbarriers;
//   * the epilogue of tile t (tcgen05.ld -> + bias [-> tanh] -> global, and the per-column sum / sum of squares the
//     next BatchNorm needs) runs after the loads and the MMA issue of tile t + 1, so the tensor core, the loads and
//     the stores of neighbouring tiles overlap.
// Layers with NI <= 8 or NO <= 8 (the first and the last of the trunk) are streams, not contractions: k_linear_narrow_*.
#pragma once
#include "common.cuh"

namespace rpc {

constexpr int kTcThreads = 256;
constexpr int kTcRows = 128;                     // rows per tile = M of the MMA
constexpr int kTcStageCols = 64;                 // columns of X per shared-memory stage (32 when the weights take > 64 KB)
constexpr int kTcMaxDim = 128;

struct TcParams {
    const float *x;                              // [K, NI]
    const float *w;                              // layout 0: [NO, NI] (forward, nn.Linear weight); 1: [NI, NO] (input gradient)
    const float *bias;                           // [NO] or NULL
    float *y;                                    // [K, NO]
    const int *k_dev;                            // device row count or NULL
    long long K;
    int NI, NO, w_layout;
    int scols;                                   // columns per stage: 64 or 32
    const float *in_mean, *in_invstd, *in_gamma, *in_beta;   // [NI] each or NULL: prologue x' = relu(bn(x)), the
                                                 // previous layer's BatchNorm + ReLU in k_bn_relu_fwd's operation order
    int epi_tanh;                                // epilogue: y = tanh(y)
    double *moments;                             // [2 * NO] or NULL: += column sums and sums of squares of y
};

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint32_t to_tf32(float v)
{
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
    return r;
}
// v = hi + lo with both parts representable in TF32 (lo exact up to 2^-22 |v|)
__device__ __forceinline__ void split_tf32(float v, uint32_t &hi, uint32_t &lo)
{
    hi = to_tf32(v);
    lo = to_tf32(__fsub_rn(v, __uint_as_float(hi)));
}

// shared-memory matrix descriptor, K-major, no swizzle: start address, leading (K chunk) and stride (8-row group)
// byte offsets in 16-byte units, descriptor version 1 (sm_100)
__device__ __forceinline__ uint64_t tc_smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes)
{
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)(lbo_bytes >> 4) << 16) | ((uint64_t)(sbo_bytes >> 4) << 32) |
           (1ull << 46);
}
// instruction descriptor: D = fp32, A = B = TF32, both K-major, M = 128, N = n
__device__ __forceinline__ uint32_t tc_instr_desc(int n)
{
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(kTcRows >> 4) << 24);
}
__device__ __forceinline__ void tc_mma_tf32(uint32_t tmem_d, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate)
{
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tc_commit(uint32_t bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void tc_mbar_wait(uint32_t bar, uint32_t parity)
{
    uint32_t ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    } while (!ok);
}
// 16 consecutive fp32 columns of this thread's TMEM lane
__device__ __forceinline__ void tc_ld16(uint32_t taddr, float (&v)[16])
{
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// relu(batch_norm(x)) with the statistics given: the operation order of k_bn_relu_fwd (bnrelu.cuh), so that the mask
// the backward recomputes from y is the one the forward applied; NaN goes through like ATen's relu
__device__ __forceinline__ float bn_relu_value(float x, float mean, float invstd, float ga, float be)
{
    const float t = __fadd_rn(__fmul_rn(__fmul_rn(__fsub_rn(x, mean), invstd), ga), be);
    return t < 0.0f ? 0.0f : t;
}

// column totals of a [32 lanes] x [32 columns] block held one row per lane: after the butterfly lane j holds column j
__device__ __forceinline__ float warp_column_sums32(float (&v)[32])
{
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) {
        const bool up = lane & d;
#pragma unroll
        for (int i = 0; i < d; ++i) {
            const float send = up ? v[i] : v[i + d];
            const float keep = up ? v[i + d] : v[i];
            v[i] = keep + __shfl_xor_sync(0xFFFFFFFFu, send, d);
        }
    }
    return v[0];
}

// dynamic shared memory of k_linear_tc (128-byte aligned by the launch):
//   [0, 2 * wbytes)                  B hi, B lo: NOp rows x NI columns, chunk-major (see header comment)
//   [.., + 4 * stage bytes)          X stages: (stage 0 hi, stage 0 lo, stage 1 hi, stage 1 lo)
//   then mean / invstd / gamma / beta [4 * 128] floats, 4 mbarriers, the TMEM base address
__host__ __device__ inline int tc_stage_cols(int NI, int NOp) { return 2 * NOp * NI * 4 > 64 * 1024 ? 32 : kTcStageCols; }
__host__ __device__ inline size_t tc_smem_bytes(int NI, int NOp)
{
    return 2 * (size_t)NOp * NI * 4 + 4 * (size_t)kTcRows * tc_stage_cols(NI, NOp) * 4 + 4 * kTcMaxDim * 4 + 4 * 8 + 16;
}

__global__ void __launch_bounds__(kTcThreads, 1)
k_linear_tc(const __grid_constant__ TcParams p)
{
    extern __shared__ __align__(128) unsigned char tc_sm[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int NI = p.NI, NO = p.NO;
    const int NOp = (NO + 31) & ~31;                               // N of the MMA (16 | N at M = 128; 32 for the epilogue's column split)
    const uint32_t wbytes = (uint32_t)NOp * NI * 4;
    unsigned char *sB = tc_sm;
    unsigned char *sA = tc_sm + 2 * wbytes;
    const int scols = p.scols;
    const uint32_t stage_bytes = (uint32_t)kTcRows * scols * 4;      // one of hi / lo
    float4 *s_bn = reinterpret_cast<float4 *>(sA + 4 * stage_bytes);      // (mean, invstd, gamma, beta) per input column
    unsigned long long *s_bar = reinterpret_cast<unsigned long long *>(s_bn + kTcMaxDim);   // [0,1] stage free, [2,3] acc full
    uint32_t *s_tmem = reinterpret_cast<uint32_t *>(s_bar + 4);
    const uint32_t sB_u = smem_u32(sB), sA_u = smem_u32(sA), bar_u = smem_u32(s_bar);
    long long K = p.K;
    if (p.k_dev) K = min(K, (long long)*p.k_dev);
    const long long ntiles = (K + kTcRows - 1) / kTcRows;
    const int tmem_cols = NOp <= 32 ? 64 : NOp <= 64 ? 128 : 256;                      // two accumulators

    // ---- set-up: barriers, tensor memory, weights (hi / lo, canonical layout), prologue constants
    if (tid == 0) {
        for (int i = 0; i < 4; ++i) tc_mbar_init(bar_u + 8 * i, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        __syncwarp();
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)), "r"(tmem_cols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    {
        const int nch = NI >> 2;                                       // 16-byte chunks per row
        for (int e = tid; e < NOp * nch; e += kTcThreads) {
            int n, c;
            float v[4] = {0.f, 0.f, 0.f, 0.f};
            if (p.w_layout == 0) {                                     // B[n][k] = w[n * NI + k]: chunks of a row are contiguous
                n = e / nch; c = e - n * nch;
                if (n < NO) {
                    const float4 t = __ldg(reinterpret_cast<const float4 *>(p.w + (size_t)n * NI) + c);
                    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
                }
            } else {                                                   // B[n][k] = w[k * NO + n]: n fastest
                c = e / NOp; n = e - c * NOp;
                if (n < NO)
                    for (int q = 0; q < 4; ++q) v[q] = __ldg(p.w + (size_t)(4 * c + q) * NO + n);
            }
            uint32_t hi[4], lo[4];
            for (int q = 0; q < 4; ++q) split_tf32(v[q], hi[q], lo[q]);
            const uint32_t off = (uint32_t)c * (NOp * 16) + (uint32_t)(n >> 3) * 128 + (uint32_t)(n & 7) * 16;
            *reinterpret_cast<uint4 *>(sB + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<uint4 *>(sB + wbytes + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        }
        if (p.in_mean)
            for (int c = tid; c < NI; c += kTcThreads)
                s_bn[c] = make_float4(__ldg(p.in_mean + c), __ldg(p.in_invstd + c), __ldg(p.in_gamma + c), __ldg(p.in_beta + c));
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *s_tmem;
    const uint32_t idesc = tc_instr_desc(NOp);
    const int nstage = (NI + scols - 1) / scols;                    // stages per tile (1 .. 4)
    uint32_t uses0 = 0, uses1 = 0;                                    // fills of stage buffer 0 / 1 so far
    // per-thread column totals for the moments: lane j of warp w owns columns (w >> 2) * NOp / 2 + 32 * g + j
    double m_sum[2] = {0.0, 0.0}, m_sq[2] = {0.0, 0.0};

    // ---- epilogue of one tile: TMEM -> registers -> (+ bias, tanh) -> global; column moments
    auto epilogue = [&](long long tile, int it) {
        const int acc = it & 1;
        tc_mbar_wait(bar_u + 8 * (2 + acc), (uint32_t)(it >> 1) & 1u);
        tc_fence_after();
        const int q = warp & 3, h = warp >> 2;
        const long long row = tile * kTcRows + q * 32 + lane;
        const bool live = row < K;
        const int half = NOp >> 1;
        for (int g = 0; g * 32 < half; ++g) {
            const int c0 = h * half + g * 32;                       // first column of this group
            const int nv = min(32, half - g * 32);                  // 16 or 32 columns
            float v[32];
            {
                float t[16];
                tc_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * NOp + c0), t);
#pragma unroll
                for (int i = 0; i < 16; ++i) v[i] = t[i];
                if (nv > 16) tc_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * NOp + c0 + 16), t);
#pragma unroll
                for (int i = 0; i < 16; ++i) v[16 + i] = nv > 16 ? t[i] : 0.0f;
            }
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                const int c = c0 + i;
                float y = v[i];
                if (p.bias && c < NO) y = __fadd_rn(y, __ldg(p.bias + c));
                if (p.epi_tanh) y = tanhf(y);
                v[i] = (live && c < NO && i < nv) ? y : 0.0f;
            }
            if (live) {
                float *dst = p.y + row * NO + c0;
                if ((NO & 3) == 0) {
#pragma unroll
                    for (int i = 0; i < 32; i += 4)
                        if (i < nv && c0 + i < NO) *reinterpret_cast<float4 *>(dst + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
                } else {
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                        if (i < nv && c0 + i < NO) dst[i] = v[i];
                }
            }
            if (p.moments) {
                float sq[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) sq[i] = v[i] * v[i];
                const float s1 = warp_column_sums32(v), s2 = warp_column_sums32(sq);
                if (g == 0) { m_sum[0] += (double)s1; m_sq[0] += (double)s2; }
                else        { m_sum[1] += (double)s1; m_sq[1] += (double)s2; }
            }
        }
        tc_fence_before();
    };

    // ---- main loop: loads + MMA issue of tile `it`, then the epilogue of tile `it - 1`
    const int r8 = lane & 7, cq = lane >> 3;
    int it = 0;
    long long prev_tile = -1;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        const int acc = it & 1;
        for (int s = 0; s < nstage; ++s) {
            const int buf = (it * nstage + s) & 1;
            uint32_t &uses = buf ? uses1 : uses0;
            if (uses > 0) tc_mbar_wait(bar_u + 8 * buf, (uses - 1) & 1u);        // the MMAs that read this stage are done
            const int col0 = s * scols;
            const int ncol = min(scols, NI - col0);                               // multiple of 8
            const int ncg = (ncol + 15) >> 4;                                     // groups of 4 chunks (64 bytes of a row)
            unsigned char *a_hi = sA + (size_t)buf * 2 * stage_bytes, *a_lo = a_hi + stage_bytes;
            // a warp instruction covers 8 rows x 4 chunks: 64 contiguous bytes per row from global memory, and per
            // quarter-warp 128 contiguous bytes of shared memory (conflict-free 16-byte stores)
            const int nj = 16 * ncg;                                              // (row group, chunk group) pairs
            for (int j0 = warp; j0 < nj; j0 += 8 * 4) {
                float4 v[4];
                int rg[4], ch[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int j = j0 + 8 * u;
                    rg[u] = j / ncg;
                    ch[u] = (j - rg[u] * ncg) * 4 + cq;
                    const long long row = tile * kTcRows + rg[u] * 8 + r8;
                    v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (j < nj && row < K && ch[u] * 4 < ncol)
                        v[u] = __ldg(reinterpret_cast<const float4 *>(p.x + row * NI + col0) + ch[u]);
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    if (j0 + 8 * u >= nj || ch[u] * 4 >= ncol) continue;
                    float e[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
                    if (p.in_mean) {
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            const float4 b = s_bn[col0 + ch[u] * 4 + q];
                            e[q] = bn_relu_value(e[q], b.x, b.y, b.z, b.w);
                        }
                    }
                    uint32_t hi[4], lo[4];
#pragma unroll
                    for (int q = 0; q < 4; ++q) split_tf32(e[q], hi[q], lo[q]);
                    const uint32_t off = (uint32_t)ch[u] * 2048u + (uint32_t)rg[u] * 128u + (uint32_t)r8 * 16u;
                    *reinterpret_cast<uint4 *>(a_hi + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                    *reinterpret_cast<uint4 *>(a_lo + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            tc_fence_before();
            __syncthreads();
            if (tid == 0) {
                tc_fence_after();
                const uint32_t a_hi_u = sA_u + (uint32_t)buf * 2 * stage_bytes, a_lo_u = a_hi_u + stage_bytes;
                const uint32_t d = tmem_base + (uint32_t)(acc * NOp);
                for (int k = 0; k < (ncol >> 3); ++k) {                           // K = 8 per instruction: two 16-byte chunks
                    const uint32_t ao = (uint32_t)k * 2 * 2048u;
                    const uint32_t bo = (uint32_t)((col0 >> 2) + 2 * k) * (uint32_t)(NOp * 16);
                    const uint64_t ah = tc_smem_desc(a_hi_u + ao, 2048u, 128u), al = tc_smem_desc(a_lo_u + ao, 2048u, 128u);
                    const uint64_t bh = tc_smem_desc(sB_u + bo, (uint32_t)NOp * 16u, 128u);
                    const uint64_t bl = tc_smem_desc(sB_u + wbytes + bo, (uint32_t)NOp * 16u, 128u);
                    tc_mma_tf32(d, al, bh, idesc, (s | k) ? 1u : 0u);             // small terms first
                    tc_mma_tf32(d, ah, bl, idesc, 1u);
                    tc_mma_tf32(d, ah, bh, idesc, 1u);
                }
                tc_commit(bar_u + 8 * buf);                                       // stage free when these MMAs have read it
                if (s == nstage - 1) tc_commit(bar_u + 8 * (2 + acc));            // accumulator complete
            }
            ++uses;
        }
        if (prev_tile >= 0) epilogue(prev_tile, it - 1);
        prev_tile = tile;
    }
    if (prev_tile >= 0) epilogue(prev_tile, it - 1);

    // ---- moments out, tensor memory back
    if (p.moments) {
        const int h = warp >> 2, half = NOp >> 1;
#pragma unroll
        for (int g = 0; g < 2; ++g) {
            const int c = h * half + g * 32 + lane;
            if (g * 32 + lane < half && c < NO) {
                atomicAdd(p.moments + c, g ? m_sum[1] : m_sum[0]);
                atomicAdd(p.moments + NO + c, g ? m_sq[1] : m_sq[0]);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(tmem_cols) : "memory");
}

// ---------------------------------------------------------------------------------------------------------------
// Streams: the first layer (NI = C <= 8 inputs) and the last one (NO = C <= 8 outputs) of the trunk, and the input
// gradient of the last layer.  Same TcParams; fp32 FMA, weights in shared memory.
constexpr int kNarrowThreads = 256;

// NI <= 8: thread = (row, four output columns); a row's threads are neighbours, so the stores are contiguous
__global__ void __launch_bounds__(kNarrowThreads)
k_linear_narrow_in(const __grid_constant__ TcParams p)
{
    __shared__ __align__(16) float s_w[8 * kTcMaxDim];               // [c][m]
    __shared__ __align__(16) float s_b[kTcMaxDim];
    __shared__ double s_m[2 * kTcMaxDim];
    const int NI = p.NI, NO = p.NO, tid = threadIdx.x;
    for (int e = tid; e < NI * NO; e += kNarrowThreads) {
        const int c = e / NO, m = e - c * NO;
        s_w[c * NO + m] = p.w_layout == 0 ? __ldg(p.w + m * NI + c) : __ldg(p.w + c * NO + m);
    }
    for (int m = tid; m < NO; m += kNarrowThreads) s_b[m] = p.bias ? __ldg(p.bias + m) : 0.0f;
    for (int m = tid; m < 2 * NO; m += kNarrowThreads) s_m[m] = 0.0;
    __syncthreads();
    long long K = p.K;
    if (p.k_dev) K = min(K, (long long)*p.k_dev);
    const int gpr = NO >> 2;                                           // column groups per row (NO % 4 == 0)
    // the grid-wide stride is a multiple of gpr, so a thread keeps its four columns over all its rows
    const long long stride = (long long)gridDim.x * kNarrowThreads / gpr * gpr;
    const long long t0 = (long long)blockIdx.x * kNarrowThreads + tid;
    float ms[4] = {0.f, 0.f, 0.f, 0.f}, mq[4] = {0.f, 0.f, 0.f, 0.f};
    const int m0 = (int)(t0 % gpr) * 4;
    if (t0 < stride) {
        for (long long e = t0; e < K * gpr; e += stride) {
            const long long row = e / gpr;
            float acc[4] = {s_b[m0], s_b[m0 + 1], s_b[m0 + 2], s_b[m0 + 3]};
            for (int c = 0; c < NI; ++c) {
                float xv = __ldg(p.x + row * NI + c);
                if (p.in_mean) xv = bn_relu_value(xv, __ldg(p.in_mean + c), __ldg(p.in_invstd + c), __ldg(p.in_gamma + c), __ldg(p.in_beta + c));
                const float4 w4 = *reinterpret_cast<const float4 *>(s_w + c * NO + m0);
                acc[0] = fmaf(xv, w4.x, acc[0]); acc[1] = fmaf(xv, w4.y, acc[1]);
                acc[2] = fmaf(xv, w4.z, acc[2]); acc[3] = fmaf(xv, w4.w, acc[3]);
            }
            if (p.epi_tanh)
                for (int q = 0; q < 4; ++q) acc[q] = tanhf(acc[q]);
            *reinterpret_cast<float4 *>(p.y + row * NO + m0) = make_float4(acc[0], acc[1], acc[2], acc[3]);
            for (int q = 0; q < 4; ++q) { ms[q] += acc[q]; mq[q] = fmaf(acc[q], acc[q], mq[q]); }
        }
    }
    if (p.moments) {
        for (int q = 0; q < 4; ++q) { atomicAdd(s_m + m0 + q, (double)ms[q]); atomicAdd(s_m + NO + m0 + q, (double)mq[q]); }
        __syncthreads();
        for (int m = tid; m < 2 * NO; m += kNarrowThreads) atomicAdd(p.moments + m, s_m[m]);
    }
}

// NO <= 8: NI / 4 lanes per row (a power of two, 4 .. 32), one float4 of the row each; shuffle reduction
__global__ void __launch_bounds__(kNarrowThreads)
k_linear_narrow_out(const __grid_constant__ TcParams p)
{
    __shared__ __align__(16) float s_w[8 * kTcMaxDim];               // [m][k]
    __shared__ float4 s_bn[kTcMaxDim];
    const int NI = p.NI, NO = p.NO, tid = threadIdx.x;
    for (int e = tid; e < NI * NO; e += kNarrowThreads) {
        const int m = e / NI, k = e - m * NI;
        s_w[m * NI + k] = p.w_layout == 0 ? __ldg(p.w + m * NI + k) : __ldg(p.w + k * NO + m);
    }
    if (p.in_mean)
        for (int k = tid; k < NI; k += kNarrowThreads)
            s_bn[k] = make_float4(__ldg(p.in_mean + k), __ldg(p.in_invstd + k), __ldg(p.in_gamma + k), __ldg(p.in_beta + k));
    __syncthreads();
    long long K = p.K;
    if (p.k_dev) K = min(K, (long long)*p.k_dev);
    const int L = NI >> 2;                                             // lanes per row
    const int sub = tid % L;
    const long long rows_per_pass = (long long)gridDim.x * kNarrowThreads / L;
    const bool relu_in = p.in_mean != nullptr;
    for (long long row = ((long long)blockIdx.x * kNarrowThreads + tid) / L; row < (K + 31) / 32 * 32; row += rows_per_pass) {
        float4 xv = make_float4(0.f, 0.f, 0.f, 0.f);
        if (row < K) xv = __ldg(reinterpret_cast<const float4 *>(p.x + row * NI) + sub);
        float e[4] = {xv.x, xv.y, xv.z, xv.w};
        if (relu_in)
            for (int q = 0; q < 4; ++q) { const float4 b = s_bn[4 * sub + q]; e[q] = bn_relu_value(e[q], b.x, b.y, b.z, b.w); }
        float acc[8];
#pragma unroll
        for (int m = 0; m < 8; ++m) {
            acc[m] = 0.0f;
            if (m < NO) {
                const float4 w4 = *reinterpret_cast<const float4 *>(s_w + m * NI + 4 * sub);
                acc[m] = fmaf(e[0], w4.x, fmaf(e[1], w4.y, fmaf(e[2], w4.z, e[3] * w4.w)));
            }
        }
        for (int d = L >> 1; d >= 1; d >>= 1)
#pragma unroll
            for (int m = 0; m < 8; ++m)
                if (m < NO) acc[m] += __shfl_xor_sync(0xFFFFFFFFu, acc[m], d);
        if (row < K) {                                                 // every lane of the row holds all NO totals
#pragma unroll
            for (int m = 0; m < 8; ++m) {
                if (m < NO && (m % L) == sub) {
                    float v = acc[m];
                    if (p.bias) v = __fadd_rn(v, __ldg(p.bias + m));
                    if (p.epi_tanh) v = tanhf(v);
                    p.y[row * NO + m] = v;
                }
            }
        }
    }
}


// mean / invstd of a BatchNorm1d from the column moments the producing layer's epilogue accumulated; running statistics
// like torch (biased variance normalises, the unbiased one is tracked) -- what block 0 of k_bn_relu_fwd does
__global__ void k_bn_finalize(const double *__restrict__ acc, long long K, const int *__restrict__ k_dev, int H, float eps,
                              float momentum, float *running_mean, float *running_var, float *save_mean, float *save_invstd)
{
    if (k_dev) K = min(K, (long long)*k_dev);
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= H || K < 1) return;
    const double mean_d = acc[c] / (double)K;
    double var_d = acc[H + c] / (double)K - mean_d * mean_d;
    if (var_d < 0.0) var_d = 0.0;
    const float mean = (float)mean_d;
    save_mean[c] = mean;
    save_invstd[c] = (float)(1.0 / sqrt(var_d + (double)eps));
    if (running_mean) running_mean[c] = (1.0f - momentum) * running_mean[c] + momentum * mean;
    if (running_var) {
        const double unb = K > 1 ? var_d * (double)K / (double)(K - 1) : var_d;
        running_var[c] = (1.0f - momentum) * running_var[c] + momentum * (float)unb;
    }
}

// z = relu(bn(y)) with saved statistics (the backward's recomputation of a layer input for the weight gradient)
__global__ void __launch_bounds__(kNarrowThreads)
k_bn_relu_apply(const float *__restrict__ y, long long K, const int *__restrict__ k_dev, int H, const float *__restrict__ mean,
                const float *__restrict__ invstd, const float *__restrict__ gamma, const float *__restrict__ beta,
                float *__restrict__ z)
{
    if (k_dev) K = min(K, (long long)*k_dev);
    const long long n4 = K * H / 4;                                    // H % 4 == 0
    for (long long e = (long long)blockIdx.x * kNarrowThreads + threadIdx.x; e < n4; e += (long long)gridDim.x * kNarrowThreads) {
        const int c = (int)((e * 4) % H);
        const float4 v = __ldg(reinterpret_cast<const float4 *>(y) + e);
        float4 o;
        o.x = bn_relu_value(v.x, __ldg(mean + c), __ldg(invstd + c), __ldg(gamma + c), __ldg(beta + c));
        o.y = bn_relu_value(v.y, __ldg(mean + c + 1), __ldg(invstd + c + 1), __ldg(gamma + c + 1), __ldg(beta + c + 1));
        o.z = bn_relu_value(v.z, __ldg(mean + c + 2), __ldg(invstd + c + 2), __ldg(gamma + c + 2), __ldg(beta + c + 2));
        o.w = bn_relu_value(v.w, __ldg(mean + c + 3), __ldg(invstd + c + 3), __ldg(gamma + c + 3), __ldg(beta + c + 3));
        reinterpret_cast<float4 *>(z)[e] = o;
    }
}

}  // namespace rpc

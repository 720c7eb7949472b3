// wgrad_tc.cuh -- weight (and bias) gradient of the trunk's 64 <-> 128 Linear layers on the tensor cores:
//   grad_w[m, n] = sum_k g[k, m] * a[k, n]          g = dL/d(out) [K, M],  a = layer input [K, N]   (row-major)
// (autograd's addmm backward for the nn.Linear layers of /root/reference/models/adversarial/voxel_perturber.py:82-101).
// A reduction over K = 0.1 .. 1.2 M rows into a 128 x 64 (or 64 x 128) result.  The SIMT kernel of linear.cuh runs it at the
// fp32 FMA rate (2 TB/s of operands); here it is 3 x TF32 on tcgen05 with the whole result in ONE tensor-memory accumulator
// for the life of the CTA, so the kernel is a pure operand stream:
//   * both operands are "MN-major" for this product (the reduction index k is the row, the M / N index is contiguous), and
//     the canonical MN-major TF32 layout -- 32-column blocks, rows 128 bytes apart with their 32-byte chunks swizzled by
//     the row number mod 4, blocks LBO apart -- is the [block][row][32 columns] staging k_linear_tc uses with another
//     chunk permutation: a 32-row slice of [g | a] (6 blocks) is one stage, no transpose anywhere, and a K = 8 step
//     advances both descriptors by 8 rows = 1024 bytes;
//   * 8 loader warps: cp.async raw ring (3 stages) -> optional BatchNorm + ReLU on the `a` columns (the layer input is
//     recomputed from the previous pre-activation, as in rpc_linear_wgrad_bn) -> hi / lo -> one of two operand stages;
//     the bias gradient (column sums of g) is accumulated by the loaders on the way;
//   * 1 MMA thread: 4 K steps x 3 products per stage, always accumulating;
//   * at the end 4 warps read the accumulator (tcgen05.ld) and add it to the zeroed output (one fp32 red per element).
// The side of the product that is 128 wide is the MMA's M (its rows are TMEM lanes); for out = 64, in = 128 the kernel
// computes grad_w^T and the epilogue writes it transposed.
#pragma once
#include "mlp_tc.cuh"

namespace rpc {

constexpr int kWgRows = 32;                          // rows (K of the product) per stage
constexpr int kWgBlocks = 6;                         // 32-column blocks per stage: (128 + 64) / 32
constexpr int kWgBlockB = kWgRows * 128;             // one block of a stage: 4 KB
constexpr int kWgStageB = kWgBlocks * kWgBlockB;     // one of hi / lo / raw: 24 KB
constexpr int kWgRaw = 3;
constexpr int kWgLoadWarps = 8;
constexpr int kWgThreads = (kWgLoadWarps + 1) * 32;

struct WgParams {
    const float *g, *a;                              // [K, M], [K, N]
    long long K;
    int M, N;                                        // out, in: {128, 64} or {64, 128}
    const float *bn_mean, *bn_invstd, *bn_gamma, *bn_beta;   // [N] each or NULL: a = relu(bn(a_pre))
    float *gw, *gb;                                  // [M, N] zeroed, [M] zeroed or NULL
};

__host__ __device__ inline size_t wg_smem_bytes()
{
    return (size_t)(4 + kWgRaw) * kWgStageB + kTcMaxDim * 16 + 6 * 8 + 16 + 1024 + 4 * kTcMaxDim;
}

// MN-major TF32 operand: the only layout the hardware takes is "128-byte swizzle with a 32-byte base" (layout type 1):
// rows of 32 columns (128 bytes), the 32-byte chunk c of row r stored at chunk position c ^ (r & 3), 4-row groups 512 bytes
// apart (SBO), 32-column blocks LBO apart.  (With the ordinary 128-byte swizzle, type 2, the MMA silently produced zeros.)
__device__ __forceinline__ uint64_t wg_smem_desc(uint32_t saddr)
{
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)(kWgBlockB >> 4) << 16) | ((uint64_t)(512u >> 4) << 32) | (1ull << 46) |
           (1ull << 61);
}
// D = fp32, A = B = TF32, both MN-major, M = 128, N = n
__device__ __forceinline__ uint32_t wg_instr_desc(int n)
{
    return (1u << 4) | (2u << 7) | (2u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
}

__global__ void __launch_bounds__(kWgThreads, 1)
k_wgrad_tc(const __grid_constant__ WgParams p)
{
    extern __shared__ __align__(1024) unsigned char wg_sm_raw[];
    unsigned char *sm = wg_sm_raw + ((1024u - (smem_u32(wg_sm_raw) & 1023u)) & 1023u);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned char *sA = sm;                                          // 2 stages x (hi, lo)
    unsigned char *sRaw = sm + 4 * kWgStageB;
    float4 *s_bn = reinterpret_cast<float4 *>(sRaw + kWgRaw * kWgStageB);
    float *s_gb = reinterpret_cast<float *>(s_bn + kTcMaxDim);       // bias-gradient partial sums of the CTA
    unsigned long long *s_bar = reinterpret_cast<unsigned long long *>(s_gb + kTcMaxDim);   // [0,1] full, [2,3] free, [4] done
    uint32_t *s_tmem = reinterpret_cast<uint32_t *>(s_bar + 6);
    const uint32_t sA_u = smem_u32(sA), sRaw_u = smem_u32(sRaw), bar_u = smem_u32(s_bar);
    const uint32_t bar_full = bar_u, bar_free = bar_u + 16, bar_done = bar_u + 32;
    const int M = p.M, N = p.N;
    const int gblk = M >> 5;                                          // blocks of g, then the blocks of a, in a stage
    const bool g_is_a_operand = M == 128;                             // the 128-wide side is the MMA's M
    const int tmem_cols = 128;
    // equal chunks of whole stages per CTA
    const long long nst_all = (p.K + kWgRows - 1) / kWgRows;
    const long long per = (nst_all + gridDim.x - 1) / gridDim.x;
    const long long st0 = (long long)blockIdx.x * per, st1 = min(st0 + per, nst_all);
    const long long njobs = st1 > st0 ? st1 - st0 : 0;

    if (tid == 0) {
        tc_mbar_init(bar_full, kWgLoadWarps * 32); tc_mbar_init(bar_full + 8, kWgLoadWarps * 32);
        tc_mbar_init(bar_free, 1); tc_mbar_init(bar_free + 8, 1);
        tc_mbar_init(bar_done, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == kWgLoadWarps) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)), "r"(tmem_cols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (p.bn_mean)
        for (int c = tid; c < N; c += kWgThreads)
            s_bn[c] = make_float4(__ldg(p.bn_mean + c), __ldg(p.bn_invstd + c), __ldg(p.bn_gamma + c), __ldg(p.bn_beta + c));
    for (int c = tid; c < kTcMaxDim; c += kWgThreads) s_gb[c] = 0.0f;
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *s_tmem;

    if (warp < kWgLoadWarps) {
        // ================= loaders.  A stage is 32 rows x 6 blocks of 32 columns ([g | a]); a warp instruction covers 8 rows
        // x 4 chunks (64 contiguous bytes per row); unit u of a thread is block u, the same 4 columns every time.
        const int r8 = lane & 7, cq = lane >> 3;
        const int rg = warp & 3, ch = (warp >> 2) * 4 + cq;             // row group and chunk inside the block, fixed per thread
        const int r = rg * 8 + r8;
        const uint32_t off0 = (uint32_t)r * 128u + (uint32_t)((((ch >> 1) ^ (r & 3)) * 32) + (ch & 1) * 16);
        float gbs[4][4];                                                // bias gradient: this thread's columns of the g blocks
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
            for (int q = 0; q < 4; ++q) gbs[u][q] = 0.0f;
        auto issue = [&](bool live, int rb, long long stage) {
            if (live) {
                const long long row = stage * kWgRows + r;
                const bool ok = row < p.K;
#pragma unroll
                for (int u = 0; u < kWgBlocks; ++u) {
                    const float *src = u < gblk ? p.g + (ok ? row : 0) * M + u * 32 + ch * 4
                                                : p.a + (ok ? row : 0) * N + (u - gblk) * 32 + ch * 4;
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;"
                                 ::"r"(sRaw_u + (uint32_t)rb * kWgStageB + (uint32_t)u * kWgBlockB + off0), "l"(src), "r"(ok ? 16 : 0) : "memory");
                }
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
        };
        for (int d = 0; d < kWgRaw; ++d) issue(d < njobs, d, st0 + d);
        int rb = 0;
        for (long long j = 0; j < njobs; ++j) {
            const int buf = (int)(j & 1);
            const bool live_row = (st0 + j) * kWgRows + r < p.K;
            asm volatile("cp.async.wait_group %0;" ::"n"(kWgRaw - 1) : "memory");
            float4 v[kWgBlocks];
#pragma unroll
            for (int u = 0; u < kWgBlocks; ++u)
                v[u] = *reinterpret_cast<const float4 *>(sRaw + (size_t)rb * kWgStageB + (size_t)u * kWgBlockB + off0);
            issue(j + kWgRaw < njobs, rb, st0 + j + kWgRaw);
            if (j >= 2) tc_mbar_wait(bar_free + 8 * buf, (uint32_t)((j >> 1) - 1) & 1u);
            unsigned char *a_hi = sA + (size_t)buf * 2 * kWgStageB, *a_lo = a_hi + kWgStageB;
#pragma unroll
            for (int u = 0; u < kWgBlocks; ++u) {
                float e[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
                if (u < gblk) {
                    if (u < 4) {
#pragma unroll
                        for (int q = 0; q < 4; ++q) gbs[u < 4 ? u : 0][q] += e[q];
                    }
                } else if (p.bn_mean) {
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const float4 b = s_bn[(u - gblk) * 32 + ch * 4 + q];
                        e[q] = live_row ? bn_relu_value(e[q], b.x, b.y, b.z, b.w) : 0.0f;   // rows beyond K stay zero
                    }
                }
                uint32_t hi[4], lo[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) split_tf32(e[q], hi[q], lo[q]);
                *reinterpret_cast<uint4 *>(a_hi + (size_t)u * kWgBlockB + off0) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                *reinterpret_cast<uint4 *>(a_lo + (size_t)u * kWgBlockB + off0) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            tc_mbar_arrive(bar_full + 8 * buf);
            if (++rb == kWgRaw) rb = 0;
        }
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        if (p.gb) {
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (u < gblk)
#pragma unroll
                    for (int q = 0; q < 4; ++q) atomicAdd(s_gb + u * 32 + ch * 4 + q, gbs[u][q]);
        }
    } else if (lane == 0) {
        // ================= MMA issuer
        const uint32_t idesc = wg_instr_desc(64);
        const uint32_t a_blk = g_is_a_operand ? 0u : (uint32_t)gblk, b_blk = g_is_a_operand ? (uint32_t)gblk : 0u;
        for (long long j = 0; j < njobs; ++j) {
            const int buf = (int)(j & 1);
            tc_mbar_wait(bar_full + 8 * buf, (uint32_t)(j >> 1) & 1u);
            tc_fence_after();
            const uint32_t hi_u = sA_u + (uint32_t)buf * 2 * kWgStageB, lo_u = hi_u + kWgStageB;
#pragma unroll
            for (int k = 0; k < kWgRows / 8; ++k) {
                const uint32_t ao = a_blk * kWgBlockB + (uint32_t)k * 1024u, bo = b_blk * kWgBlockB + (uint32_t)k * 1024u;
                const uint64_t ah = wg_smem_desc(hi_u + ao), al = wg_smem_desc(lo_u + ao);
                const uint64_t bh = wg_smem_desc(hi_u + bo), bl = wg_smem_desc(lo_u + bo);
                tc_mma_tf32(tmem_base, al, bh, idesc, (j | k) ? 1u : 0u);
                tc_mma_tf32(tmem_base, ah, bl, idesc, 1u);
                tc_mma_tf32(tmem_base, ah, bh, idesc, 1u);
            }
            tc_commit(bar_free + 8 * buf);
        }
        if (njobs > 0) tc_commit(bar_done);
    }
    __syncthreads();                                                   // bias partial sums are in s_gb; the MMA thread has issued everything

    // ================= epilogue: the accumulator (128 lanes x 64 columns) -> += grad_w
    if (warp < 4 && njobs > 0) {
        tc_mbar_wait(bar_done, 0u);
        tc_fence_after();
        const int mi = warp * 32 + lane;                               // index along the 128-wide side
#pragma unroll
        for (int c0 = 0; c0 < 64; c0 += 32) {
            uint32_t rr[32];
            tc_ld32_issue(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, rr);
            tc_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                const int c = c0 + i;                                  // index along the 64-wide side
                float *dst = g_is_a_operand ? p.gw + (size_t)mi * N + c : p.gw + (size_t)c * N + mi;
                atomicAdd(dst, __uint_as_float(rr[i]));
            }
        }
        tc_fence_before();
    }
    if (p.gb && njobs > 0)
        for (int m = tid; m < M; m += kWgThreads) atomicAdd(p.gb + m, s_gb[m]);
    __syncthreads();
    if (warp == kWgLoadWarps)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(tmem_cols) : "memory");
}

}  // namespace rpc

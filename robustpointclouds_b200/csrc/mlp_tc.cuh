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
// One persistent CTA per SM, 17 warps with fixed roles that meet only at mbarriers:
//   * the weight matrix is split once into hi / lo and kept in shared memory in the tcgen05 K-major canonical layout
//     with 128-byte swizzle (rows of 32 TF32, 8-row groups 1024 bytes apart);
//   * 8 loader warps bring 128 rows x 32 columns of X per stage through registers (optional fused prologue: the
//     previous layer's BatchNorm + ReLU, x' = relu((x - mean) * invstd * gamma + beta), so the normalised activation never
//     exists in HBM), split them and store hi / lo into one of two shared-memory stages; four stage loads stay in flight
//     per thread (64 KB per SM);
//   * one thread issues the 3 x (columns / 8) tcgen05.mma (M = 128, N = NO, K = 8) of a stage into one of two TMEM
//     accumulators and commits to the "stage free" / "accumulator full" mbarriers;
//   * 8 epilogue warps (two per TMEM lane quadrant, alternate 32-column groups) take the finished accumulator:
//     tcgen05.ld -> + bias [-> tanh] -> a padded per-warp slab in shared memory -> full 128-byte lines to global memory
//     (st.global; one 128-byte cp.async.bulk per row was measured and is slower: 231 vs 211 us at 64 -> 128),
//     and, lane = column, the per-column sum / sum of squares the next BatchNorm needs (no shuffles).
// The first version (all warps doing load -> MMA -> epilogue in turn) ran at 1.2 TB/s: profiles/r02_linear_tc.txt.
// The first layer (NI = C <= 8) and the input gradient of the last one are streams, not contractions: k_linear_narrow_in.
// The last layer (NO = C) runs on the tensor cores with N padded to 32 (a 16-lanes-per-row shuffle kernel was 2.3 x slower).
#pragma once
#include "common.cuh"

namespace rpc {

constexpr int kTcRows = 128;                     // rows per tile = M of the MMA
constexpr int kTcMaxDim = 128;

struct TcParams {
    const float *x;                              // [K, NI]
    const float *w;                              // layout 0: [NO, NI] (forward, nn.Linear weight); 1: [NI, NO] (input gradient)
    const float *bias;                           // [NO] or NULL
    float *y;                                    // [K, NO]
    const int *k_dev;                            // device row count or NULL
    long long K;
    int NI, NO, w_layout;
    const float *in_mean, *in_invstd, *in_gamma, *in_beta;   // [NI] each or NULL: prologue x' = relu(bn(x)), the
                                                 // previous layer's BatchNorm + ReLU in k_bn_relu_fwd's operation order
    int epi_tanh;                                // epilogue: y = tanh(y)
    int dbg;                                     // experiments (RPC_TC_DBG): 1 = one MMA per K step instead of three, 2 = none,
                                                 // 4 = role cycle counters (tools/dbg_tc_stats.py), 5 = no global stores
    double *moments;                             // [2 * NO] or NULL: += column sums and sums of squares of y
    // input-gradient mode: y is dL/d(relu(bn(v))) of the layer below and `moments` receives that BatchNorm's backward
    // reductions instead -- sum g' and sum g' * vhat with g' = y where the ReLU was open (what k_bn_relu_bwd_reduce computes)
    const float *red_v, *red_mean, *red_invstd, *red_gamma, *red_beta;   // v [K, NO] and its statistics, or NULL
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

// shared-memory matrix descriptor, K-major, 128-byte swizzle: rows of 32 TF32 (128 bytes), 8-row groups 1024 bytes apart
// (SBO), the 16-byte chunk c of row r stored at chunk position c ^ (r & 7); the start address may be advanced by 32
// bytes per K = 8 step inside the row (the hardware applies the XOR to the address bits); descriptor version 1 (sm_100).
// The first version of this kernel used the unswizzled "interleaved" layout (8 x 16-byte core matrices, LBO between K
// chunks): equally correct and, measured, equally fast -- the tensor pipe is not what bounds this kernel
// (profiles/r02_linear_tc.txt; a three-stage ring was measured too and was slower: 273 vs 197 us); the swizzled layout is kept because it is the one the hardware documentation centres on.
__device__ __forceinline__ uint64_t tc_smem_desc(uint32_t saddr)
{
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | (1ull << 16) | ((uint64_t)(1024u >> 4) << 32) | (1ull << 46) | (2ull << 61);
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

// 32 consecutive fp32 columns of this thread's TMEM lane (issue only; tc_ld_wait before the registers are read)
__device__ __forceinline__ void tc_ld32_issue(uint32_t taddr, uint32_t (&r)[32])
{
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tc_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tc_mbar_arrive(uint32_t bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}

// ---- layout of k_linear_tc's dynamic shared memory
//   B hi, B lo          2 x ceil(NI / 32) blocks of NOp rows x 128 bytes (32 columns of K), 128-byte swizzle
//   X stages            2 x (hi, lo) x 128 rows x 128 bytes (32 columns) = 64 KB, 128-byte swizzle
//   raw ring            3 x 128 rows x 128 bytes of fp32 on their way from global memory (cp.async), same chunk layout
//   epilogue slabs      8 warps x 32 rows x 36 floats (padded rows: conflict-free 16-byte writes by row and reads by column)
//   BatchNorm constants of the prologue (float4 per input column), bias, 8 mbarriers, the TMEM base address
// experiment counters of CTA 0 (RPC_TC_DBG = 4): cycles spent by one loader thread, the MMA thread and one epilogue thread
__device__ unsigned long long g_tc_stats[16];

constexpr int kTcLoadWarps = 8, kTcEpiWarps = 8;               // (two epilogue warps per TMEM lane quadrant)
constexpr int kTcLoaders = kTcLoadWarps * 32;
constexpr int kTcThreads2 = (kTcLoadWarps + kTcEpiWarps + 1) * 32;     // + the MMA warp
constexpr int kTcSC = 32;                                              // columns of X per stage
constexpr int kTcStageB = kTcRows * kTcSC * 4;                         // one of hi / lo: 16 KB
constexpr int kTcSlabStride = 36;                                      // floats per slab row
constexpr int kTcRaw = 3;                                              // raw fp32 stages in flight ahead of the conversion (cp.async)

__host__ __device__ inline size_t tc_smem_bytes(int NI, int NOp)
{
    return 2 * (size_t)((NI + kTcSC - 1) / kTcSC) * NOp * 128 + (4 + kTcRaw) * (size_t)kTcStageB + (size_t)kTcEpiWarps * 32 * kTcSlabStride * 4 +
           (size_t)kTcMaxDim * 16 + kTcMaxDim * 4 + 8 * 8 + 16 + 1024;               // (+ 1024: the swizzle atoms need that alignment)
}

__global__ void __launch_bounds__(kTcThreads2, 1)
k_linear_tc(const __grid_constant__ TcParams p)
{
    extern __shared__ __align__(1024) unsigned char tc_sm_raw[];
    unsigned char *tc_sm = tc_sm_raw + ((1024u - (smem_u32(tc_sm_raw) & 1023u)) & 1023u);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int NI = p.NI, NO = p.NO;
    const int NOp = (NO + 31) & ~31;                               // N of the MMA (16 | N at M = 128; 32-column epilogue groups)
    const uint32_t wblock = (uint32_t)NOp * 128;                    // one 32-column block of B
    const uint32_t wbytes = (uint32_t)((NI + kTcSC - 1) / kTcSC) * wblock;
    unsigned char *sB = tc_sm;
    unsigned char *sA = tc_sm + 2 * wbytes;
    unsigned char *sRaw = sA + 4 * kTcStageB;                        // raw fp32 ring of the loaders
    const uint32_t sRaw_u = smem_u32(sRaw);
    float *s_slab = reinterpret_cast<float *>(sRaw + kTcRaw * kTcStageB);
    float4 *s_bn = reinterpret_cast<float4 *>(s_slab + kTcEpiWarps * 32 * kTcSlabStride);
    float *s_bias = reinterpret_cast<float *>(s_bn + kTcMaxDim);
    unsigned long long *s_bar = reinterpret_cast<unsigned long long *>(s_bias + kTcMaxDim);
    uint32_t *s_tmem = reinterpret_cast<uint32_t *>(s_bar + 8);
    const uint32_t sB_u = smem_u32(sB), sA_u = smem_u32(sA), bar_u = smem_u32(s_bar);
    // barriers: [0,1] stage full (256 loader arrivals), [2,3] stage free (MMA commit), [4,5] accumulator full (commit),
    // [6,7] accumulator free (128 epilogue arrivals)
    const uint32_t bar_afull = bar_u, bar_afree = bar_u + 16, bar_cfull = bar_u + 32, bar_cfree = bar_u + 48;
    long long K = p.K;
    if (p.k_dev) K = min(K, (long long)*p.k_dev);
    const long long ntiles = (K + kTcRows - 1) / kTcRows;
    const int tmem_cols = NOp <= 32 ? 64 : NOp <= 64 ? 128 : 256;   // two accumulators
    const int nstage = (NI + kTcSC - 1) / kTcSC;                     // stages per tile (1 .. 4)
    const long long my_tiles = blockIdx.x < ntiles ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const long long njobs = my_tiles * nstage;                       // (tile, stage) pairs of this CTA, in order

    // ---- set-up: barriers, tensor memory, weights (hi / lo, canonical layout), prologue / epilogue constants
    if (tid == 0) {
        tc_mbar_init(bar_afull, kTcLoaders); tc_mbar_init(bar_afull + 8, kTcLoaders);
        tc_mbar_init(bar_afree, 1); tc_mbar_init(bar_afree + 8, 1);
        tc_mbar_init(bar_cfull, 1); tc_mbar_init(bar_cfull + 8, 1);
        tc_mbar_init(bar_cfree, kTcEpiWarps * 32); tc_mbar_init(bar_cfree + 8, kTcEpiWarps * 32);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == kTcLoadWarps + kTcEpiWarps) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)), "r"(tmem_cols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    {
        const int nch = NI >> 2;                                       // 16-byte chunks per row
        for (int e = tid; e < NOp * nch; e += kTcThreads2) {
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
            const uint32_t off = (uint32_t)(c >> 3) * wblock + (uint32_t)n * 128 + (uint32_t)(((c & 7) ^ (n & 7)) * 16);
            *reinterpret_cast<uint4 *>(sB + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<uint4 *>(sB + wbytes + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        }
        if (p.in_mean)
            for (int c = tid; c < NI; c += kTcThreads2)
                s_bn[c] = make_float4(__ldg(p.in_mean + c), __ldg(p.in_invstd + c), __ldg(p.in_gamma + c), __ldg(p.in_beta + c));
        for (int c = tid; c < kTcMaxDim; c += kTcThreads2) s_bias[c] = (p.bias && c < NO) ? __ldg(p.bias + c) : 0.0f;
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *s_tmem;

    if (warp < kTcLoadWarps) {
        // ================= loaders: global -> (cp.async) raw ring -> (BatchNorm + ReLU) -> hi / lo -> shared-memory stage
        // A warp instruction covers 8 rows x 4 chunks: 64 contiguous bytes per row from global memory, and per quarter-warp
        // 128 contiguous bytes of shared memory (conflict-free 16-byte accesses).  The raw fp32 chunks travel with cp.async
        // into a three-deep ring (kTcRaw stages ahead of the conversion, no registers held: with the loads in registers the
        // compiler sank them to their first use under the 96-register cap of a 17-warp CTA and every stage paid the memory
        // latency -- 1730 cycles of "work" per stage, profiles/r02_linear_tc.txt); every thread reads back exactly the
        // chunks it copied, so the ring needs no barrier, only cp.async.wait_group.
        const int r8 = lane & 7, cq = lane >> 3;
        long long st_wait = 0, st_work = 0;
        auto chunk_off = [&](int u, int &ch) -> uint32_t {
            const int jj = warp + 8 * u, rg = jj >> 1;
            ch = (jj & 1) * 4 + cq;
            return (uint32_t)(rg * 8 + r8) * 128u + (uint32_t)((ch ^ r8) * 16);
        };
        auto issue = [&](bool live, int rb, long long tile, int s) {
            if (live) {
                const int col0 = s * kTcSC, ncol = min(kTcSC, NI - col0);
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    int ch;
                    const uint32_t off = chunk_off(u, ch);
                    if (ch * 4 >= ncol) continue;
                    const long long row = tile * kTcRows + ((warp + 8 * u) >> 1) * 8 + r8;
                    const bool ok = row < K;                            // rows beyond K are zero-filled (src-size 0)
                    const float *src = p.x + (ok ? row : 0) * NI + col0 + ch * 4;
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(sRaw_u + (uint32_t)rb * kTcStageB + off),
                                 "l"(src), "r"(ok ? 16 : 0) : "memory");
                }
            }
            asm volatile("cp.async.commit_group;" ::: "memory");          // (empty groups keep the count uniform)
        };
        long long tile_i = blockIdx.x, tile_c = blockIdx.x;
        int s_i = 0, s_c = 0;
        for (int d = 0; d < kTcRaw; ++d) {
            issue(d < njobs, d, tile_i, s_i);
            if (++s_i == nstage) { s_i = 0; tile_i += gridDim.x; }
        }
        int rb = 0;
        for (long long j = 0; j < njobs; ++j) {
            const int col0 = s_c * kTcSC, ncol = min(kTcSC, NI - col0);
            const int buf = (int)(j & 1);
            asm volatile("cp.async.wait_group %0;" ::"n"(kTcRaw - 1) : "memory");   // this thread's chunks of job j have landed
            float4 v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                int ch;
                const uint32_t off = chunk_off(u, ch);
                v[u] = *reinterpret_cast<const float4 *>(sRaw + (size_t)rb * kTcStageB + off);
            }
            issue(j + kTcRaw < njobs, rb, tile_i, s_i);                   // refill the slot just read
            if (++s_i == nstage) { s_i = 0; tile_i += gridDim.x; }
            const long long t0 = clock64();
            if (j >= 2) tc_mbar_wait(bar_afree + 8 * buf, (uint32_t)((j >> 1) - 1) & 1u);     // the MMAs that read this stage are done
            const long long t1 = clock64();
            st_wait += t1 - t0;
            unsigned char *a_hi = sA + (size_t)buf * 2 * kTcStageB, *a_lo = a_hi + kTcStageB;
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                int ch;
                const uint32_t off = chunk_off(u, ch);
                if (ch * 4 >= ncol) continue;
                float e[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
                if (p.in_mean) {
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const float4 b = s_bn[col0 + ch * 4 + q];
                        e[q] = bn_relu_value(e[q], b.x, b.y, b.z, b.w);
                    }
                }
                uint32_t hi[4], lo[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) split_tf32(e[q], hi[q], lo[q]);
                *reinterpret_cast<uint4 *>(a_hi + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                *reinterpret_cast<uint4 *>(a_lo + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            tc_mbar_arrive(bar_afull + 8 * buf);
            st_work += clock64() - t1;
            if (++s_c == nstage) { s_c = 0; tile_c += gridDim.x; }
            if (++rb == kTcRaw) rb = 0;
        }
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        if (p.dbg == 4 && blockIdx.x == 0 && tid == 0) { g_tc_stats[0] = st_wait; g_tc_stats[1] = st_work; g_tc_stats[2] = njobs; }
    } else if (warp == kTcLoadWarps + kTcEpiWarps) {
        // ================= MMA issuer: one thread
        if (lane == 0) {
            const uint32_t idesc = tc_instr_desc(NOp);
            long long mw_c = 0, mw_a = 0, mi = 0;
            long long it = 0;
            int s = 0;
            for (long long j = 0; j < njobs; ++j) {
                const int acc = (int)(it & 1), buf = (int)(j & 1);
                const int col0 = s * kTcSC, ncol = min(kTcSC, NI - col0);
                const long long t0 = clock64();
                if (s == 0 && it >= 2) tc_mbar_wait(bar_cfree + 8 * acc, (uint32_t)((it >> 1) - 1) & 1u);   // epilogue has drained it
                const long long t1 = clock64();
                tc_mbar_wait(bar_afull + 8 * buf, (uint32_t)(j >> 1) & 1u);
                const long long t2 = clock64();
                mw_c += t1 - t0; mw_a += t2 - t1;
                tc_fence_after();
                const uint32_t a_hi_u = sA_u + (uint32_t)buf * 2 * kTcStageB, a_lo_u = a_hi_u + kTcStageB;
                const uint32_t d = tmem_base + (uint32_t)(acc * NOp);
                for (int k = 0; k < (ncol >> 3); ++k) {                           // K = 8 per instruction: 32 bytes of every row
                    const uint32_t ko = (uint32_t)k * 32u, bo = (uint32_t)s * wblock + ko;
                    const uint64_t ah = tc_smem_desc(a_hi_u + ko), al = tc_smem_desc(a_lo_u + ko);
                    const uint64_t bh = tc_smem_desc(sB_u + bo), bl = tc_smem_desc(sB_u + wbytes + bo);
                    if (p.dbg == 0) {
                        tc_mma_tf32(d, al, bh, idesc, (s | k) ? 1u : 0u);         // small terms first
                        tc_mma_tf32(d, ah, bl, idesc, 1u);
                        tc_mma_tf32(d, ah, bh, idesc, 1u);
                    } else if (p.dbg == 1) {
                        tc_mma_tf32(d, ah, bh, idesc, (s | k) ? 1u : 0u);
                    }
                }
                tc_commit(bar_afree + 8 * buf);                                   // stage free when these MMAs have read it
                if (s == nstage - 1) tc_commit(bar_cfull + 8 * acc);              // accumulator complete
                mi += clock64() - t2;
                if (++s == nstage) { s = 0; ++it; }
            }
            if (p.dbg == 4 && blockIdx.x == 0) { g_tc_stats[3] = mw_c; g_tc_stats[4] = mw_a; g_tc_stats[5] = mi; }
        }
        __syncwarp();
    } else {
        // ================= epilogue: TMEM -> registers -> (+ bias, tanh) -> padded slab -> full-line stores, column moments
        const int q = warp & 3;                                    // TMEM lane quadrant of this warp (warp id % 4)
        const int half = (warp - kTcLoadWarps) >> 2;               // the quadrant's two warps take alternate column groups
        float *slab = s_slab + (size_t)(warp - kTcLoadWarps) * 32 * kTcSlabStride;
        double m_sum[2] = {0.0, 0.0}, m_sq[2] = {0.0, 0.0};        // lane = column (2 * u + half) * 32 + lane
        const int ngrp = NOp >> 5;
        long long ew = 0, el = 0, es = 0;
        for (long long it = 0; it < my_tiles; ++it) {
            const long long tile = blockIdx.x + it * gridDim.x;
            const int acc = (int)(it & 1);
            const long long e0 = clock64();
            tc_mbar_wait(bar_cfull + 8 * acc, (uint32_t)(it >> 1) & 1u);
            const long long e1 = clock64();
            ew += e1 - e0;
            tc_fence_after();
            const long long row0 = tile * kTcRows + q * 32;
            const int nrow = (int)max(0LL, min(32LL, K - row0));   // live rows of this warp's slab
            uint32_t r[2][32];
#pragma unroll
            for (int u = 0; u < 2; ++u) {                          // all of this warp's columns first: the accumulator is
                const int g = 2 * u + half;                        // released before the slab work starts
                if (g < ngrp) tc_ld32_issue(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * NOp + g * 32), r[u]);
            }
            tc_ld_wait();
            tc_fence_before();
            tc_mbar_arrive(bar_cfree + 8 * acc);
            const long long e2 = clock64();
            el += e2 - e1;
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                const int g = 2 * u + half;
                if (g >= ngrp) break;
                __syncwarp();                                      // the previous group's slab reads are done
                const bool live_row = lane < nrow;                 // rows beyond K go into the slab as zeros (moments)
#pragma unroll
                for (int i = 0; i < 32; i += 4) {
                    const float4 b4 = *reinterpret_cast<const float4 *>(s_bias + g * 32 + i);
                    float o[4] = {__fadd_rn(__uint_as_float(r[u][i]), b4.x), __fadd_rn(__uint_as_float(r[u][i + 1]), b4.y),
                                  __fadd_rn(__uint_as_float(r[u][i + 2]), b4.z), __fadd_rn(__uint_as_float(r[u][i + 3]), b4.w)};
                    if (p.epi_tanh) {
#pragma unroll
                        for (int c = 0; c < 4; ++c) o[c] = tanhf(o[c]);
                    }
                    *reinterpret_cast<float4 *>(slab + lane * kTcSlabStride + i) =
                        live_row ? make_float4(o[0], o[1], o[2], o[3]) : make_float4(0.f, 0.f, 0.f, 0.f);
                }
                __syncwarp();
                if (p.moments && p.red_v) {                        // BatchNorm-backward reductions of the layer below
                    const int col = g * 32 + lane;
                    const bool okc = col < NO;
                    const float mean = okc ? __ldg(p.red_mean + col) : 0.f, invstd = okc ? __ldg(p.red_invstd + col) : 0.f;
                    const float ga = okc ? __ldg(p.red_gamma + col) : 0.f, be = okc ? __ldg(p.red_beta + col) : 0.f;
                    float s1[4] = {0.f, 0.f, 0.f, 0.f}, s2[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                    for (int r0 = 0; r0 < 32; r0 += 8) {
                        float vv[8];
#pragma unroll
                        for (int q = 0; q < 8; ++q)            // a row of the group is 128 contiguous bytes across the lanes
                            vv[q] = (okc && r0 + q < nrow) ? __ldg(p.red_v + (row0 + r0 + q) * NO + col) : 0.0f;
#pragma unroll
                        for (int q = 0; q < 8; ++q) {
                            const float gv = slab[(r0 + q) * kTcSlabStride + lane];
                            const float xh = __fmul_rn(__fsub_rn(vv[q], mean), invstd);
                            const float t = __fadd_rn(__fmul_rn(xh, ga), be);
                            const float gp = t <= 0.0f ? 0.0f : gv;    // threshold_backward's rule, as k_bn_relu_bwd_reduce
                            s1[q & 3] += gp;
                            s2[q & 3] = fmaf(gp, xh, s2[q & 3]);
                        }
                    }
                    m_sum[u] += (double)((s1[0] + s1[1]) + (s1[2] + s1[3]));
                    m_sq[u] += (double)((s2[0] + s2[1]) + (s2[2] + s2[3]));
                } else if (p.moments) {                            // lane = column: four independent partial sums
                    float s1[4] = {0.f, 0.f, 0.f, 0.f}, s2[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                    for (int rr = 0; rr < 32; ++rr) {
                        const float t = slab[rr * kTcSlabStride + lane];
                        s1[rr & 3] += t;
                        s2[rr & 3] = fmaf(t, t, s2[rr & 3]);
                    }
                    m_sum[u] += (double)((s1[0] + s1[1]) + (s1[2] + s1[3]));
                    m_sq[u] += (double)((s2[0] + s2[1]) + (s2[2] + s2[3]));
                }
                const int c4 = lane & 7;
                if (p.dbg == 5) {                                  // experiment: no global stores (203 -> 109 us at 64 -> 128)
                } else if ((NO & 3) == 0) {
                    if (g * 32 + c4 * 4 < NO) {
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const int rr = i * 4 + (lane >> 3);
                            if (rr < nrow)
                                *reinterpret_cast<float4 *>(p.y + (row0 + rr) * NO + g * 32 + c4 * 4) =
                                    *reinterpret_cast<const float4 *>(slab + rr * kTcSlabStride + c4 * 4);
                        }
                    }
                } else if (g * 32 + lane < NO) {                    // odd widths (the 5-wide output layer): lane = column
                    for (int rr = 0; rr < nrow; ++rr) p.y[(row0 + rr) * NO + g * 32 + lane] = slab[rr * kTcSlabStride + lane];
                }
            }
            es += clock64() - e2;
        }
        if (p.dbg == 4 && blockIdx.x == 0 && warp == kTcLoadWarps && lane == 0) {
            g_tc_stats[6] = ew; g_tc_stats[7] = el; g_tc_stats[8] = es; g_tc_stats[9] = my_tiles;
        }
        if (p.moments) {
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                const int g = 2 * u + half, c = g * 32 + lane;
                if (g < ngrp && c < NO) {
                    atomicAdd(p.moments + c, m_sum[u]);
                    atomicAdd(p.moments + NO + c, m_sq[u]);
                }
            }
        }
    }

    // ---- tensor memory back
    tc_fence_before();
    __syncthreads();
    if (warp == kTcLoadWarps + kTcEpiWarps)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(tmem_cols) : "memory");
}

// ---------------------------------------------------------------------------------------------------------------
// Stream: the first layer of the trunk (NI = C <= 8 inputs) and the input gradient of the last layer.  Same TcParams;
// fp32 FMA, weights in shared memory.
constexpr int kNarrowThreads = 256;

// NI <= 8: thread = (row, four output columns); a row's threads are neighbours, so the stores are contiguous.  Four rows
// are in flight per thread.
constexpr int kNarrowU = 4;
__global__ void __launch_bounds__(kNarrowThreads)
k_linear_narrow_in(const __grid_constant__ TcParams p)
{
    __shared__ __align__(16) float s_w[8 * kTcMaxDim];               // [c][m]
    __shared__ __align__(16) float s_b[kTcMaxDim];
    __shared__ float4 s_bn[8];
    __shared__ double s_m[2 * kTcMaxDim];
    const int NI = p.NI, NO = p.NO, tid = threadIdx.x;
    for (int e = tid; e < NI * NO; e += kNarrowThreads) {
        const int c = e / NO, m = e - c * NO;
        s_w[c * NO + m] = p.w_layout == 0 ? __ldg(p.w + m * NI + c) : __ldg(p.w + c * NO + m);
    }
    for (int m = tid; m < NO; m += kNarrowThreads) s_b[m] = p.bias ? __ldg(p.bias + m) : 0.0f;
    for (int m = tid; m < 2 * NO; m += kNarrowThreads) s_m[m] = 0.0;
    if (p.in_mean && tid < NI)
        s_bn[tid] = make_float4(__ldg(p.in_mean + tid), __ldg(p.in_invstd + tid), __ldg(p.in_gamma + tid), __ldg(p.in_beta + tid));
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
        const long long total = K * gpr;
        for (long long e0 = t0; e0 < total; e0 += kNarrowU * stride) {
            float xv[kNarrowU][8];
#pragma unroll
            for (int u = 0; u < kNarrowU; ++u) {
                const long long e = e0 + u * stride;
                const long long row = e / gpr;
#pragma unroll
                for (int c = 0; c < 8; ++c) xv[u][c] = (e < total && c < NI) ? __ldg(p.x + row * NI + c) : 0.0f;
            }
#pragma unroll
            for (int u = 0; u < kNarrowU; ++u) {
                const long long e = e0 + u * stride;
                if (e >= total) break;
                const long long row = e / gpr;
                float acc[4] = {s_b[m0], s_b[m0 + 1], s_b[m0 + 2], s_b[m0 + 3]};
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    if (c >= NI) break;
                    float t = xv[u][c];
                    if (p.in_mean) { const float4 b = s_bn[c]; t = bn_relu_value(t, b.x, b.y, b.z, b.w); }
                    const float4 w4 = *reinterpret_cast<const float4 *>(s_w + c * NO + m0);
                    acc[0] = fmaf(t, w4.x, acc[0]); acc[1] = fmaf(t, w4.y, acc[1]);
                    acc[2] = fmaf(t, w4.z, acc[2]); acc[3] = fmaf(t, w4.w, acc[3]);
                }
                if (p.epi_tanh)
                    for (int q = 0; q < 4; ++q) acc[q] = tanhf(acc[q]);
                *reinterpret_cast<float4 *>(p.y + row * NO + m0) = make_float4(acc[0], acc[1], acc[2], acc[3]);
#pragma unroll
                for (int q = 0; q < 4; ++q) { ms[q] += acc[q]; mq[q] = fmaf(acc[q], acc[q], mq[q]); }
            }
        }
    }
    if (p.moments) {
        for (int q = 0; q < 4; ++q) { atomicAdd(s_m + m0 + q, (double)ms[q]); atomicAdd(s_m + NO + m0 + q, (double)mq[q]); }
        __syncthreads();
        for (int m = tid; m < 2 * NO; m += kNarrowThreads) atomicAdd(p.moments + m, s_m[m]);
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

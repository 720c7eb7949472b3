// allreduce.cuh -- one-shot all-reduce (sum) of the perturber's flat parameter-gradient buffer over NVLink peer
// memory (the only exchange of the path; the reference gets it from DDP, custom_hook.py:24-28).
//
// The buffer is small (34 641 fp32 = 139 KB for C = 4), so the collective is pure latency: NCCL's all-reduce
// costs ~17 us at 2 GPUs -- more than any phase of the forward.  Here ONE kernel per rank, with launch
// parameters that never change (so it sits in the step's CUDA graph):
//   0. the call number e comes from a device counter (completed calls + 1); staging buffer = e & 1;
//   1. CTA b copies ITS slice of the local gradients into the rank's peer-mapped staging buffer, then tells every
//      peer "slice b of rank r is staged for call e" (system-scope release store of e into the peer's flag block)
//      and waits until every peer has said the same (acquire loads, bounded by a timeout);
//   2. out[i] = sum_r stage_r[i] over the same slice, peers read through NVLink with uncached 16-byte loads, all
//      of them in flight before the first add, added in rank order (every rank gets bit-identical sums);
//   3. the last CTA to finish publishes e as the number of completed calls.
// Staging is double-buffered: a rank can only start call e + 2, which overwrites buffer e & 1, after every peer
// has entered call e + 1, i.e. finished reading in call e -- no second barrier.  Epochs only grow, flags never
// need a reset.
#pragma once
#include "common.cuh"

namespace rpc {

constexpr int kArCtas = 64;                       // at most; one float4 per thread at the named sizes
constexpr int kArThreads = 256;
constexpr int kArLLCtas = 136;                    // low-latency variant: one element per thread at the named sizes
constexpr int kArMaxWorld = 16;

struct ArParams {
    float *stage[2][kArMaxWorld];        // peer-mapped staging buffers [buffer][rank]
    uint32_t *flags[kArMaxWorld];        // peer-mapped flag blocks [kArCtas][kArMaxWorld], index = rank
    const float *src;                    // local values
    float *out;                          // local result (may alias src)
    uint32_t *ctr;                       // local: [0] completed calls, [1] CTA ticket
    int *status;                         // local: set to 1 on timeout
    long long n;
    int rank, world;
};

__global__ void __launch_bounds__(kArThreads)
k_allreduce_oneshot(const __grid_constant__ ArParams p)
{
    __shared__ uint32_t s_e;
    const int tid = threadIdx.x, b = blockIdx.x;
    if (tid == 0) s_e = *reinterpret_cast<volatile uint32_t *>(p.ctr) + 1u;
    __syncthreads();
    const uint32_t e = s_e;
    const int buf = (int)(e & 1u);
    const long long n4 = p.n >> 2;
    const int tail = (int)(p.n & 3);
    // 1. stage this CTA's slice
    float *mine = p.stage[buf][p.rank];
    for (long long i = (long long)b * kArThreads + tid; i < n4; i += (long long)gridDim.x * kArThreads)
        reinterpret_cast<float4 *>(mine)[i] = __ldg(reinterpret_cast<const float4 *>(p.src) + i);
    if (b == 0 && tid < tail) mine[(n4 << 2) + tid] = __ldg(p.src + (n4 << 2) + tid);
    __syncthreads();
    if (tid < p.world) {
        uint32_t *dst = p.flags[tid] + b * kArMaxWorld + p.rank;       // my slot in peer tid's block
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(dst), "r"(e) : "memory");
        const uint32_t *src = p.flags[p.rank] + b * kArMaxWorld + tid; // peer tid's slot in my block
        unsigned long long t0, t1;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t0));
        for (;;) {
            uint32_t v;
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(src) : "memory");
            if ((int32_t)(v - e) >= 0) break;
            asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t1));
            if (t1 - t0 > 2000000000ull) { *p.status = 1; break; }     // 2 s: a peer died; do not hang the GPU
        }
    }
    __syncthreads();
    // 2. reduce the same slice
    for (long long i = (long long)b * kArThreads + tid; i < n4; i += (long long)gridDim.x * kArThreads) {
        float4 v[kArMaxWorld];
#pragma unroll
        for (int r = 0; r < kArMaxWorld; ++r)                          // all peers' loads in flight before the first add
            if (r < p.world) v[r] = __ldcv(reinterpret_cast<const float4 *>(p.stage[buf][r]) + i);
        float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int r = 0; r < kArMaxWorld; ++r)
            if (r < p.world) {
                s.x = __fadd_rn(s.x, v[r].x); s.y = __fadd_rn(s.y, v[r].y);
                s.z = __fadd_rn(s.z, v[r].z); s.w = __fadd_rn(s.w, v[r].w);
            }
        reinterpret_cast<float4 *>(p.out)[i] = s;
    }
    if (b == 0 && tid < tail) {
        const long long i = (n4 << 2) + tid;
        float s = 0.0f;
        for (int r = 0; r < p.world; ++r) s = __fadd_rn(s, __ldcv(p.stage[buf][r] + i));
        p.out[i] = s;
    }
    // 3. completion
    __syncthreads();
    if (tid == 0) {
        const uint32_t t = atomicAdd(p.ctr + 1, 1u);
        if (t == gridDim.x - 1) {
            p.ctr[1] = 0u;
            __threadfence();
            *reinterpret_cast<volatile uint32_t *>(p.ctr) = e;
        }
    }
}

// ---- low-latency variant: data and flag travel in the same 8-byte store -------------------------------------------
// k_allreduce_oneshot above needs three dependent trips over NVLink per call (release flag stores, acquire flag
// loads, then the remote data loads): ~16 us per step at 2-8 GPUs.  Here every rank PUSHES: element i of rank r goes
// as one 8-byte word {bits(value), call number e} straight into slot [r][i] of every peer's receive area (and of its
// own), an aligned 8-byte store cannot tear, so a receiver that polls its LOCAL slot until the call number matches
// has the value -- no fence, no flag round trip, one one-way NVLink latency in total.  Sums are taken in rank order
// (bit-identical on every rank).  Receive areas are double-buffered by the parity of e, for the same reason as above:
// a rank can only start call e + 2 after every peer has sent for call e + 1, i.e. has finished polling call e.
struct ArLLParams {
    uint2 *recv[2][kArMaxWorld];         // [buffer][destination rank]: that rank's receive area, [world][npad] words
    const float *src;
    float *out;                          // may alias src
    uint32_t *ctr;                       // local: [0] completed calls, [1] CTA ticket
    int *status;                         // local: set to 1 on timeout
    long long n, npad;
    int rank, world;
};

__global__ void __launch_bounds__(kArThreads)
k_allreduce_ll(const __grid_constant__ ArLLParams p)
{
    __shared__ uint32_t s_e;
    const int tid = threadIdx.x;
    if (tid == 0) s_e = *reinterpret_cast<volatile uint32_t *>(p.ctr) + 1u;
    __syncthreads();
    const uint32_t e = s_e;
    const int buf = (int)(e & 1u);
    const long long stride = (long long)gridDim.x * kArThreads;
    const long long i0 = (long long)blockIdx.x * kArThreads + tid;
    // 1. push my values to everybody (peers first, my own area last)
    for (long long i = i0; i < p.n; i += stride) {
        const uint32_t bits = __float_as_uint(__ldg(p.src + i));
#pragma unroll
        for (int k = 1; k <= kArMaxWorld; ++k) {
            if (k <= p.world) {
                const int r = (p.rank + k) % p.world;
                uint2 *dst = p.recv[buf][r] + (long long)p.rank * p.npad + i;
                asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" ::"l"(dst), "r"(bits), "r"(e) : "memory");
            }
        }
    }
    // 2. poll my own area, add in rank order
    const uint2 *mine = p.recv[buf][p.rank];
    unsigned long long t0;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t0));
    for (long long i = i0; i < p.n; i += stride) {
        uint32_t v[kArMaxWorld];
        for (;;) {                                                  // all sources polled together: one L2 round trip
            bool all = true;
#pragma unroll
            for (int r = 0; r < kArMaxWorld; ++r) {
                if (r < p.world) {
                    uint32_t f;
                    asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];"
                                 : "=r"(v[r]), "=r"(f) : "l"(mine + (long long)r * p.npad + i) : "memory");
                    all &= f == e;
                }
            }
            if (all) break;
            unsigned long long t1;
            asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t1));
            if (t1 - t0 > 2000000000ull) { *p.status = 1; break; }   // 2 s: a peer died; do not hang the GPU
        }
        float s = 0.0f;
#pragma unroll
        for (int r = 0; r < kArMaxWorld; ++r)
            if (r < p.world) s = __fadd_rn(s, __uint_as_float(v[r]));
        p.out[i] = s;
    }
    // 3. completion
    __syncthreads();
    if (tid == 0) {
        const uint32_t t = atomicAdd(p.ctr + 1, 1u);
        if (t == gridDim.x - 1) {
            p.ctr[1] = 0u;
            __threadfence();
            *reinterpret_cast<volatile uint32_t *>(p.ctr) = e;
        }
    }
}

}  // namespace rpc

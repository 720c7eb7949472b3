// Microbenchmark: chip-wide throughput of random 32-byte-sector loads and random atomics on an L2-resident table,
// the access pattern of the insert / rank phases.  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o atomics atomics.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t mix(uint32_t h) { h *= 0x9E3779B1u; h ^= h >> 15; h *= 0x85EBCA77u; h ^= h >> 13; return h; }

template <int MODE>
__global__ void __launch_bounds__(1024, 2) k(unsigned long long *tab, uint32_t size, int n, int reps, unsigned long long *sink, int hot)
{
    unsigned long long acc = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        for (int r = 0; r < reps; ++r) {
            uint32_t key = mix((uint32_t)i * 31u + r);
            if (hot) key = mix(key % (uint32_t)hot);                   // `hot` distinct addresses: same-address contention
            const uint32_t s = __umulhi(key, size);
            if (MODE == 0) acc += __ldcg(tab + s);                                         // 8-byte load
            if (MODE == 1) { ulonglong2 v = __ldcg(reinterpret_cast<const ulonglong2 *>(tab) + (s >> 1)); acc += v.x ^ v.y; }
            if (MODE == 2) acc += atomicCAS(tab + s, ~0ull, ((unsigned long long)key << 32) | (uint32_t)i);
            if (MODE == 3) acc += atomicMin(tab + s, ((unsigned long long)key << 32) | (uint32_t)i);
            if (MODE == 4) acc += atomicMin(reinterpret_cast<unsigned int *>(tab) + 2 * s, key);
            if (MODE == 5) atomicMin(reinterpret_cast<unsigned int *>(tab) + 2 * s, key);  // no return value: RED
            if (MODE == 6) acc += atomicOr(reinterpret_cast<unsigned int *>(tab) + 2 * s, 1u << (key & 31));
            if (MODE == 7) tab[s] = key;                                                   // 8-byte store
            if (MODE == 8) acc += atomicExch(tab + s, ((unsigned long long)key << 32) | (uint32_t)i);
        }
    }
    if (acc == 0x123456789ull) *sink = acc;
}

template <int MODE>
void run(const char *name, unsigned long long *tab, uint32_t size, int n, int reps, unsigned long long *sink, int hot)
{
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    float best = 1e9f;
    for (int t = 0; t < 6; ++t) {
        cudaMemset(tab, 0xFF, 8ull * size);
        cudaEventRecord(a);
        k<MODE><<<296, 1024>>>(tab, size, n, reps, sink, hot);
        cudaEventRecord(b);
        cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b);
        if (t > 0 && ms < best) best = ms;
    }
    printf("%-28s n=%8d reps=%d table=%6.1f MB hot=%7d: %8.1f us  %7.1f G ops/s\n", name, n, reps, 8.0 * size / 1e6, hot,
           best * 1e3, (double)n * reps / (best * 1e-3) / 1e9);
}

int main()
{
    unsigned long long *tab, *sink;
    const uint32_t maxsize = 64u << 20;
    cudaMalloc(&tab, 8ull * maxsize); cudaMalloc(&sink, 8);
    for (uint32_t size : {1080000u, 8u << 20, 64u << 20}) {
        for (int n : {720000, 4 * 720000}) {
            run<0>("ld.8B", tab, size, n, 1, sink, 0);
            run<1>("ld.16B", tab, size, n, 1, sink, 0);
            run<7>("st.8B", tab, size, n, 1, sink, 0);
            run<2>("atomicCAS.64", tab, size, n, 1, sink, 0);
            run<3>("atomicMin.64", tab, size, n, 1, sink, 0);
            run<8>("atomicExch.64", tab, size, n, 1, sink, 0);
            run<4>("atomicMin.32", tab, size, n, 1, sink, 0);
            run<5>("red.min.32 (no return)", tab, size, n, 1, sink, 0);
            run<6>("atomicOr.32", tab, size, n, 1, sink, 0);
        }
    }
    // dependent chains (reps serial ops per thread) and same-address contention
    run<3>("atomicMin.64 chain", tab, 1080000u, 720000, 4, sink, 0);
    run<1>("ld.16B chain", tab, 1080000u, 720000, 4, sink, 0);
    for (int hot : {90000, 10000, 1000, 100, 1}) run<3>("atomicMin.64 contended", tab, 1080000u, 720000, 1, sink, hot);
    for (int hot : {90000, 1000, 1}) run<4>("atomicMin.32 contended", tab, 1080000u, 720000, 1, sink, hot);
    for (int hot : {90000, 1000, 1}) run<5>("red.min.32 contended", tab, 1080000u, 720000, 1, sink, hot);
    return 0;
}

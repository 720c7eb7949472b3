// microbenchmark: random-address L2 traffic rates on B200 (what bounds the hash phases)
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t mix(uint32_t x){ x*=0x9E3779B1u; x^=x>>15; x*=0x85EBCA77u; x^=x>>13; return x; }
template<int MODE> __global__ void __launch_bounds__(1024,1) k(unsigned long long* t64, uint32_t* t32, uint32_t size, int per, uint32_t* sink)
{
    uint32_t acc=0; uint32_t g = blockIdx.x*1024+threadIdx.x;
    for(int it=0; it<per; it+=4){
        uint32_t a[4]; unsigned long long r[4];
        #pragma unroll
        for(int k2=0;k2<4;++k2) a[k2]=__umulhi(mix(g*64u+it+k2), size);
        #pragma unroll
        for(int k2=0;k2<4;++k2){
            if(MODE==0) r[k2]=atomicCAS(&t64[a[k2]], ~0ull, (unsigned long long)g);
            if(MODE==1) r[k2]=atomicCAS(&t32[a[k2]], ~0u, g);
            if(MODE==2) r[k2]=atomicMin(&t32[a[k2]], g);
            if(MODE==3){ atomicMin(&t32[a[k2]], g); r[k2]=0; }
            if(MODE==4) r[k2]=atomicMin(&t64[a[k2]], (unsigned long long)g);
            if(MODE==5){ atomicMin(&t64[a[k2]], (unsigned long long)g); r[k2]=0; }
            if(MODE==6) r[k2]=__ldcg(&t64[a[k2]]);
            if(MODE==7){ t32[a[k2]]=g; r[k2]=0; }
            if(MODE==8) r[k2]=__ldcg(&t32[a[k2]]);
            if(MODE==9) r[k2]=atomicAdd(&t32[a[k2]], 1u);
            if(MODE==10) r[k2]=atomicExch(&t64[a[k2]], (unsigned long long)g);
        }
        #pragma unroll
        for(int k2=0;k2<4;++k2) acc+=(uint32_t)r[k2];
    }
    if(acc==0x12345678u) sink[0]=acc;
}
template<int MODE> void run(const char* nm, unsigned long long* t64, uint32_t* t32, uint32_t size, int per, uint32_t* sink, int grid){
    cudaEvent_t a,b; cudaEventCreate(&a); cudaEventCreate(&b);
    float best=1e9;
    for(int rep=0;rep<5;++rep){
        cudaMemset(t64,0xFF,(size_t)size*8); cudaMemset(t32,0xFF,(size_t)size*4);
        cudaEventRecord(a); k<MODE><<<grid,1024>>>(t64,t32,size,per,sink); cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms,a,b); if(ms<best) best=ms;
    }
    double ops=(double)grid*1024*per;
    printf("%-28s size %8u ops %8.0f  %8.2f us  %7.2f Gop/s  %6.2f cyc/op/SM@1.9GHz\n", nm, size, ops, best*1e3, ops/best/1e6, best*1e-3*1.9e9/(ops/grid));
}
int main(){
    int grid=148; uint32_t* sink; cudaMalloc(&sink,4);
    for(uint32_t size : {1u<<20, 1u<<23}){
        unsigned long long* t64; uint32_t* t32; cudaMalloc(&t64,(size_t)size*8); cudaMalloc(&t32,(size_t)size*4);
        for(int per : {4, 16}){
            printf("--- per-thread ops %d\n", per);
            run<0>("CAS u64 (ret)",t64,t32,size,per,sink,grid);
            run<1>("CAS u32 (ret)",t64,t32,size,per,sink,grid);
            run<2>("min u32 (ret)",t64,t32,size,per,sink,grid);
            run<3>("min u32 (red)",t64,t32,size,per,sink,grid);
            run<4>("min u64 (ret)",t64,t32,size,per,sink,grid);
            run<5>("min u64 (red)",t64,t32,size,per,sink,grid);
            run<9>("add u32 (ret)",t64,t32,size,per,sink,grid);
            run<10>("exch u64 (ret)",t64,t32,size,per,sink,grid);
            run<6>("load u64",t64,t32,size,per,sink,grid);
            run<8>("load u32",t64,t32,size,per,sink,grid);
            run<7>("store u32",t64,t32,size,per,sink,grid);
        }
        cudaFree(t64); cudaFree(t32);
    }
    return 0;
}

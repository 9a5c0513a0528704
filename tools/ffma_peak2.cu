// FP32 peak microbenchmark v2 (sm_100a): pure FMA streams with loop-invariant operands.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("CUDA error %s at %d\n",cudaGetErrorString(e),__LINE__); exit(1);} }while(0)
typedef unsigned long long u64;
__device__ __forceinline__ void fma2(u64& d, u64 a, u64 b){ asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(d) : "l"(a), "l"(b)); }
__device__ __forceinline__ u64 pack(float x,float y){ u64 r; asm("mov.b64 %0, {%1,%2};":"=l"(r):"f"(x),"f"(y)); return r;}
__device__ __forceinline__ float fold(u64 v){ float x,y; asm("mov.b64 {%0,%1}, %2;":"=f"(x),"=f"(y):"l"(v)); return x+y; }

// MODE 0: scalar 8x8 ; 1: packed, a broadcast (a,a) ; 2: packed, both operands true pairs (4x(2) x 8 as acc[i/2 pair][j]) 
template<int MODE,int ITERS>
__global__ void __launch_bounds__(256,2) kern(float* out, const float* in, int n) {
    float a[8], b[8];
    for (int i=0;i<8;i++){ a[i]=in[(threadIdx.x+i)%n]; b[i]=in[(threadIdx.x*3+i)%n]; }
    float s=0;
    if (MODE==0) {
        float acc[8][8];
        for (int i=0;i<8;i++) for(int j=0;j<8;j++) acc[i][j]=(float)(i+j);
        #pragma unroll 1
        for (int it=0; it<ITERS; ++it) {
            #pragma unroll
            for (int u=0;u<4;u++)
            #pragma unroll
            for (int i=0;i<8;i++)
                #pragma unroll
                for (int j=0;j<8;j++) asm volatile("fma.rn.f32 %0, %1, %2, %0;":"+f"(acc[i][j]):"f"(a[i]),"f"(b[j]));
        }
        for(int i=0;i<8;i++) for(int j=0;j<8;j++) s+=acc[i][j];
    } else if (MODE==1) {
        u64 acc[8][4], b2[4];
        for (int j=0;j<4;j++) b2[j]=pack(b[2*j],b[2*j+1]);
        for (int i=0;i<8;i++) for(int j=0;j<4;j++) acc[i][j]=pack((float)i,(float)j);
        #pragma unroll 1
        for (int it=0; it<ITERS; ++it) {
            #pragma unroll
            for (int u=0;u<4;u++)
            #pragma unroll
            for (int i=0;i<8;i++) { u64 aa=pack(a[i],a[i]);
                #pragma unroll
                for (int j=0;j<4;j++) fma2(acc[i][j], aa, b2[j]); }
        }
        for(int i=0;i<8;i++) for(int j=0;j<4;j++) s+=fold(acc[i][j]);
    } else {
        u64 acc[8][4], b2[4], a2[8];
        for (int j=0;j<4;j++) b2[j]=pack(b[2*j],b[2*j+1]);
        for (int i=0;i<8;i++) a2[i]=pack(a[i],a[(i+1)&7]);
        for (int i=0;i<8;i++) for(int j=0;j<4;j++) acc[i][j]=pack((float)i,(float)j);
        #pragma unroll 1
        for (int it=0; it<ITERS; ++it) {
            #pragma unroll
            for (int u=0;u<4;u++)
            #pragma unroll
            for (int i=0;i<8;i++)
                #pragma unroll
                for (int j=0;j<4;j++) fma2(acc[i][j], a2[i], b2[j]);
        }
        for(int i=0;i<8;i++) for(int j=0;j<4;j++) s+=fold(acc[i][j]);
    }
    out[blockIdx.x*blockDim.x+threadIdx.x]=s;
}
template<typename F> float timeit(F f, int reps){
    cudaEvent_t e0,e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    f(); f(); CK(cudaDeviceSynchronize()); float best=1e30f;
    for(int r=0;r<reps;r++){ CK(cudaEventRecord(e0)); f(); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); float ms; CK(cudaEventElapsedTime(&ms,e0,e1)); if(ms<best)best=ms;}
    return best;
}
int main(){
    CK(cudaSetDevice(0)); cudaDeviceProp p; CK(cudaGetDeviceProperties(&p,0));
    int nsm=p.multiProcessorCount; int n=1024; float *in,*out; CK(cudaMalloc(&in,n*4)); CK(cudaMalloc(&out,(size_t)nsm*8*256*4*4)); CK(cudaMemset(in,0,n*4));
    constexpr int IT=4096;
    for (int occ=1; occ<=4; occ*=2) {
        int grid=nsm*occ*4; double flops=(double)grid*256*IT*4*64*2;
        // limit residency via dynamic smem: occ CTAs per SM
        size_t smem = occ==1? 120*1024 : (occ==2? 100*1024 : 50*1024);
        CK(cudaFuncSetAttribute(kern<0,IT>,cudaFuncAttributeMaxDynamicSharedMemorySize,(int)smem));
        CK(cudaFuncSetAttribute(kern<1,IT>,cudaFuncAttributeMaxDynamicSharedMemorySize,(int)smem));
        CK(cudaFuncSetAttribute(kern<2,IT>,cudaFuncAttributeMaxDynamicSharedMemorySize,(int)smem));
        float t0=timeit([&]{kern<0,IT><<<grid,256,smem>>>(out,in,n);},5);
        float t1=timeit([&]{kern<1,IT><<<grid,256,smem>>>(out,in,n);},5);
        float t2=timeit([&]{kern<2,IT><<<grid,256,smem>>>(out,in,n);},5);
        printf("CTAs/SM<=%d (256 thr): scalar %.2f TF | packed-bcast %.2f TF | packed-pairs %.2f TF\n", occ>2?2:occ, flops/t0*1e-9, flops/t1*1e-9, flops/t2*1e-9);
    }
    CK(cudaGetLastError()); return 0;
}

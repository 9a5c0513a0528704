// FP32 peak microbenchmark v3: packed-FMA throughput vs register-tile shape / issue order / scalar mix.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("CUDA error %s at %d\n",cudaGetErrorString(e),__LINE__); exit(1);} }while(0)
typedef unsigned long long u64;
__device__ __forceinline__ void fma2(u64& d, u64 a, u64 b){ asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(d) : "l"(a), "l"(b)); }
__device__ __forceinline__ void fma1(float& d, float a, float b){ asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(d) : "f"(a), "f"(b)); }
__device__ __forceinline__ u64 pack(float x,float y){ u64 r; asm("mov.b64 %0, {%1,%2};":"=l"(r):"f"(x),"f"(y)); return r;}
__device__ __forceinline__ float fold(u64 v){ float x,y; asm("mov.b64 {%0,%1}, %2;":"=f"(x),"=f"(y):"l"(v)); return x+y; }

// TI scalar A values (broadcast), TJ packed B pairs; ORDER 0: i outer, j inner; 1: j outer, i inner.
// SCAL: extra scalar FFMA per SCAL packed ones (0 = none)
template<int TI,int TJ,int ORDER,int SCAL,int ITERS,int MINB>
__global__ void __launch_bounds__(256,MINB) kern(float* out, const float* in) {
    float a[TI]; u64 b2[TJ]; u64 acc[TI][TJ]; float sacc[4]={0,1,2,3};
    for (int i=0;i<TI;i++) a[i]=in[(threadIdx.x+i)&1023];
    for (int j=0;j<TJ;j++) b2[j]=pack(in[(threadIdx.x*3+2*j)&1023],in[(threadIdx.x*3+2*j+1)&1023]);
    for (int i=0;i<TI;i++) for(int j=0;j<TJ;j++) acc[i][j]=pack((float)i,(float)j);
    #pragma unroll 1
    for (int it=0; it<ITERS; ++it) {
        #pragma unroll
        for (int u=0;u<4;u++) {
            int cnt=0;
            if (ORDER==0) {
                #pragma unroll
                for (int i=0;i<TI;i++){ u64 aa=pack(a[i],a[i]);
                    #pragma unroll
                    for (int j=0;j<TJ;j++){ fma2(acc[i][j],aa,b2[j]); if (SCAL && (++cnt%SCAL)==0) fma1(sacc[(cnt/SCAL)&3],a[i],a[(i+1)%TI]); } }
            } else {
                #pragma unroll
                for (int j=0;j<TJ;j++)
                    #pragma unroll
                    for (int i=0;i<TI;i++){ u64 aa=pack(a[i],a[i]); fma2(acc[i][j],aa,b2[j]); if (SCAL && (++cnt%SCAL)==0) fma1(sacc[(cnt/SCAL)&3],a[i],a[(i+1)%TI]); }
            }
        }
    }
    float s=sacc[0]+sacc[1]+sacc[2]+sacc[3];
    for(int i=0;i<TI;i++) for(int j=0;j<TJ;j++) s+=fold(acc[i][j]);
    out[blockIdx.x*blockDim.x+threadIdx.x]=s;
}
template<typename K> double run(K k, int grid, double flops_per_thread, float* out, const float* in){
    cudaEvent_t e0,e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    k<<<grid,256>>>(out,in); k<<<grid,256>>>(out,in); CK(cudaDeviceSynchronize()); float best=1e30f;
    for(int r=0;r<4;r++){ CK(cudaEventRecord(e0)); k<<<grid,256>>>(out,in); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); float ms; CK(cudaEventElapsedTime(&ms,e0,e1)); if(ms<best)best=ms;}
    return flops_per_thread*grid*256/(best*1e-3)*1e-12;
}
#define RUN(TI,TJ,ORD,SC,MB) { double fl=(double)IT*4*(TI*TJ*4.0 + (SC? (TI*TJ/SC)*2.0:0)); \
   printf("TI=%2d TJ=%d order=%d scalar-mix=1/%d minBlocks=%d : %.2f TFLOP/s\n",TI,TJ,ORD,SC,MB, run(kern<TI,TJ,ORD,SC,IT,MB>, nsm*MB*4, fl, out, in)); }
int main(){
    CK(cudaSetDevice(0)); cudaDeviceProp p; CK(cudaGetDeviceProperties(&p,0)); int nsm=p.multiProcessorCount;
    float *in,*out; CK(cudaMalloc(&in,4096)); CK(cudaMalloc(&out,(size_t)nsm*16*256*4)); CK(cudaMemset(in,0,4096));
    constexpr int IT=2048;
    RUN(8,4,0,0,2) RUN(8,4,1,0,2) RUN(4,4,0,0,2) RUN(4,8,0,0,2) RUN(16,2,0,0,2) RUN(16,2,1,0,2) RUN(4,4,0,0,4) RUN(2,4,0,0,4)
    RUN(8,4,0,8,2) RUN(8,4,0,4,2) RUN(8,4,0,2,2) RUN(8,4,0,16,2) RUN(8,8,0,0,1) RUN(16,4,0,0,1) RUN(8,4,0,0,1)
    CK(cudaGetLastError()); return 0;
}

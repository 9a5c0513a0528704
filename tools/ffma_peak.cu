// FP32 FFMA peak microbenchmark for sm_100a: measures the roofline denominator for the
// fused distance+argmin kernel (MEASURED_PEAKS.json has no FP32 figure).
// Variants: scalar FFMA 8x8 outer product; packed fma.rn.f32x2 8x8 outer product; pure chains.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("CUDA error %s at %d\n",cudaGetErrorString(e),__LINE__); exit(1);} }while(0)

template<int ITERS>
__global__ void __launch_bounds__(256,2) k_scalar(float* out, const float* in, int n) {
    float a[8], b[8], acc[8][8];
    for (int i=0;i<8;i++){ a[i]=in[(threadIdx.x+i)%n]; b[i]=in[(threadIdx.x*3+i)%n]; }
    for (int i=0;i<8;i++) for(int j=0;j<8;j++) acc[i][j]=0.f;
    for (int it=0; it<ITERS; ++it) {
        #pragma unroll
        for (int u=0;u<8;u++) {
            #pragma unroll
            for (int i=0;i<8;i++)
                #pragma unroll
                for (int j=0;j<8;j++) acc[i][j]=fmaf(a[i],b[j],acc[i][j]);
            // rotate operands a little so the compiler cannot hoist
            #pragma unroll
            for (int i=0;i<8;i++){ a[i]+=1e-9f; }
        }
    }
    float s=0; for(int i=0;i<8;i++) for(int j=0;j<8;j++) s+=acc[i][j];
    out[blockIdx.x*blockDim.x+threadIdx.x]=s;
}

__device__ __forceinline__ void fma2(unsigned long long& d, unsigned long long a, unsigned long long b){
    asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(d) : "l"(a), "l"(b));
}
__device__ __forceinline__ unsigned long long pack(float x,float y){ unsigned long long r; asm("mov.b64 %0, {%1,%2};":"=l"(r):"f"(x),"f"(y)); return r;}

template<int ITERS>
__global__ void __launch_bounds__(256,2) k_packed(float* out, const float* in, int n) {
    float a[8]; unsigned long long b2[4], acc[8][4];
    for (int i=0;i<8;i++){ a[i]=in[(threadIdx.x+i)%n]; }
    for (int j=0;j<4;j++) b2[j]=pack(in[(threadIdx.x*3+2*j)%n], in[(threadIdx.x*3+2*j+1)%n]);
    for (int i=0;i<8;i++) for(int j=0;j<4;j++) acc[i][j]=0ull;
    for (int it=0; it<ITERS; ++it) {
        #pragma unroll
        for (int u=0;u<8;u++) {
            #pragma unroll
            for (int i=0;i<8;i++) {
                unsigned long long aa = pack(a[i],a[i]);
                #pragma unroll
                for (int j=0;j<4;j++) fma2(acc[i][j], aa, b2[j]);
            }
            #pragma unroll
            for (int i=0;i<8;i++){ a[i]+=1e-9f; }
        }
    }
    float s=0; for(int i=0;i<8;i++) for(int j=0;j<4;j++){ float x,y; asm("mov.b64 {%0,%1}, %2;":"=f"(x),"=f"(y):"l"(acc[i][j])); s+=x+y; }
    out[blockIdx.x*blockDim.x+threadIdx.x]=s;
}

template<typename F> float timeit(F f, int reps){
    cudaEvent_t e0,e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    f(); f(); CK(cudaDeviceSynchronize());
    float best=1e30f, tot=0;
    for(int r=0;r<reps;r++){ CK(cudaEventRecord(e0)); f(); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); float ms; CK(cudaEventElapsedTime(&ms,e0,e1)); if(ms<best)best=ms; tot+=ms;}
    printf("   best %.3f ms  mean %.3f ms\n",best,tot/reps);
    return best;
}

int main(int argc,char**argv){
    int dev=0; CK(cudaSetDevice(dev)); cudaDeviceProp p; CK(cudaGetDeviceProperties(&p,dev));
    printf("device %s SMs %d clock %d kHz\n",p.name,p.multiProcessorCount,p.clockRate);
    int nsm=p.multiProcessorCount; int grid=nsm*2*8; int n=1024;
    float *in,*out; CK(cudaMalloc(&in,n*4)); CK(cudaMalloc(&out,(size_t)grid*256*4)); CK(cudaMemset(in,0,n*4));
    constexpr int IT=2048;
    double flops=(double)grid*256*IT*8*64*2;
    printf("scalar FFMA 8x8:\n"); float t1=timeit([&]{k_scalar<IT><<<grid,256>>>(out,in,n);},10);
    printf("   => %.2f TFLOP/s (best)\n",flops/t1*1e-9);
    printf("packed FFMA2 8x(4x2):\n"); float t2=timeit([&]{k_packed<IT><<<grid,256>>>(out,in,n);},10);
    printf("   => %.2f TFLOP/s (best)\n",flops/t2*1e-9);
    // sustained: run ~3 s back-to-back
    for (int v=0; v<2; ++v){
        cudaEvent_t e0,e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
        int reps = (int)(3000.0f/(v==0?t1:t2))+1; CK(cudaEventRecord(e0));
        for(int r=0;r<reps;r++){ if(v==0) k_scalar<IT><<<grid,256>>>(out,in,n); else k_packed<IT><<<grid,256>>>(out,in,n);} 
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); float ms; CK(cudaEventElapsedTime(&ms,e0,e1));
        printf("sustained %s: %.2f TFLOP/s over %.1f s\n", v==0?"scalar":"packed", flops*reps/ms*1e-9, ms*1e-3);
    }
    CK(cudaGetLastError());
    return 0;
}

// peak.cu — on-device FP32 FMA peak measurement with kernel 2's instruction mix (roofline denominator).
// MEASURED_PEAKS.json carries HBM and bf16 tensor figures only; the fused distance kernel is bound by the
// FP32 pipe, so its peak is measured here: an 8x8 outer-product of register operands, packed (FFMA2) and scalar.
#include "common.cuh"
#include "internal.h"

namespace f2p {

template <int PACKED, int ITERS>
__global__ void __launch_bounds__(256, 2) fma_peak_kernel(float* out, const float* in) {
    float a[8], b[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { a[i] = in[(threadIdx.x + i) & 1023]; b[i] = in[(threadIdx.x * 3 + i) & 1023]; }
    float s = 0.f;
    if (PACKED) {
        u64 acc[8][4], b2[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) b2[j] = pack2(b[2 * j], b[2 * j + 1]);
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = pack2((float)i, (float)j);
#pragma unroll 1
        for (int it = 0; it < ITERS; ++it) {
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int j = 0; j < 4; ++j)   // packed operand outer: the order that reaches the nominal rate
#pragma unroll
                    for (int i = 0; i < 8; ++i)
                        asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(acc[i][j]) : "l"(pack2(a[i], a[i])), "l"(b2[j]));
        }
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) { float x, y; unpack2(acc[i][j], x, y); s += x + y; }
    } else {
        float acc[8][8];
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[i][j] = (float)(i + j);
#pragma unroll 1
        for (int it = 0; it < ITERS; ++it) {
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int i = 0; i < 8; ++i)
#pragma unroll
                    for (int j = 0; j < 8; ++j) asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(acc[i][j]) : "f"(a[i]), "f"(b[j]));
        }
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 8; ++j) s += acc[i][j];
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int measure_fp32_peak(double* packed, double* scalar, cudaStream_t st) {
    int dev = 0, sm = 0;
    F2P_CUDA(cudaGetDevice(&dev));
    F2P_CUDA(cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, dev));
    constexpr int IT = 2048;
    const int grid = sm * 2 * 4;
    float *in = nullptr, *out = nullptr;
    F2P_CUDA(cudaMalloc(&in, 1024 * sizeof(float)));
    F2P_CUDA(cudaMalloc(&out, (size_t)grid * 256 * sizeof(float)));
    F2P_CUDA(cudaMemsetAsync(in, 0, 1024 * sizeof(float), st));
    cudaEvent_t e0, e1;
    F2P_CUDA(cudaEventCreate(&e0)); F2P_CUDA(cudaEventCreate(&e1));
    const double flops = (double)grid * 256 * IT * 4 * 64 * 2;
    double res[2] = {0, 0};
    for (int v = 0; v < 2; ++v) {
        float best = 1e30f;
        for (int rep = 0; rep < 6; ++rep) {
            F2P_CUDA(cudaEventRecord(e0, st));
            if (v == 0) fma_peak_kernel<1, IT><<<grid, 256, 0, st>>>(out, in);
            else fma_peak_kernel<0, IT><<<grid, 256, 0, st>>>(out, in);
            F2P_CUDA(cudaEventRecord(e1, st));
            F2P_CUDA(cudaEventSynchronize(e1));
            float ms = 0.f;
            F2P_CUDA(cudaEventElapsedTime(&ms, e0, e1));
            if (rep >= 2 && ms < best) best = ms;
        }
        res[v] = flops / (best * 1e-3) * 1e-12;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(in); cudaFree(out);
    if (packed) *packed = res[0];
    if (scalar) *scalar = res[1];
    return 0;
}

}  // namespace f2p

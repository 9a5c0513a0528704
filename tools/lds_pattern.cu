// Shared-memory LDS.128 throughput for the lane->address patterns a register-tiled kernel can use.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/lds_pattern tools/lds_pattern.cu
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(const int* __restrict__ lane_off, float* out, int iters, long long* cyc) {
    __shared__ __align__(16) float s[8192];
    for (int i = threadIdx.x; i < 8192; i += blockDim.x) s[i] = (float)i;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const float* p = s + lane_off[lane];            // float offset, multiple of 4
    float4 acc = make_float4(0, 0, 0, 0);
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            float4 v;
            asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                         : "r"((unsigned)__cvta_generic_to_shared(p + ((u * 512 + it * 128) & 4095))));
            acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
    }
    long long t1 = clock64();
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc.x + acc.y + acc.z + acc.w;
}
int main() {
    struct Pat { const char* name; int off[32]; } pats[16];
    int np = 0;
    auto add = [&](const char* n, auto f) { pats[np].name = n; for (int l = 0; l < 32; ++l) pats[np].off[l] = 4 * f(l); ++np; };
    add("32 distinct (512B contiguous)", [](int l) { return l; });
    add("b now: 8 distinct, lane&7", [](int l) { return l & 7; });
    add("a now: 4 distinct, lane>>3", [](int l) { return l >> 3; });
    add("b 8x4: 4 distinct, lane&3", [](int l) { return l & 3; });
    add("a 8x4: 8 distinct, lane>>2", [](int l) { return l >> 2; });
    add("b 16x2: 2 distinct, lane&1", [](int l) { return l & 1; });
    add("a 16x2: 16 distinct, lane>>1", [](int l) { return l >> 1; });
    add("b 2x16: 16 distinct, lane&15", [](int l) { return l & 15; });
    add("a 2x16: 2 distinct, lane>>4", [](int l) { return l >> 4; });
    add("1 distinct", [](int l) { return 0; });
    add("a half-warp 2x8: (l>>4)*2+((l>>3)&1)", [](int l) { return ((l >> 4) << 1) | ((l >> 3) & 1); });
    int* d_off; float* d_out; long long* d_c;
    cudaMalloc(&d_off, 128); cudaMalloc(&d_out, 148 * 8 * 256 * 4); cudaMalloc(&d_c, 8);
    for (int i = 0; i < np; ++i) {
        cudaMemcpy(d_off, pats[i].off, 128, cudaMemcpyHostToDevice);
        const int iters = 2000;
        k<<<148, 256>>>(d_off, d_out, iters, d_c);
        k<<<148, 256>>>(d_off, d_out, iters, d_c);
        cudaDeviceSynchronize();
        long long c; cudaMemcpy(&c, d_c, 8, cudaMemcpyDeviceToHost);
        // 8 warps per SM (1 CTA/SM), each issues iters*16 LDS.128
        printf("%-42s %6.2f SM-cycles per warp-LDS.128\n", pats[i].name, (double)c / (iters * 16.0 * 8.0));
    }
    return 0;
}

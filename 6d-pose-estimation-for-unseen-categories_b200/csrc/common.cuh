// common.cuh — shared device/host helpers for libf2p_sm100.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/f2p.h"

namespace f2p {

// ---------------------------------------------------------------------------------------------------
// Tile geometry of the k-major operand layout shared by kernels 1 and 2.
// Every pair's point axis is padded to a multiple of TILE so a TMA box never straddles two pairs.
// ---------------------------------------------------------------------------------------------------
constexpr int TILE = 128;      // rows (CAD vertices) and columns (scene points) per CTA tile
constexpr int KC = 16;         // k-rows per TMA stage of the streamed operand
constexpr int NSTAGE = 4;      // TMA pipeline depth
constexpr int NN_THREADS = 256;
constexpr unsigned F2P_ROW0_ONLY = 0x80000000u;  // internal flag: single-kernel entry points write indices only

static inline int round_up(int v, int m) { return (v + m - 1) / m * m; }

// Per-pair record the kernels read (built on the host from F2PPairDesc, uploaded once per call).
struct PairDev {
    int64_t ex_off, ey_off, c_off, cad_off, pc_off, out_off;  // element offsets into the batch base pointers
    int n1, n2, ldx, ldy, ncad, npc;
    int x_col0, y_col0;      // first column of this pair in the concatenated xT / yT operands (multiples of TILE)
    int nrt, nct;            // row tiles (over n1) and column tiles (over n2)
    int nsplit;              // row splits per column tile (work items per column tile)
    int item0;               // first NN work item of this pair (prefix sum of nct*nsplit)
    int xtile0, ytile0;      // prefix sums of nrt / nct (prep-kernel work items)
    int cand0;               // first candidate slot of this pair (prefix sum of 5*n2), spatial only
    float thr[4];
    // tensor-pipe screen (tc.cu): CAD rows re-packed K-major in 256-row tiles
    int tc_nrt;              // 256-row tiles over n1
    int tc_nsplit;           // row splits per column tile
    int tc_item0;            // first screen work item of this pair (prefix sum of nct*tc_nsplit)
    int tc_xtile0;           // prefix sum of 2*tc_nrt (128-row packing tiles)
    int xa_row0;             // first row of this pair in the packed XA operand (multiple of 256)
};

// thread-local error text (f2p_last_error)
void set_error(const char* fmt, ...);
const char* get_error();

#define F2P_CUDA(call)                                                                       \
    do {                                                                                     \
        cudaError_t e__ = (call);                                                            \
        if (e__ != cudaSuccess) {                                                            \
            f2p::set_error("%s failed at %s:%d: %s", #call, __FILE__, __LINE__, cudaGetErrorString(e__)); \
            return (int)e__;                                                                 \
        }                                                                                    \
    } while (0)

#ifdef __CUDACC__
// ---------------------------------------------------------------------------------------------------
// PTX wrappers: mbarrier, TMA (cp.async.bulk.tensor), packed FP32 (fma/add .f32x2 — sm_100 only).
// ---------------------------------------------------------------------------------------------------
typedef unsigned long long u64;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// one non-blocking probe of a phase: true when the phase with the given parity has completed
__device__ __forceinline__ bool mbar_test(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void prefetch_l1(const void* p) {
    asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
}
// 2-D tiled TMA load: box at (c0 = inner coordinate, c1 = outer coordinate) -> smem, completes on `bar`.
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}

__device__ __forceinline__ u64 pack2(float lo, float hi) {
    u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r;
}
__device__ __forceinline__ void unpack2(u64 v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
// d = a*b + d on two FP32 lanes, each IEEE round-to-nearest fused (bit-identical to scalar fmaf per lane)
__device__ __forceinline__ void fma2(u64& d, u64 a, u64 b) {
    asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(d) : "l"(a), "l"(b));
}
// same, volatile: keeps the source issue order (operand-reuse pattern) through nvcc's scheduler
__device__ __forceinline__ void fma2v(u64& d, u64 a, u64 b) {
    asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(d) : "l"(a), "l"(b));
}
__device__ __forceinline__ u64 add2(u64 a, u64 b) {
    u64 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r;
}

// Maps a flat work-item id to its pair: uniform batches divide, ragged batches binary-search the prefix.
__device__ __forceinline__ int find_pair(int item, const PairDev* __restrict__ pairs, int npairs, int uniform_items,
                                         int PairDev::*prefix) {
    if (uniform_items > 0) return item / uniform_items;
    int lo = 0, hi = npairs - 1;
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (pairs[mid].*prefix <= item) lo = mid; else hi = mid - 1;
    }
    return lo;
}
#endif  // __CUDACC__

// Host-side: encode a 2-D FP32 tensor map {inner = cols, outer = rows} with box {box_cols, box_rows}.
// swizzle128 != 0: CU_TENSOR_MAP_SWIZZLE_128B (box_cols * 4 must be 128), the layout tcgen05 descriptors expect.
int make_tensor_map_2d(CUtensorMap* map, const float* base, uint64_t cols, uint64_t rows, uint64_t row_stride_elems,
                       uint32_t box_cols, uint32_t box_rows, int swizzle128 = 0);

}  // namespace f2p

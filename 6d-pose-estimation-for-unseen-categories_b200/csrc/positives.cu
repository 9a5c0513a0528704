// positives.cu — ground-truth correspondence list of a sample: all (CAD vertex, scene point)
// pairs closer than r under the ground-truth pose.
//
// Replaces `find_positives` (dataset/object.py:281-288: np.linalg.norm(pc1[:,None]-pc2, axis=2) <= r, np.argwhere) and
// `get_overlap` (:311-317) without the [n1,n2,3] and [n1,n2] temporaries.  Same all-pairs access pattern as kernel 2,
// but 3-D points, a radius test instead of an arg-min, and an ORDERED variable-length result.
// Arithmetic contract (NumPy, in the arrays' own dtype T = float64 or float32; pinned by oracle/f2p_oracle.py):
//   d_c = fl(a_c - b_c); q_c = fl(d_c*d_c); s = fl(fl(q0+q1)+q2) (add.reduce over 3 items is sequential);
//   dist = sqrt_rn(s); positive iff dist <= T(r).  Pairs are listed in row-major (i, then j) order as int64 [N,2].
// Two passes keep the order deterministic: per-row counts -> exclusive scan -> ordered write (warp = one row,
// ballot + popc ranks inside a 32-column step).
#include <string.h>
#include "common.cuh"
#include "internal.h"

namespace f2p {

template <typename T> struct Pt { T x, y, z; };

template <typename T> __device__ __forceinline__ bool within(const Pt<T> a, const T* __restrict__ b, T r);
template <> __device__ __forceinline__ bool within<double>(const Pt<double> a, const double* __restrict__ b, double r) {
    const double dx = __dsub_rn(a.x, b[0]), dy = __dsub_rn(a.y, b[1]), dz = __dsub_rn(a.z, b[2]);
    const double s = __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
    return __dsqrt_rn(s) <= r;
}
template <> __device__ __forceinline__ bool within<float>(const Pt<float> a, const float* __restrict__ b, float r) {
    const float dx = __fsub_rn(a.x, b[0]), dy = __fsub_rn(a.y, b[1]), dz = __fsub_rn(a.z, b[2]);
    const float s = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
    return __fsqrt_rn(s) <= r;
}

constexpr int PW = 8;   // warps (= rows) per CTA

// WRITE = false: rowcnt[i] = positives of row i, overlap flags.  WRITE = true: pairs at rowoff[i] in column order.
template <typename T, bool WRITE>
__global__ void __launch_bounds__(PW * 32)
positives_kernel(const T* __restrict__ pc1, int n1, const T* __restrict__ pc2, int n2, T r, int* __restrict__ rowcnt,
                 const int64_t* __restrict__ rowoff, int64_t* __restrict__ pairs, int64_t cap,
                 int8_t* __restrict__ ov12, int8_t* __restrict__ ov21) {
    const int lane = threadIdx.x & 31;
    const int i = blockIdx.x * PW + (threadIdx.x >> 5);
    if (i >= n1) return;
    const Pt<T> a = {pc1[3 * (size_t)i], pc1[3 * (size_t)i + 1], pc1[3 * (size_t)i + 2]};
    int64_t pos = WRITE ? rowoff[i] : 0;
    int cnt = 0;
    for (int j0 = 0; j0 < n2; j0 += 32) {
        const int j = j0 + lane;
        const bool hit = j < n2 && within<T>(a, pc2 + 3 * (size_t)j, r);
        const unsigned bal = __ballot_sync(0xffffffffu, hit);
        if (WRITE) {
            if (hit) {
                const int64_t o = pos + __popc(bal & ((1u << lane) - 1));
                if (o < cap) { pairs[2 * o] = i; pairs[2 * o + 1] = j; }
            }
            pos += __popc(bal);
        } else {
            if (hit && ov21) ov21[j] = 1;
            cnt += __popc(bal);
        }
    }
    if (!WRITE && lane == 0) { rowcnt[i] = cnt; if (ov12) ov12[i] = cnt > 0 ? 1 : 0; }
}

// exclusive scan of the row counts (single CTA; n1 is at most a few 100k)
__global__ void __launch_bounds__(1024)
positives_scan_kernel(const int* __restrict__ rowcnt, int n1, int64_t* __restrict__ rowoff, int64_t* __restrict__ total) {
    __shared__ int64_t s_w[32];
    __shared__ int64_t s_run;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_run = 0;
    __syncthreads();
    for (int base = 0; base < n1; base += 1024) {
        const int i = base + tid;
        const int64_t v = i < n1 ? rowcnt[i] : 0;
        int64_t x = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const int64_t y = __shfl_up_sync(0xffffffffu, x, d); if (lane >= d) x += y; }
        if (lane == 31) s_w[warp] = x;
        __syncthreads();
        int64_t woff = 0;
        for (int w = 0; w < warp; ++w) woff += s_w[w];
        const int64_t run = s_run;
        if (i < n1) rowoff[i] = run + woff + x - v;
        __syncthreads();
        if (tid == 1023) s_run = run + woff + x;
        __syncthreads();
    }
    if (tid == 0) *total = s_run;
}

int launch_rowscan(const int* rowcnt, int n, int64_t* rowoff, int64_t* total, cudaStream_t st) {
    positives_scan_kernel<<<1, 1024, 0, st>>>(rowcnt, n, rowoff, total);
    F2P_CUDA(cudaGetLastError());
    return 0;
}

template <typename T>
static int run_count(const void* pc1, int n1, const void* pc2, int n2, double r, int64_t* total, int8_t* ov12, int8_t* ov21,
                     unsigned char* ws, cudaStream_t st) {
    int* rowcnt = reinterpret_cast<int*>(ws);
    int64_t* rowoff = reinterpret_cast<int64_t*>(ws + (((size_t)n1 * sizeof(int) + 255) & ~(size_t)255));
    if (ov21) F2P_CUDA(cudaMemsetAsync(ov21, 0, (size_t)n2, st));
    if (n1 > 0)
        positives_kernel<T, false><<<(n1 + PW - 1) / PW, PW * 32, 0, st>>>(static_cast<const T*>(pc1), n1, static_cast<const T*>(pc2),
                                                                         n2, (T)r, rowcnt, nullptr, nullptr, 0, ov12, ov21);
    positives_scan_kernel<<<1, 1024, 0, st>>>(rowcnt, n1, rowoff, total);
    F2P_CUDA(cudaGetLastError());
    return 0;
}

template <typename T>
static int run_write(const void* pc1, int n1, const void* pc2, int n2, double r, int64_t* pairs, int64_t cap,
                     unsigned char* ws, cudaStream_t st) {
    const int64_t* rowoff = reinterpret_cast<const int64_t*>(ws + (((size_t)n1 * sizeof(int) + 255) & ~(size_t)255));
    if (n1 > 0)
        positives_kernel<T, true><<<(n1 + PW - 1) / PW, PW * 32, 0, st>>>(static_cast<const T*>(pc1), n1, static_cast<const T*>(pc2),
                                                                        n2, (T)r, nullptr, rowoff, pairs, cap, nullptr, nullptr);
    F2P_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace f2p

using namespace f2p;

extern "C" {

size_t f2p_positives_workspace_bytes(int n1) {
    const size_t n = (size_t)(n1 > 0 ? n1 : 1);
    return ((n * sizeof(int) + 255) & ~(size_t)255) + ((n * sizeof(int64_t) + 255) & ~(size_t)255);
}

int f2p_find_positives_count(const void* pc1, int n1, const void* pc2, int n2, int is_f64, double r, int64_t* total_dev,
                             int8_t* ov12, int8_t* ov21, void* ws, size_t ws_bytes, void* stream) {
    if ((!pc1 && n1 > 0) || (!pc2 && n2 > 0) || !total_dev || !ws) { set_error("null pointer"); return F2P_EINVAL; }
    if (n1 < 0 || n2 < 0) { set_error("negative size"); return F2P_EINVAL; }
    if (ws_bytes < f2p_positives_workspace_bytes(n1)) { set_error("workspace too small"); return F2P_ENOSPACE; }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    return is_f64 ? run_count<double>(pc1, n1, pc2, n2, r, total_dev, ov12, ov21, static_cast<unsigned char*>(ws), st)
                  : run_count<float>(pc1, n1, pc2, n2, r, total_dev, ov12, ov21, static_cast<unsigned char*>(ws), st);
}

int f2p_find_positives_write(const void* pc1, int n1, const void* pc2, int n2, int is_f64, double r, int64_t* pairs_dev,
                             int64_t capacity, void* ws, size_t ws_bytes, void* stream) {
    if ((!pc1 && n1 > 0) || (!pc2 && n2 > 0) || (!pairs_dev && capacity > 0) || !ws) { set_error("null pointer"); return F2P_EINVAL; }
    if (n1 < 0 || n2 < 0 || capacity < 0) { set_error("negative size"); return F2P_EINVAL; }
    if (ws_bytes < f2p_positives_workspace_bytes(n1)) { set_error("workspace too small"); return F2P_ENOSPACE; }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    return is_f64 ? run_write<double>(pc1, n1, pc2, n2, r, pairs_dev, capacity, static_cast<unsigned char*>(ws), st)
                  : run_write<float>(pc1, n1, pc2, n2, r, pairs_dev, capacity, static_cast<unsigned char*>(ws), st);
}

}  // extern "C"

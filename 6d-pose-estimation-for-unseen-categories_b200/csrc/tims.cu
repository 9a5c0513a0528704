// tims.cu — pairwise-consistency front half of the TEASER++ solve that consumes the point map.
//
// Replaces, for the configuration scripts/test_teaser.py:363-370 uses (estimate_scaling = False, cbar2 = 1):
//   RobustRegistrationSolver::computeTIMs   TEASER-plusplus/teaser/src/registration.cc:481-519  (src and dst)
//   ScaleInliersSelector::solveForScale     registration.cc:427-445
//   the inlier-graph edge loop              registration.cc:582-590
// i.e. for every pair i < j of the N correspondences: the translation-invariant measurements s = src_j - src_i and
// d = dst_j - dst_i, the test | |s| - |d| | <= beta (beta = 2 * noise_bound * sqrt(cbar2)), and an edge (i, j) of the
// inlier graph when it holds.  The reference materialises both [3, N(N-1)/2] TIM matrices and a boolean mask before
// walking them serially; here the test is fused with the differences (same all-pairs pattern as the vote kernel, FP64)
// and only what the next step needs is written: the edge list in the reference's insertion order (TIM order: i
// ascending, then j), the vertex degrees, and — on request — the mask and the TIM matrices themselves.
//
// Arithmetic contract (FP64): t_c = fl(v_j[c] - v_i[c]); n = sqrt_rn(fl(fl(t_0^2 + t_1^2) + t_2^2));
// edge iff fl(|fl(n_src - n_dst)|) <= beta.
// Parity: the mask is pinned to the reference's own fixture for this configuration (test/teaser/data/registration_test/
// {objectIn,sceneIn,fixed_scale_inliers}.csv of registration-test.cc:300-362 -> tests/golden/teaser_golden.npz: 168
// points, 14028 TIMs, 2008 inliers, all reproduced).  Eigen (the reference's only dependency on this path) is not in the
// image, so the reference code itself cannot be compiled here, and the summation order of Eigen's 3-element
// colwise().sum() depends on its vectorisation flags ((t0^2+t1^2)+t2^2 with SSE2 packets, t0^2+(t1^2+t2^2) without).
// The restatement in oracle/f2p_oracle.py uses the packet order; an edge can differ from the other order only when the
// statistic is within one ulp of beta (the fixture's closest statistic is 3e-5 away, so it does not decide this).
#include <string.h>
#include "common.cuh"
#include "internal.h"

namespace f2p {

constexpr int TW = 8;   // warps (= rows i) per CTA

__device__ __forceinline__ double tim_norm(const double ax, const double ay, const double az, const double* __restrict__ b) {
    const double tx = __dsub_rn(b[0], ax), ty = __dsub_rn(b[1], ay), tz = __dsub_rn(b[2], az);
    return __dsqrt_rn(__dadd_rn(__dadd_rn(__dmul_rn(tx, tx), __dmul_rn(ty, ty)), __dmul_rn(tz, tz)));
}

// WRITE = false: rowcnt[i] = edges (i, j > i), degree, optional mask.  WRITE = true: edges at rowoff[i], j ascending.
template <bool WRITE>
__global__ void __launch_bounds__(TW * 32)
tim_graph_kernel(const double* __restrict__ src, const double* __restrict__ dst, int n, double beta,
                 int* __restrict__ rowcnt, const int64_t* __restrict__ rowoff, int32_t* __restrict__ edges, int64_t cap,
                 int32_t* __restrict__ degree, uint8_t* __restrict__ mask) {
    const int lane = threadIdx.x & 31;
    // rows are handed out from both ends (row i has n-1-i columns): CTA b takes rows b*TW/2.. and the mirrored ones
    const int w = blockIdx.x * TW + (threadIdx.x >> 5);
    const int i = (w & 1) ? n - 1 - (w >> 1) : (w >> 1);
    if (w >= n) return;
    const double sx = src[3 * (size_t)i], sy = src[3 * (size_t)i + 1], sz = src[3 * (size_t)i + 2];
    const double dx = dst[3 * (size_t)i], dy = dst[3 * (size_t)i + 1], dz = dst[3 * (size_t)i + 2];
    const int64_t seg = (int64_t)i * n - (int64_t)i * (i + 1) / 2;     // first TIM of row i (registration.cc:497)
    int64_t pos = WRITE ? rowoff[i] : 0;
    int cnt = 0;
    for (int j0 = i + 1; j0 < n; j0 += 32) {
        const int j = j0 + lane;
        bool hit = false;
        if (j < n) {
            const double ns = tim_norm(sx, sy, sz, src + 3 * (size_t)j), nd = tim_norm(dx, dy, dz, dst + 3 * (size_t)j);
            hit = fabs(__dsub_rn(ns, nd)) <= beta;
        }
        const unsigned bal = __ballot_sync(0xffffffffu, hit);
        if (WRITE) {
            if (hit) {
                const int64_t o = pos + __popc(bal & ((1u << lane) - 1));
                if (o < cap) { edges[2 * o] = i; edges[2 * o + 1] = j; }
            }
            pos += __popc(bal);
        } else {
            if (hit && degree) atomicAdd(degree + j, 1);
            if (mask && j < n) mask[seg + (j - i - 1)] = hit ? 1 : 0;
            cnt += __popc(bal);
        }
    }
    if (!WRITE && lane == 0) { rowcnt[i] = cnt; if (degree && cnt) atomicAdd(degree + i, cnt); }
}

// vtilde[:, seg(i) + (j-i-1)] = v_j - v_i, map = (i, j)   (registration.cc:481-519; column-major [3, T] / [2, T])
__global__ void __launch_bounds__(TW * 32)
tims_kernel(const double* __restrict__ v, int n, double* __restrict__ vtilde, int32_t* __restrict__ map) {
    const int lane = threadIdx.x & 31;
    const int i = blockIdx.x * TW + (threadIdx.x >> 5);
    if (i >= n - 1) return;
    const double ax = v[3 * (size_t)i], ay = v[3 * (size_t)i + 1], az = v[3 * (size_t)i + 2];
    const int64_t seg = (int64_t)i * n - (int64_t)i * (i + 1) / 2;
    for (int j = i + 1 + lane; j < n; j += 32) {
        const int64_t t = seg + (j - i - 1);
        const double* b = v + 3 * (size_t)j;
        vtilde[3 * t] = __dsub_rn(b[0], ax); vtilde[3 * t + 1] = __dsub_rn(b[1], ay); vtilde[3 * t + 2] = __dsub_rn(b[2], az);
        if (map) { map[2 * t] = i; map[2 * t + 1] = j; }
    }
}

}  // namespace f2p

using namespace f2p;

extern "C" {

size_t f2p_tim_workspace_bytes(int n) { return f2p_positives_workspace_bytes(n); }   // row counts + row offsets

int f2p_tim_graph_count(const double* src, const double* dst, int n, double beta, int64_t* total_dev, int32_t* degree,
                        uint8_t* mask, void* ws_v, size_t ws_bytes, void* stream) {
    if ((n > 0 && (!src || !dst)) || !total_dev || !ws_v) { set_error("null pointer"); return F2P_EINVAL; }
    if (n < 0 || !(beta >= 0.0)) { set_error("n=%d beta=%g", n, beta); return F2P_EINVAL; }
    if (ws_bytes < f2p_tim_workspace_bytes(n)) { set_error("workspace too small"); return F2P_ENOSPACE; }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    unsigned char* ws = static_cast<unsigned char*>(ws_v);
    int* rowcnt = reinterpret_cast<int*>(ws);
    int64_t* rowoff = reinterpret_cast<int64_t*>(ws + (((size_t)n * sizeof(int) + 255) & ~(size_t)255));
    if (degree && n > 0) F2P_CUDA(cudaMemsetAsync(degree, 0, sizeof(int32_t) * (size_t)n, st));
    if (n > 0)
        tim_graph_kernel<false><<<(n + TW - 1) / TW, TW * 32, 0, st>>>(src, dst, n, beta, rowcnt, nullptr, nullptr, 0, degree, mask);
    const int rc = launch_rowscan(rowcnt, n, rowoff, total_dev, st);
    if (rc) return rc;
    F2P_CUDA(cudaGetLastError());
    return 0;
}

int f2p_tim_graph_write(const double* src, const double* dst, int n, double beta, int32_t* edges, int64_t capacity,
                        void* ws_v, size_t ws_bytes, void* stream) {
    if ((n > 0 && (!src || !dst)) || (!edges && capacity > 0) || !ws_v) { set_error("null pointer"); return F2P_EINVAL; }
    if (n < 0 || capacity < 0 || !(beta >= 0.0)) { set_error("n=%d capacity=%lld beta=%g", n, (long long)capacity, beta); return F2P_EINVAL; }
    if (ws_bytes < f2p_tim_workspace_bytes(n)) { set_error("workspace too small"); return F2P_ENOSPACE; }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    unsigned char* ws = static_cast<unsigned char*>(ws_v);
    const int64_t* rowoff = reinterpret_cast<const int64_t*>(ws + (((size_t)n * sizeof(int) + 255) & ~(size_t)255));
    if (n > 0)
        tim_graph_kernel<true><<<(n + TW - 1) / TW, TW * 32, 0, st>>>(src, dst, n, beta, nullptr, rowoff, edges, capacity, nullptr, nullptr);
    F2P_CUDA(cudaGetLastError());
    return 0;
}

int f2p_compute_tims(const double* v, int n, double* vtilde, int32_t* map, void* stream) {
    if (n < 0) { set_error("n < 0"); return F2P_EINVAL; }
    if (n < 2) return 0;
    if (!v || !vtilde) { set_error("null pointer"); return F2P_EINVAL; }
    tims_kernel<<<(n - 1 + TW - 1) / TW, TW * 32, 0, static_cast<cudaStream_t>(stream)>>>(v, n, vtilde, map);
    F2P_CUDA(cudaGetLastError());
    return 0;
}

}  // extern "C"

// inlier.cu — inlier ratio of predicted correspondences under the ground-truth alignment, for a whole batch.
//
// Replaces `compute_inlier_ratio` (utils/utils.py:81-105), called once per sample right after the solver in
// scripts/eval.py:88: gather CAD[p0], PC_aligned[p1], dist = sqrt(sum((a-b)^2)), inliers = #(dist < thr),
// ratio = inliers / M.  Arithmetic contract (pinned by oracle/f2p_oracle.c: f2po_inlier_count):
//   d_c = fl(a_c - b_c); q_c = fl(d_c*d_c); s = fl(fl(q0+q1)+q2); dist = sqrt_rn(s); inlier iff dist < thr (FP32);
//   ratio = fl32(inliers) / fl32(M)  (torch: int64 tensor / int -> float32 true division).
// Index rows follow torch indexing: negative values wrap once (the spatial solver's row 1 carries the reference's
// int16 wrap); anything still out of range marks the pair invalid (inliers = -1, ratio = NaN).
#include <string.h>
#include <vector>
#include "common.cuh"
#include "internal.h"

namespace f2p {

struct InlierPair {
    const float* cad; const float* pc;   // [ncad,3], [npc,3] (PC under the ground-truth transform)
    const int64_t* map;                  // dense int64 [2,M] (row stride M)
    int ncad, npc;
    int M;                               // used when counts == nullptr
    float thr;
};

constexpr int IN_T = 256;
constexpr int IN_COLS = 2048;            // columns per CTA

// grid = (ceil(maxM / IN_COLS), npairs)
__global__ void __launch_bounds__(IN_T)
inlier_count_kernel(const InlierPair* __restrict__ ips, const int32_t* __restrict__ counts, int32_t* __restrict__ inliers) {
    const InlierPair& ip = ips[blockIdx.y];
    const int M = counts ? counts[blockIdx.y] : ip.M;
    const int c0 = blockIdx.x * IN_COLS;
    if (c0 >= M) return;
    int n = 0;
    bool bad = false;
    for (int c = c0 + threadIdx.x; c < min(M, c0 + IN_COLS); c += IN_T) {
        int64_t i = ip.map[c], j = ip.map[(int64_t)M + c];
        if (i < 0) i += ip.ncad;
        if (j < 0) j += ip.npc;
        if (i < 0 || i >= ip.ncad || j < 0 || j >= ip.npc) { bad = true; continue; }
        const float* a = ip.cad + 3 * i; const float* b = ip.pc + 3 * j;
        const float dx = __fsub_rn(a[0], b[0]), dy = __fsub_rn(a[1], b[1]), dz = __fsub_rn(a[2], b[2]);
        const float s = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
        n += (__fsqrt_rn(s) < ip.thr) ? 1 : 0;
    }
    n = __reduce_add_sync(0xffffffffu, n);
    const bool anybad = __any_sync(0xffffffffu, bad);
    if ((threadIdx.x & 31) == 0) {
        if (anybad) atomicOr(reinterpret_cast<int*>(inliers + blockIdx.y), (int)0x80000000);   // sign bit: invalid
        else if (n) atomicAdd(inliers + blockIdx.y, n);
    }
}

__global__ void inlier_ratio_kernel(const InlierPair* __restrict__ ips, int npairs, const int32_t* __restrict__ counts,
                                    int32_t* __restrict__ inliers, float* __restrict__ ratio) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npairs) return;
    const int M = counts ? counts[p] : ips[p].M;
    const int n = inliers[p];
    if (n < 0) { inliers[p] = -1; if (ratio) ratio[p] = __int_as_float(0x7fc00000); return; }
    if (ratio) ratio[p] = M > 0 ? __fdiv_rn((float)n, (float)M) : 0.f;   // utils/utils.py:95: empty -> 0
}

}  // namespace f2p

using namespace f2p;

extern "C" {

size_t f2p_inlier_workspace_bytes(int npairs) {
    return (sizeof(InlierPair) * (size_t)(npairs > 0 ? npairs : 1) + 255) & ~(size_t)255;
}

int f2p_inlier_ratio_batched(const float* cad, const float* pc_aligned, const int64_t* maps, const F2PPairDesc* d,
                             int npairs, int spatial, const int32_t* counts_dev, const float* thr_host,
                             int32_t* inliers_dev, float* ratio_dev, void* ws, size_t ws_bytes, void* stream) {
    if (npairs == 0) return 0;
    if (!cad || !pc_aligned || !maps || !d || !thr_host || !inliers_dev || !ws) { set_error("null pointer"); return F2P_EINVAL; }
    if (npairs < 0) { set_error("npairs < 0"); return F2P_EINVAL; }
    if (ws_bytes < f2p_inlier_workspace_bytes(npairs)) { set_error("workspace too small"); return F2P_ENOSPACE; }
    std::vector<InlierPair> ips((size_t)npairs);
    int maxM = 0;
    for (int i = 0; i < npairs; ++i) {
        if (d[i].ncad < 1 || d[i].npc < 1 || d[i].n2 < 0) { set_error("pair %d: ncad=%d npc=%d n2=%d", i, d[i].ncad, d[i].npc, d[i].n2); return F2P_EINVAL; }
        InlierPair& p = ips[i];
        p.cad = cad + d[i].cad_off; p.pc = pc_aligned + d[i].pc_off; p.map = maps + d[i].out_off;
        p.ncad = d[i].ncad; p.npc = d[i].npc; p.thr = thr_host[i];
        p.M = d[i].n2;                                   // naive map [2,n2]; spatial maps take M' from counts_dev
        const int cap = spatial ? 5 * d[i].n2 : d[i].n2;
        if (cap > maxM) maxM = cap;
    }
    if (spatial && !counts_dev) { set_error("spatial maps need counts_dev (M' per pair)"); return F2P_EINVAL; }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    F2P_CUDA(cudaMemcpyAsync(ws, ips.data(), sizeof(InlierPair) * (size_t)npairs, cudaMemcpyHostToDevice, st));
    F2P_CUDA(cudaMemsetAsync(inliers_dev, 0, sizeof(int32_t) * (size_t)npairs, st));
    const InlierPair* ipd = static_cast<const InlierPair*>(ws);
    for (int base = 0; maxM > 0 && base < npairs; base += 65535) {      // grid.y limit
        const int np = npairs - base < 65535 ? npairs - base : 65535;
        dim3 grid((unsigned)((maxM + IN_COLS - 1) / IN_COLS), (unsigned)np);
        inlier_count_kernel<<<grid, IN_T, 0, st>>>(ipd + base, spatial ? counts_dev + base : nullptr, inliers_dev + base);
    }
    inlier_ratio_kernel<<<(npairs + 127) / 128, 128, 0, st>>>(ipd, npairs, spatial ? counts_dev : nullptr, inliers_dev, ratio_dev);
    F2P_CUDA(cudaGetLastError());
    return 0;
}

}  // extern "C"

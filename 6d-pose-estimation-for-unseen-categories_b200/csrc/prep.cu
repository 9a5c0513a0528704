// prep.cu — kernel 1: spectral projection into the k-major operand layout with fused ATen-ordered row norms.
//
// Replaces `torch.matmul(evecs_x, C12.t())` (naive.py:20, spacial_filtering.py:19) and the
// `x.pow(2).sum(-1)` norms that torch.cdist's matmul path forms (naive.py:33 -> ATen _euclidean_dist).
//
// Arithmetic contract (pinned by oracle/f2p_oracle.c):
//   feat[i,a] = fma(e[i,k-1],C[a,k-1], ... fma(e[i,0],C[a,0], 0))   one accumulator, b ascending
//   xT[a][i]  = -2*feat[i,a]                                        exact
//   nrm[i]    = sum_a feat[i,a]^2 in ATen's 8-lane x ILP-4 order (squares rounded first)
#include "common.cuh"

namespace f2p {

constexpr int PREP_THREADS = 256;
constexpr int EPAD = TILE + 1;  // padded row of the transposed evecs tile: conflict-free transposed stores

// ATen `x.pow(2).sum(-1)` of one point whose k values sit at p[c * stride]: 8 lanes x ILP-4 partial sums (k < 8: four
// interleaved scalar chains), squares rounded first.  `scale` is an exact power of two (-0.5 recovers feat from -2*feat).
__device__ __forceinline__ float aten_sqnorm(const float* p, int64_t stride, int k, float scale) {
    if (k < 8) {
        // scalar_inner_sum: 4 interleaved chains, leftovers into chain 0
        float q[4] = {0.f, 0.f, 0.f, 0.f};
        const int s4 = k / 4;
        for (int rr = 0; rr < s4; ++rr)
#pragma unroll
            for (int l = 0; l < 4; ++l) { const float v = scale * p[(4 * rr + l) * stride]; q[l] = __fadd_rn(q[l], __fmul_rn(v, v)); }
        for (int t2 = 4 * s4; t2 < k; ++t2) { const float v = scale * p[t2 * stride]; q[0] = __fadd_rn(q[0], __fmul_rn(v, v)); }
        q[0] = __fadd_rn(q[0], q[1]); q[0] = __fadd_rn(q[0], q[2]); q[0] = __fadd_rn(q[0], q[3]);
        return q[0];
    }
    const int nv = k / 8, g = nv / 4;
    float P[4][8];
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int l = 0; l < 8; ++l) P[j][l] = 0.f;
    for (int rr = 0; rr < g; ++rr) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int l = 0; l < 8; ++l) {
                const float v = scale * p[(32 * rr + 8 * j + l) * stride];
                P[j][l] = __fadd_rn(P[j][l], __fmul_rn(v, v));
            }
    }
    for (int v8 = 4 * g; v8 < nv; ++v8)
#pragma unroll
        for (int l = 0; l < 8; ++l) {
            const float v = scale * p[(8 * v8 + l) * stride];
            P[0][l] = __fadd_rn(P[0][l], __fmul_rn(v, v));
        }
#pragma unroll
    for (int l = 0; l < 8; ++l) {
        P[0][l] = __fadd_rn(P[0][l], P[1][l]);
        P[0][l] = __fadd_rn(P[0][l], P[2][l]);
        P[0][l] = __fadd_rn(P[0][l], P[3][l]);
    }
    float f = 0.f;
    for (int t2 = 8 * nv; t2 < k; ++t2) { const float v = scale * p[t2 * stride]; f = __fadd_rn(f, __fmul_rn(v, v)); }
#pragma unroll
    for (int l = 0; l < 8; ++l) f = __fadd_rn(f, P[0][l]);
    return f;
}

// grid.x = total row tiles over all pairs (x operand) or column tiles (y operand).
// smem: eT[k][EPAD] + Cs[2][k][8] (+ fT[k][EPAD] when C is given: the projected rows, kept for the fused norms)
template <bool IS_X>
__global__ void __launch_bounds__(PREP_THREADS)
project_kernel(const float* __restrict__ evecs, const float* __restrict__ Cmat, const PairDev* __restrict__ pairs,
               int npairs, int uniform_tiles, int k, float* __restrict__ outT, int64_t ld_total,
               float* __restrict__ nrm, int use_ft) {
    extern __shared__ float smem[];
    float* eT = smem;                 // [k][EPAD]
    float* Cs = smem + (((size_t)k * EPAD + 3) & ~(size_t)3);  // 2 x [k][8], 16-byte aligned
    float* fT = Cs + (size_t)k * 16;  // [k][EPAD]
    const int tid = threadIdx.x;
    const int pi = find_pair(blockIdx.x, pairs, npairs, uniform_tiles, IS_X ? &PairDev::xtile0 : &PairDev::ytile0);
    const PairDev pd = pairs[pi];
    const int tile = blockIdx.x - (IS_X ? pd.xtile0 : pd.ytile0);
    const int n = IS_X ? pd.n1 : pd.n2;
    const int ldin = IS_X ? pd.ldx : pd.ldy;
    const float* src = evecs + (IS_X ? pd.ex_off : pd.ey_off);
    const int row0 = tile * TILE;
    const int64_t col0 = (int64_t)(IS_X ? pd.x_col0 : pd.y_col0) + row0;

    // coalesced read of the [TILE][k] row-major tile, transposed into smem
    for (int idx = tid; idx < TILE * k; idx += PREP_THREADS) {
        const int r = idx / k, b = idx - r * k;
        const int gi = row0 + r;
        eT[b * EPAD + r] = (gi < n) ? __ldg(src + (int64_t)gi * ldin + b) : 0.f;
    }
    __syncthreads();

    const int r = tid & (TILE - 1);
    const int half = tid >> 7;  // 0/1: which of two interleaved 8-wide output chunks this thread computes

    if (Cmat == nullptr) {
        // transpose only: outT[c][col0+r] = e[r][c]
        for (int c = half; c < k; c += 2) outT[(int64_t)c * ld_total + col0 + r] = eT[c * EPAD + r];
        // norms of the rows as they are (padding rows: +inf for the streamed operand so they never win, else 0)
        if (tid < TILE) nrm[col0 + tid] = (row0 + tid < n) ? aten_sqnorm(eT + tid, EPAD, k, 1.f) : (IS_X ? __int_as_float(0x7f800000) : 0.f);
        return;
    }
    const float* Cp = Cmat + pd.c_off;
    for (int a0 = 0; a0 < k; a0 += 16) {
        // stage Cs[h][b][q] = C[a0 + 8h + q][b] for both halves
        __syncthreads();
        for (int idx = tid; idx < 16 * k; idx += PREP_THREADS) {
            const int q = idx / k, b = idx - q * k;
            const int a = a0 + q;
            Cs[(q >> 3) * (k * 8) + b * 8 + (q & 7)] = (a < k) ? __ldg(Cp + (int64_t)a * k + b) : 0.f;
        }
        __syncthreads();
        const float* Ch = Cs + half * (k * 8);
        float acc[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) acc[q] = 0.f;
        for (int b = 0; b < k; ++b) {
            const float e = eT[b * EPAD + r];
            const float4 c0 = *reinterpret_cast<const float4*>(Ch + b * 8);
            const float4 c1 = *reinterpret_cast<const float4*>(Ch + b * 8 + 4);
            acc[0] = __fmaf_rn(e, c0.x, acc[0]); acc[1] = __fmaf_rn(e, c0.y, acc[1]);
            acc[2] = __fmaf_rn(e, c0.z, acc[2]); acc[3] = __fmaf_rn(e, c0.w, acc[3]);
            acc[4] = __fmaf_rn(e, c1.x, acc[4]); acc[5] = __fmaf_rn(e, c1.y, acc[5]);
            acc[6] = __fmaf_rn(e, c1.z, acc[6]); acc[7] = __fmaf_rn(e, c1.w, acc[7]);
        }
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            const int a = a0 + half * 8 + q;
            if (a < k) { outT[(int64_t)a * ld_total + col0 + r] = -2.0f * acc[q]; if (use_ft) fT[a * EPAD + r] = acc[q]; }
        }
    }
    // fused row norms of the projected points: from the shared-memory copy, or (k too large for a second [k][129]
    // buffer) from this CTA's own stores, which the barrier makes visible
    __syncthreads();
    if (tid < TILE) {
        float v = IS_X ? __int_as_float(0x7f800000) : 0.f;
        if (row0 + tid < n)
            v = use_ft ? aten_sqnorm(fT + tid, EPAD, k, 1.f) : aten_sqnorm(outT + col0 + tid, ld_total, k, -0.5f);
        nrm[col0 + tid] = v;
    }
}

// Host launchers -----------------------------------------------------------------------------------
int launch_project(bool is_x, const float* evecs, const float* Cmat, const PairDev* pairs_dev, int npairs,
                   int uniform_tiles, int total_tiles, int k, float* outT, int64_t ld_total, float* nrm,
                   cudaStream_t stream) {
    if (total_tiles <= 0) return 0;
    const size_t base = ((((size_t)k * EPAD + 3) & ~(size_t)3) + (size_t)k * 16) * sizeof(float);
    const int use_ft = Cmat && base + (size_t)k * EPAD * sizeof(float) <= 112 * 1024;   // keep 2 CTAs per SM
    const size_t smem = base + (use_ft ? (size_t)k * EPAD * sizeof(float) : 0);
    if (is_x) {
        static bool attr_done = false;
        if (!attr_done) { F2P_CUDA(cudaFuncSetAttribute(project_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024)); attr_done = true; }
        project_kernel<true><<<total_tiles, PREP_THREADS, smem, stream>>>(evecs, Cmat, pairs_dev, npairs, uniform_tiles, k, outT, ld_total, nrm, use_ft);
    } else {
        static bool attr_done = false;
        if (!attr_done) { F2P_CUDA(cudaFuncSetAttribute(project_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024)); attr_done = true; }
        project_kernel<false><<<total_tiles, PREP_THREADS, smem, stream>>>(evecs, Cmat, pairs_dev, npairs, uniform_tiles, k, outT, ld_total, nrm, use_ft);
    }
    F2P_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace f2p

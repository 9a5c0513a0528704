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
constexpr int ELD = TILE + 4;   // row of the transposed evecs tile: float4-aligned; the loader's lane mapping keeps stores conflict-free

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

// 4-byte asynchronous global -> shared copy (LDGSTS): no register staging, so a thread keeps all its loads in flight;
// `live == false` zero-fills the destination without reading the source.
__device__ __forceinline__ void cp_async4(float* smem_dst, const float* gsrc, bool live) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(live ? 4 : 0) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// [TILE][k] row-major tile of points -> eT[b][r] (row stride ELD), issued as asynchronous copies (the caller waits).
// A warp reads 8 consecutive b of 4 consecutive rows per step (whole 32-byte sectors) and its 32 transposed stores hit
// 32 different banks ((4*b + r) mod 32); warps walk 4-row blocks, lanes keep their (b, r) offsets: no integer
// divisions, and every lane has its TILE*k/256 loads outstanding at once (a register-staged loop kept 4: the kernel
// then ran at a quarter of the bytes in flight that HBM latency asks for).  Rows >= n are zero-filled.
__device__ __forceinline__ void load_tile_T_async(float* eT, const float* __restrict__ src, int ldin, int row0, int n, int k) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int bl = lane & 7, rl = lane >> 3;
    for (int rb = warp; rb < TILE / 4; rb += PREP_THREADS / 32) {
        const int r = rb * 4 + rl, gi = row0 + r;
        const bool live = gi < n;
        const float* p = live ? src + (int64_t)gi * ldin : src;
        for (int b = bl; b < k; b += 8) cp_async4(eT + b * ELD + r, p + b, live);
    }
}

// grid.x = total row tiles over all pairs (x operand) or column tiles (y operand).
// Transpose only (the operand that is not projected): outT[c][col0+r] = e[r][c], norms of the rows as they are.
// smem: eT[k][ELD]
template <bool IS_X>
__device__ __forceinline__ void
transpose_body(int bid, float* smem, const float* __restrict__ evecs, const PairDev* __restrict__ pairs, int npairs,
               int uniform_tiles, int k, float* __restrict__ outT, int64_t ld_total, float* __restrict__ nrm,
               u64* __restrict__ best_init, int* __restrict__ tilecnt) {
    float* eT = smem;                 // [k][ELD]
    const int tid = threadIdx.x;
    const int pi = find_pair(bid, pairs, npairs, uniform_tiles, IS_X ? &PairDev::xtile0 : &PairDev::ytile0);
    const PairDev pd = pairs[pi];
    const int tile = bid - (IS_X ? pd.xtile0 : pd.ytile0);
    const int n = IS_X ? pd.n1 : pd.n2;
    const int ldin = IS_X ? pd.ldx : pd.ldy;
    const float* src = evecs + (IS_X ? pd.ex_off : pd.ey_off);
    const int row0 = tile * TILE;
    const int64_t col0 = (int64_t)(IS_X ? pd.x_col0 : pd.y_col0) + row0;

    load_tile_T_async(eT, src, ldin, row0, n, k);
    cp_async_wait_all();
    __syncthreads();
    const int r = tid & (TILE - 1);
    for (int c = tid >> 7; c < k; c += 2) outT[(int64_t)c * ld_total + col0 + r] = eT[c * ELD + r];
    // padding rows: +inf for the streamed operand so they never win, else 0
    if (tid < TILE) nrm[col0 + tid] = (row0 + tid < n) ? aten_sqnorm(eT + tid, ELD, k, 1.f) : (IS_X ? __int_as_float(0x7f800000) : 0.f);
    // fused arg-min state of this column tile (saves the init launch of kernel 2): merge keys and the arrival counter
    if (best_init && tid >= TILE) best_init[col0 + tid - TILE] = ~0ull;
    if (tilecnt && tid == 0) tilecnt[bid] = 0;
}

template <bool IS_X>
__global__ void __launch_bounds__(PREP_THREADS)
transpose_kernel(const float* __restrict__ evecs, const PairDev* __restrict__ pairs, int npairs, int uniform_tiles, int k,
                 float* __restrict__ outT, int64_t ld_total, float* __restrict__ nrm, u64* __restrict__ best_init,
                 int* __restrict__ tilecnt) {
    extern __shared__ __align__(16) float smem[];
    transpose_body<IS_X>(blockIdx.x, smem, evecs, pairs, npairs, uniform_tiles, k, outT, ld_total, nrm, best_init, tilecnt);
}

// Projection: feat = e . C^T for a [TILE][k] tile of points, written k-major as -2*feat, with the fused row norms.
// A register-tiled FP32 product: warp = 8 consecutive outputs a (one pass covers 8 warps x 8 = 64 outputs), lane = 4
// consecutive points, 4 x 8 accumulators per thread held as packed pairs along a.  Every output keeps the reference's
// chain — one accumulator, b ascending, fused multiply-add (`fma.rn.f32x2` = two independent IEEE FMAs).  Per b a
// thread issues one LDS.128 of points (4 wavefronts per warp), two broadcast LDS.128 of C and 16 packed FMAs.
// smem: eT[k][ELD] + Cs[k][CLD] (+ fT[k][ELD] when it fits: the projected rows, kept for the fused norms)
constexpr int CLD = 64 + 4;     // one pass of outputs + pad (staging stores: 4-way conflicts instead of 32-way)

template <bool IS_X>
__device__ __forceinline__ void
project_body(int bid, float* smem, const float* __restrict__ evecs, const float* __restrict__ Cmat,
             const PairDev* __restrict__ pairs, int npairs, int uniform_tiles, int k, float* __restrict__ outT,
             int64_t ld_total, float* __restrict__ nrm, int use_ft) {
    float* eT = smem;                          // [k][ELD]
    float* Cs = eT + (size_t)k * ELD;          // [k][CLD]: Cs[b][q] = C[a0 + q][b]
    float* fT = Cs + (size_t)k * CLD;          // [k][ELD]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int pi = find_pair(bid, pairs, npairs, uniform_tiles, IS_X ? &PairDev::xtile0 : &PairDev::ytile0);
    const PairDev pd = pairs[pi];
    const int tile = bid - (IS_X ? pd.xtile0 : pd.ytile0);
    const int n = IS_X ? pd.n1 : pd.n2;
    const int ldin = IS_X ? pd.ldx : pd.ldy;
    const float* src = evecs + (IS_X ? pd.ex_off : pd.ey_off);
    const int row0 = tile * TILE;
    const int64_t col0 = (int64_t)(IS_X ? pd.x_col0 : pd.y_col0) + row0;

    load_tile_T_async(eT, src, ldin, row0, n, k);

    const float* Cp = Cmat + pd.c_off;
    const float* ep = eT + 4 * lane;
    const float* cp = Cs + 8 * warp;
    for (int a0 = 0; a0 < k; a0 += 64) {
        if (a0) __syncthreads();               // the previous pass has finished reading Cs
        for (int q = warp; q < 64; q += PREP_THREADS / 32) {      // warp = one row of C, lanes along b (no divisions)
            const int a = a0 + q;
            const bool live = a < k;
            const float* crow = live ? Cp + (int64_t)a * k : Cp;
            for (int b = lane; b < k; b += 32) cp_async4(Cs + b * CLD + q, crow + b, live);
        }
        cp_async_wait_all();                   // first pass: the point tile too
        __syncthreads();
        const int aw = a0 + 8 * warp;          // this warp's 8 outputs
        if (aw >= k) continue;
        u64 acc[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = 0ull;
#pragma unroll 2
        for (int b = 0; b < k; ++b) {
            const float4 e = *reinterpret_cast<const float4*>(ep + b * ELD);
            const ulonglong2 c0 = *reinterpret_cast<const ulonglong2*>(cp + b * CLD);
            const ulonglong2 c1 = *reinterpret_cast<const ulonglong2*>(cp + b * CLD + 4);
            const float ev[4] = {e.x, e.y, e.z, e.w};
            const u64 cv[4] = {c0.x, c0.y, c1.x, c1.y};
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const u64 ee = pack2(ev[i], ev[i]);
#pragma unroll
                for (int j = 0; j < 4; ++j) fma2(acc[i][j], ee, cv[j]);
            }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float lo[4], hi[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) unpack2(acc[i][j], lo[i], hi[i]);
            const int a = aw + 2 * j;
            if (a < k) {
                *reinterpret_cast<float4*>(outT + (int64_t)a * ld_total + col0 + 4 * lane) =
                    make_float4(-2.0f * lo[0], -2.0f * lo[1], -2.0f * lo[2], -2.0f * lo[3]);
                if (use_ft) *reinterpret_cast<float4*>(fT + a * ELD + 4 * lane) = make_float4(lo[0], lo[1], lo[2], lo[3]);
            }
            if (a + 1 < k) {
                *reinterpret_cast<float4*>(outT + (int64_t)(a + 1) * ld_total + col0 + 4 * lane) =
                    make_float4(-2.0f * hi[0], -2.0f * hi[1], -2.0f * hi[2], -2.0f * hi[3]);
                if (use_ft) *reinterpret_cast<float4*>(fT + (a + 1) * ELD + 4 * lane) = make_float4(hi[0], hi[1], hi[2], hi[3]);
            }
        }
    }
    // fused row norms of the projected points: from the shared-memory copy, or (k too large for a second [k][ELD]
    // buffer) from this CTA's own stores, which the barrier makes visible
    __syncthreads();
    if (tid < TILE) {
        float v = IS_X ? __int_as_float(0x7f800000) : 0.f;
        if (row0 + tid < n)
            v = use_ft ? aten_sqnorm(fT + tid, ELD, k, 1.f) : aten_sqnorm(outT + col0 + tid, ld_total, k, -0.5f);
        nrm[col0 + tid] = v;
    }
}

template <bool IS_X>
__global__ void __launch_bounds__(PREP_THREADS)
project_kernel(const float* __restrict__ evecs, const float* __restrict__ Cmat, const PairDev* __restrict__ pairs,
               int npairs, int uniform_tiles, int k, float* __restrict__ outT, int64_t ld_total,
               float* __restrict__ nrm, int use_ft) {
    extern __shared__ __align__(16) float smem[];
    project_body<IS_X>(blockIdx.x, smem, evecs, Cmat, pairs, npairs, uniform_tiles, k, outT, ld_total, nrm, use_ft);
}

// Both operands of a batch in ONE launch: CTAs [0, total_xtiles) project the CAD eigenbasis, the rest transpose the scene
// eigenbasis (and reset the arg-min merge state of their column tile).  A single 5k x 2k pair is 5 short kernels; each
// launch boundary costs ~2.5 us of drain + launch latency next to ~15 us of work.
__global__ void __launch_bounds__(PREP_THREADS)
prep_kernel(const float* __restrict__ ex, const float* __restrict__ Cmat, const float* __restrict__ ey,
            const PairDev* __restrict__ pairs, int npairs, int uni_xtiles, int total_xtiles, int uni_ytiles, int k,
            float* __restrict__ xT, int64_t ld1_total, float* __restrict__ nx, float* __restrict__ yT, int64_t ld2_total,
            float* __restrict__ ny, int use_ft, u64* __restrict__ best_init, int* __restrict__ tilecnt) {
    extern __shared__ __align__(16) float smem[];
    // lets a dependent launch (the small-problem arg-min kernel) start its prologue now; it still waits for this whole
    // grid (griddepcontrol.wait) before reading anything written here
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    if ((int)blockIdx.x < total_xtiles)
        project_body<true>(blockIdx.x, smem, ex, Cmat, pairs, npairs, uni_xtiles, k, xT, ld1_total, nx, use_ft);
    else
        transpose_body<false>(blockIdx.x - total_xtiles, smem, ey, pairs, npairs, uni_ytiles, k, yT, ld2_total, ny,
                              best_init, tilecnt);
}

// Host launchers -----------------------------------------------------------------------------------
template <bool IS_X>
static int launch_project_t(const float* evecs, const float* Cmat, const PairDev* pairs_dev, int npairs, int uniform_tiles,
                            int total_tiles, int k, float* outT, int64_t ld_total, float* nrm, cudaStream_t stream,
                            u64* best_init = nullptr, int* tilecnt = nullptr) {
    if (Cmat == nullptr) {
        const size_t smem = (size_t)k * ELD * sizeof(float);
        // (the attribute is per device: set on every launch, as nn.cu does — a process-wide flag broke the second GPU)
        F2P_CUDA(cudaFuncSetAttribute(transpose_kernel<IS_X>, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));
        transpose_kernel<IS_X><<<total_tiles, PREP_THREADS, smem, stream>>>(evecs, pairs_dev, npairs, uniform_tiles, k, outT, ld_total, nrm, best_init, tilecnt);
    } else {
        const size_t base = (size_t)k * (ELD + CLD) * sizeof(float);
        const int use_ft = base + (size_t)k * ELD * sizeof(float) <= 112 * 1024;   // keep 2 CTAs per SM
        const size_t smem = base + (use_ft ? (size_t)k * ELD * sizeof(float) : 0);
        F2P_CUDA(cudaFuncSetAttribute(project_kernel<IS_X>, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));
        project_kernel<IS_X><<<total_tiles, PREP_THREADS, smem, stream>>>(evecs, Cmat, pairs_dev, npairs, uniform_tiles, k, outT, ld_total, nrm, use_ft);
    }
    F2P_CUDA(cudaGetLastError());
    return 0;
}

constexpr int PREP_FUSE_MAX_TILES = 8 * 148;

int launch_prep(const float* ex, const float* Cmat, const float* ey, const PairDev* pairs_dev, int npairs, int uni_xtiles,
                int total_xtiles, int uni_ytiles, int total_ytiles, int k, float* xT, int64_t ld1_total, float* nx,
                float* yT, int64_t ld2_total, float* ny, unsigned long long* best_init, int* tilecnt, cudaStream_t stream) {
    if (total_xtiles + total_ytiles <= 0) return 0;
    if (total_xtiles + total_ytiles > PREP_FUSE_MAX_TILES) {
        // big batches: two launches, so the transposing CTAs keep their own (smaller) shared-memory footprint — the
        // launch boundary is noise there (256 pairs: 0.33 ms as two launches, 0.38 ms as one)
        int rc = total_xtiles > 0 ? launch_project_t<true>(ex, Cmat, pairs_dev, npairs, uni_xtiles, total_xtiles, k, xT, ld1_total, nx, stream) : 0;
        if (rc || total_ytiles <= 0) return rc;
        return launch_project_t<false>(ey, nullptr, pairs_dev, npairs, uni_ytiles, total_ytiles, k, yT, ld2_total, ny, stream,
                                       reinterpret_cast<u64*>(best_init), tilecnt);
    }
    const size_t base = (size_t)k * (ELD + CLD) * sizeof(float);
    const int use_ft = base + (size_t)k * ELD * sizeof(float) <= 112 * 1024;   // keep 2 CTAs per SM
    const size_t smem = base + (use_ft ? (size_t)k * ELD * sizeof(float) : 0);
    F2P_CUDA(cudaFuncSetAttribute(prep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));
    prep_kernel<<<total_xtiles + total_ytiles, PREP_THREADS, smem, stream>>>(ex, Cmat, ey, pairs_dev, npairs, uni_xtiles,
        total_xtiles, uni_ytiles, k, xT, ld1_total, nx, yT, ld2_total, ny, use_ft, best_init, tilecnt);
    F2P_CUDA(cudaGetLastError());
    return 0;
}

int launch_project(bool is_x, const float* evecs, const float* Cmat, const PairDev* pairs_dev, int npairs,
                   int uniform_tiles, int total_tiles, int k, float* outT, int64_t ld_total, float* nrm,
                   cudaStream_t stream) {
    if (total_tiles <= 0) return 0;
    return is_x ? launch_project_t<true>(evecs, Cmat, pairs_dev, npairs, uniform_tiles, total_tiles, k, outT, ld_total, nrm, stream)
                : launch_project_t<false>(evecs, Cmat, pairs_dev, npairs, uniform_tiles, total_tiles, k, outT, ld_total, nrm, stream);
}

}  // namespace f2p

// nn.cu — kernel 2: fused all-pairs squared distance + running arg-min / top-5 over k-major operands.
//
// Replaces torch.cdist + argmin(dim=-2) (naive.py:33-34) and torch.cdist + sort(dim=-2)[:5]
// (spacial_filtering.py:32-35) without ever materialising the [V1,V2] matrix.
//
// Arithmetic contract (SURVEY 8(a).1-3,-5,-6, pinned by oracle/f2p_oracle.c):
//   acc = 0; acc = fma(xT[c][i], yT[c][j], acc) for c = 0..k-1 (xT already holds -2*feat_x, exact);
//   acc = acc + nx[i]; acc = acc + ny[j]; d = sqrt_rn(max(acc, 0));
//   arg-min over i with the lowest index winning exact ties in d; top-5 ordered by (d, i).
// Packed fma.rn.f32x2 is two independent IEEE fused multiply-adds, so pairing columns changes no bit.
//
// Structure: one CTA owns a 128-column tile of one pair and streams 128-row tiles of the other operand.
// The column operand (k x 128) is TMA-loaded once and stays in shared memory; the row operand arrives in
// KC-row TMA stages through an mbarrier ring.  256 threads, each an 8x8 register tile (32 packed accumulators),
// running (distance, index) minima live in registers across all row tiles of the CTA.
#include <stdlib.h>
#include "common.cuh"

namespace f2p {

struct NNParams {
    const PairDev* pairs;
    int npairs;
    int uniform_items;   // >0: every pair has this many work items (pair = item / uniform_items)
    int k;
    const float* nx;     // [ld1_total] row norms (+inf on padding)
    const float* ny;     // [ld2_total] column norms
    u64* best;           // MODE 0: [ld2_total] packed (d bits << 32 | i) keys, pre-set to ~0
    u64* part5;          // MODE 1: [total_items][TILE][5] keys
};

constexpr uint32_t IDX_NONE = 0x7fffffffu;

__device__ __forceinline__ u64 make_key(float d, uint32_t i) {
    return ((u64)(__float_as_uint(d) & 0x7fffffffu) << 32) | (u64)i;
}
__device__ __forceinline__ u64 shfl_xor_u64(u64 v, int m) {
    return __shfl_xor_sync(0xffffffffu, v, m);
}
// insert `key` into the ascending list l[0..4], dropping the largest
__device__ __forceinline__ void insert5(u64 (&l)[5], u64 key) {
#pragma unroll
    for (int r = 0; r < 5; ++r) {
        const bool p = key < l[r];
        const u64 t = l[r];
        l[r] = p ? key : t;
        key = p ? t : key;
    }
}

// MODE 0 arg-min, 1 top-5, 2 = measurement-only variant without the distance epilogue (never used by the API).
template <int MODE, int KC, int NSTAGE>
__global__ void __launch_bounds__(NN_THREADS, MODE == 1 ? 1 : 2)
nn_kernel(const __grid_constant__ CUtensorMap mapX, const __grid_constant__ CUtensorMap mapY, const NNParams prm) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    const int k = prm.k;
    const int nkc = (k + KC - 1) / KC;
    float* sB = reinterpret_cast<float*>(smem_raw);                 // [nkc*KC][TILE]
    float* sA = sB + (size_t)nkc * KC * TILE;                        // NSTAGE x [KC][TILE]
    uint64_t* bars = reinterpret_cast<uint64_t*>(sA + (size_t)NSTAGE * KC * TILE);
    uint64_t* fullB = bars;            // 1
    uint64_t* full = bars + 1;         // NSTAGE
    uint64_t* empty = bars + 1 + NSTAGE;  // NSTAGE
    u64* red = reinterpret_cast<u64*>(bars + 16);                    // MODE 0: [8][TILE]; MODE 1: [8][TILE][5]

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tx = tid & 15, ty = tid >> 4;

    const int item = blockIdx.x;
    const int pi = find_pair(item, prm.pairs, prm.npairs, prm.uniform_items, &PairDev::item0);
    const PairDev* pdp = prm.pairs + pi;
    const int pd_nsplit = pdp->nsplit, pd_nrt = pdp->nrt, pd_n2 = pdp->n2;
    const int local = item - pdp->item0;
    const int ct = local / pd_nsplit, sp = local - ct * pd_nsplit;
    const int t_begin = (int)(((long long)sp * pd_nrt) / pd_nsplit);
    const int t_end = (int)(((long long)(sp + 1) * pd_nrt) / pd_nsplit);
    const int ntiles = t_end - t_begin;
    const int xcol = pdp->x_col0, ycol = pdp->y_col0 + ct * TILE;
    const int total_stages = ntiles * nkc;

    if (tid == 0) {
        tma_prefetch_desc(&mapX);
        tma_prefetch_desc(&mapY);
        mbar_init(fullB, 1);
        for (int s = 0; s < NSTAGE; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], NN_THREADS / 32); }
        fence_barrier_init();
    }
    __syncthreads();

    auto issue = [&](int s) {
        const int slot = s % NSTAGE, t = s / nkc, kc = s - t * nkc;
        mbar_arrive_expect_tx(&full[slot], KC * TILE * sizeof(float));
        tma_load_2d(sA + (size_t)slot * KC * TILE, &mapX, xcol + (t_begin + t) * TILE, kc * KC, &full[slot]);
    };
    if (tid == 0) {
        mbar_arrive_expect_tx(fullB, (uint32_t)(nkc * KC * TILE * sizeof(float)));
        for (int kc = 0; kc < nkc; ++kc) tma_load_2d(sB + (size_t)kc * KC * TILE, &mapY, ycol, kc * KC, fullB);
        for (int s = 0; s < NSTAGE && s < total_stages; ++s) issue(s);
    }

    // column norms for this thread's 8 columns, as packed pairs
    u64 ny2[4];
    {
        const float4 n0 = __ldg(reinterpret_cast<const float4*>(prm.ny + ycol + tx * 4));
        const float4 n1 = __ldg(reinterpret_cast<const float4*>(prm.ny + ycol + 64 + tx * 4));
        ny2[0] = pack2(n0.x, n0.y); ny2[1] = pack2(n0.z, n0.w);
        ny2[2] = pack2(n1.x, n1.y); ny2[3] = pack2(n1.z, n1.w);
    }

    // running minima
    float bs[8]; uint32_t bi[8];          // MODE 0
    u64 l5[MODE == 1 ? 8 : 1][5];         // MODE 1
    const float INF = __int_as_float(0x7f800000);
#pragma unroll
    for (int j = 0; j < 8; ++j) { bs[j] = INF; bi[j] = IDX_NONE; }
    if (MODE == 1) {
#pragma unroll
        for (int j = 0; j < 8; ++j)
#pragma unroll
            for (int r = 0; r < 5; ++r) l5[MODE == 1 ? j : 0][r] = ~0ull;
    }

    mbar_wait(fullB, 0);

    int s = 0;
    for (int t = 0; t < ntiles; ++t) {
        const int rowbase = (t_begin + t) * TILE;
        const float4 nxa = __ldg(reinterpret_cast<const float4*>(prm.nx + xcol + rowbase + ty * 4));
        const float4 nxb = __ldg(reinterpret_cast<const float4*>(prm.nx + xcol + rowbase + 64 + ty * 4));
        u64 acc[8][4];
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = 0ull;

        for (int kc = 0; kc < nkc; ++kc, ++s) {
            const int slot = s % NSTAGE;
            const uint32_t ph = (uint32_t)((s / NSTAGE) & 1);
            if (tid == 0 && s >= 1) {
                const int sn = s - 1 + NSTAGE;
                if (sn < total_stages) {
                    mbar_wait(&empty[(s - 1) % NSTAGE], (uint32_t)(((s - 1) / NSTAGE) & 1));
                    issue(sn);
                }
            }
            mbar_wait(&full[slot], ph);
            const float* A = sA + (size_t)slot * KC * TILE + ty * 4;
            const float* B = sB + (size_t)kc * KC * TILE + tx * 4;
            const int klen = min(KC, k - kc * KC);
            auto step = [&](int kk) {
                const float4 a0 = *reinterpret_cast<const float4*>(A + kk * TILE);
                const float4 a1 = *reinterpret_cast<const float4*>(A + kk * TILE + 64);
                const ulonglong2 b0 = *reinterpret_cast<const ulonglong2*>(B + kk * TILE);
                const ulonglong2 b1 = *reinterpret_cast<const ulonglong2*>(B + kk * TILE + 64);
                const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
                // issue order matters on sm_100: consecutive FFMA2 sharing the PACKED (64-bit) operand hit the operand
                // reuse cache and run at the nominal FP32 rate (74 TFLOP/s measured); sharing only the scalar
                // broadcast operand caps at 63 TFLOP/s (profiles/r01_ffma_peak_v3.log).  Column pair outer, row inner.
                const u64 bp[4] = {b0.x, b0.y, b1.x, b1.y};
#pragma unroll
                for (int j = 0; j < 4; ++j)
#pragma unroll
                    for (int i = 0; i < 8; ++i) fma2v(acc[i][j], pack2(av[i], av[i]), bp[j]);
            };
            if (klen == KC) {
#pragma unroll
                for (int kk = 0; kk < KC; ++kk) step(kk);
            } else {
                for (int kk = 0; kk < klen; ++kk) step(kk);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[slot]);
        }

        // epilogue: (acc + nx) + ny, clamp, sqrt_rn, running (d, i) minimum — rows ascend within a thread
        const float nxv[8] = {nxa.x, nxa.y, nxa.z, nxa.w, nxb.x, nxb.y, nxb.z, nxb.w};
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const uint32_t row = (uint32_t)(rowbase + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4)));
            const u64 nn = pack2(nxv[i], nxv[i]);
#pragma unroll
            for (int jp = 0; jp < 4; ++jp) {
                float v0, v1;
                unpack2(add2(add2(acc[i][jp], nn), ny2[jp]), v0, v1);
                if (MODE == 2) { bs[2 * jp] = fminf(bs[2 * jp], v0); bs[2 * jp + 1] = fminf(bs[2 * jp + 1], v1); continue; }
                const float d0 = __fsqrt_rn(fmaxf(v0, 0.f));
                const float d1 = __fsqrt_rn(fmaxf(v1, 0.f));
                if (MODE == 0) {
                    const bool p0 = d0 < bs[2 * jp], p1 = d1 < bs[2 * jp + 1];
                    bs[2 * jp] = p0 ? d0 : bs[2 * jp];             bi[2 * jp] = p0 ? row : bi[2 * jp];
                    bs[2 * jp + 1] = p1 ? d1 : bs[2 * jp + 1];     bi[2 * jp + 1] = p1 ? row : bi[2 * jp + 1];
                } else {
                    const u64 k0 = make_key(d0, row), k1 = make_key(d1, row);
                    if (k0 < l5[MODE == 1 ? 2 * jp : 0][4]) insert5(l5[MODE == 1 ? 2 * jp : 0], k0);
                    if (k1 < l5[MODE == 1 ? 2 * jp + 1 : 0][4]) insert5(l5[MODE == 1 ? 2 * jp + 1 : 0], k1);
                }
            }
        }
    }

    // ---- cross-thread reduction: 16 threads (2 lanes x 8 warps) share each column -------------------
    __syncthreads();
    if (MODE != 1) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            u64 key = make_key(bs[j], bi[j]);
            const u64 o = shfl_xor_u64(key, 16);
            key = o < key ? o : key;
            const int col = (j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4));
            if (lane < 16) red[warp * TILE + col] = key;
        }
        __syncthreads();
        if (tid < TILE) {
            u64 m = red[tid];
#pragma unroll
            for (int w = 1; w < 8; ++w) { const u64 o = red[w * TILE + tid]; m = o < m ? o : m; }
            if (ct * TILE + tid < pd_n2) atomicMin(prm.best + ycol + tid, m);
        }
    } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            u64 mine[5];
#pragma unroll
            for (int r = 0; r < 5; ++r) mine[r] = l5[MODE == 1 ? j : 0][r];
#pragma unroll
            for (int r = 0; r < 5; ++r) {
                const u64 o = shfl_xor_u64(l5[MODE == 1 ? j : 0][r], 16);
                insert5(mine, o);
            }
            const int col = (j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4));
            if (lane < 16) {
#pragma unroll
                for (int r = 0; r < 5; ++r) red[((size_t)warp * TILE + col) * 5 + r] = mine[r];
            }
        }
        __syncthreads();
        if (tid < TILE) {
            u64 m[5];
#pragma unroll
            for (int r = 0; r < 5; ++r) m[r] = red[(size_t)tid * 5 + r];
            for (int w = 1; w < 8; ++w)
#pragma unroll
                for (int r = 0; r < 5; ++r) insert5(m, red[((size_t)w * TILE + tid) * 5 + r]);
#pragma unroll
            for (int r = 0; r < 5; ++r) prm.part5[((size_t)item * TILE + tid) * 5 + r] = m[r];
        }
    }
}

// ---- finalize: keys -> int64 index maps ---------------------------------------------------------------
__device__ __forceinline__ int64_t row1_value(int j, unsigned flags) {
    // linspace(0,V2-1,V2).type(int16) promoted to int64 (naive.py:21): wraps modulo 2^16 on the reference's CPU path
    return (flags & F2P_INT16_COMPAT) ? (int64_t)(int16_t)(uint16_t)(j & 0xffff) : (int64_t)j;
}

// grid.x = total column tiles; one thread per column.  out[pair] = int64 [2,n2].
__global__ void __launch_bounds__(TILE)
finalize_argmin_kernel(const u64* __restrict__ best, const PairDev* __restrict__ pairs, int npairs, int uniform_tiles,
                       unsigned flags, int64_t* __restrict__ out) {
    const int pi = find_pair(blockIdx.x, pairs, npairs, uniform_tiles, &PairDev::ytile0);
    const PairDev* pd = pairs + pi;
    const int j = (blockIdx.x - pd->ytile0) * TILE + threadIdx.x;
    const int n2 = pd->n2;
    if (j >= n2) return;
    uint32_t idx = (uint32_t)(best[pd->y_col0 + j] & 0xffffffffull);
    if (idx == IDX_NONE) idx = 0;  // all distances were +inf/NaN: out of contract, mirror argmin's first index
    int64_t* o = out + pd->out_off;
    o[j] = (int64_t)idx;
    if (!(flags & F2P_ROW0_ONLY)) o[(int64_t)n2 + j] = row1_value(j, flags);
}

// Merges the nsplit partial top-5 lists of every column and writes nn_query's [2, 5*n2] candidate layout
// (spacial_filtering.py:36-38): column 5j+r = (r-th nearest CAD row of scene point j, j).
__global__ void __launch_bounds__(TILE)
finalize_top5_kernel(const u64* __restrict__ part5, const PairDev* __restrict__ pairs, int npairs, int uniform_tiles,
                     unsigned flags, int64_t* __restrict__ cand, int cand_is_out_layout) {
    const int pi = find_pair(blockIdx.x, pairs, npairs, uniform_tiles, &PairDev::ytile0);
    const PairDev* pd = pairs + pi;
    const int ct = blockIdx.x - pd->ytile0;
    const int j = ct * TILE + threadIdx.x;
    const int n2 = pd->n2;
    if (j >= n2) return;
    const int nsplit = pd->nsplit;
    const size_t item_base = (size_t)pd->item0 + (size_t)ct * nsplit;
    u64 m[5];
#pragma unroll
    for (int r = 0; r < 5; ++r) m[r] = part5[((item_base)*TILE + threadIdx.x) * 5 + r];
    for (int sp = 1; sp < nsplit; ++sp)
#pragma unroll
        for (int r = 0; r < 5; ++r) insert5(m, part5[((item_base + sp) * TILE + threadIdx.x) * 5 + r]);
    const int64_t M = 5ll * n2;
    int64_t* o = cand + (cand_is_out_layout ? pd->out_off : 2ll * pd->cand0);
    const int64_t r1 = row1_value(j, flags);
#pragma unroll
    for (int r = 0; r < 5; ++r) {
        o[5ll * j + r] = (int64_t)(uint32_t)(m[r] & 0xffffffffull);
        if (!(flags & F2P_ROW0_ONLY)) o[M + 5ll * j + r] = r1;
    }
}

__global__ void fill_u64_kernel(u64* p, size_t n, u64 v) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

// ---- host launchers -----------------------------------------------------------------------------------
static size_t nn_smem_bytes_v(int k, int mode, int kc, int nst) {
    const int nkc = (k + kc - 1) / kc;
    size_t b = ((size_t)nkc * kc * TILE + (size_t)nst * kc * TILE) * sizeof(float) + 16 * sizeof(uint64_t);
    b += (mode != 1 ? (size_t)8 * TILE : (size_t)8 * TILE * 5) * sizeof(u64);
    return b;
}
size_t nn_smem_bytes(int k, int mode) { return nn_smem_bytes_v(k, mode, KC, NSTAGE); }

template <int MODE, int KCT, int NST>
static int launch_variant(const float* xT, int64_t ld1_total, const float* yT, int64_t ld2_total, int k, const NNParams& prm,
                          int total_items, cudaStream_t stream) {
    CUtensorMap mapX, mapY;
    int rc = make_tensor_map_2d(&mapX, xT, (uint64_t)ld1_total, (uint64_t)k, (uint64_t)ld1_total, TILE, KCT);
    if (rc) return rc;
    rc = make_tensor_map_2d(&mapY, yT, (uint64_t)ld2_total, (uint64_t)k, (uint64_t)ld2_total, TILE, KCT);
    if (rc) return rc;
    const size_t smem = nn_smem_bytes_v(k, MODE, KCT, NST);
    F2P_CUDA(cudaFuncSetAttribute(nn_kernel<MODE, KCT, NST>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    nn_kernel<MODE, KCT, NST><<<total_items, NN_THREADS, smem, stream>>>(mapX, mapY, prm);
    return 0;
}

int launch_nn(int mode, const float* xT, int64_t ld1_total, const float* nx, const float* yT, int64_t ld2_total,
              const float* ny, int k, const PairDev* pairs_dev, int npairs, int uniform_items, int total_items,
              u64* best, u64* part5, cudaStream_t stream) {
    if (total_items <= 0) return 0;
    NNParams prm;
    prm.pairs = pairs_dev; prm.npairs = npairs; prm.uniform_items = uniform_items; prm.k = k;
    prm.nx = nx; prm.ny = ny; prm.best = best; prm.part5 = part5;
    int rc = 0;
    if (mode == 0) {
        fill_u64_kernel<<<(unsigned)((ld2_total + 255) / 256), 256, 0, stream>>>(best, (size_t)ld2_total, ~0ull);
        // F2P_NN_VARIANT: developer switch for pipeline-shape experiments (profiles/); default = 0
        static int variant = -1;
        if (variant < 0) { const char* e = getenv("F2P_NN_VARIANT"); variant = e ? atoi(e) : 0; }
        switch (variant) {
            case 1: rc = launch_variant<0, 32, 3>(xT, ld1_total, yT, ld2_total, k, prm, total_items, stream); break;
            case 2: rc = launch_variant<0, 8, 8>(xT, ld1_total, yT, ld2_total, k, prm, total_items, stream); break;
            case 3: rc = launch_variant<2, 16, 4>(xT, ld1_total, yT, ld2_total, k, prm, total_items, stream); break;
            case 4: rc = launch_variant<2, 32, 3>(xT, ld1_total, yT, ld2_total, k, prm, total_items, stream); break;
            default: rc = launch_variant<0, KC, NSTAGE>(xT, ld1_total, yT, ld2_total, k, prm, total_items, stream);
        }
    } else {
        rc = launch_variant<1, KC, NSTAGE>(xT, ld1_total, yT, ld2_total, k, prm, total_items, stream);
    }
    if (rc) return rc;
    F2P_CUDA(cudaGetLastError());
    return 0;
}

int launch_finalize_argmin(const u64* best, const PairDev* pairs_dev, int npairs, int uniform_tiles, int total_ytiles,
                           unsigned flags, int64_t* out, cudaStream_t stream) {
    if (total_ytiles <= 0) return 0;
    finalize_argmin_kernel<<<total_ytiles, TILE, 0, stream>>>(best, pairs_dev, npairs, uniform_tiles, flags, out);
    F2P_CUDA(cudaGetLastError());
    return 0;
}

int launch_finalize_top5(const u64* part5, const PairDev* pairs_dev, int npairs, int uniform_tiles, int total_ytiles,
                         unsigned flags, int64_t* cand, int cand_is_out_layout, cudaStream_t stream) {
    if (total_ytiles <= 0) return 0;
    finalize_top5_kernel<<<total_ytiles, TILE, 0, stream>>>(part5, pairs_dev, npairs, uniform_tiles, flags, cand, cand_is_out_layout);
    F2P_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace f2p

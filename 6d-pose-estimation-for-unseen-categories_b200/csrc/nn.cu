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
#include "internal.h"

namespace f2p {

struct NNParams {
    const PairDev* pairs;
    int npairs;
    int uniform_items;   // >0: every pair has this many work items (pair = item / uniform_items)
    int k;
    const float* nx;     // [ld1_total] row norms (+inf on padding)
    const float* ny;     // [ld2_total] column norms
    u64* best;           // MODE 0/2: [ld2_total] packed (value bits << 32 | i) keys, pre-set to ~0
    u64* part5;          // MODE 1: [total_items][TILE][5] keys
    uint32_t* best2;     // MODE 2: [ld2_total] bits of the second-smallest clamped d2 per column, pre-set to +inf bits
    const int* ncols;    // MODE 0 re-check pass: per-pair number of (compacted) columns to process, or null = n2.
                         // Gated kernels behind the tensor-pipe screen (tc.cu) read it as [total column tiles] flags and
                         // process only the tiles the screen flagged (null = everything) — the struct keeps its size, the
                         // gated kernels' register allocation at the 128-register cap is sensitive to it
    int* tileflag;       // [total column tiles]: gated kernel sets 1 when a candidate list overflowed; MODE 0/1 launched
                         // as the exact fallback process only flagged column tiles (null = process everything)
    int* tilecnt;        // gated arg-min, fused finalise: [total column tiles] arrival counters (pre-set to 0), or null
    int64_t* out;        // fused finalise: the batch's int64 result buffer (PairDev::out_off)
    unsigned flags;      // fused finalise: F2P_INT16_COMPAT / F2P_ROW0_ONLY
};

constexpr uint32_t IDX_NONE = 0x7fffffffu;

__device__ __forceinline__ int64_t row1_value(int j, unsigned flags) {
    // linspace(0,V2-1,V2).type(int16) promoted to int64 (naive.py:21): wraps modulo 2^16 on the reference's CPU path
    return (flags & F2P_INT16_COMPAT) ? (int64_t)(int16_t)(uint16_t)(j & 0xffff) : (int64_t)j;
}
constexpr int PRODUCER_TID = NN_THREADS - 32;   // lane 0 of the last (highest-priority) warp issues all TMA loads

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

// MODE 0: exact arg-min (sqrt_rn per element, reference semantics by construction).
// MODE 1: exact top-5 ordered by (distance, index).
// MODE 2: raw first pass of the two-phase arg-min: tracks the smallest clamped d2 with its row and the second-smallest
//         value per column, no square roots.  Columns whose two smallest d2 fall into the same sqrt_rn class are
//         re-checked by a MODE 0 pass over the compacted list of those columns (finalize_raw_kernel decides).
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
    // Thread -> (ty, tx) of the 16x16 grid of 8x8 register tiles.  Shared memory serves LDS.128 per HALF-warp, one
    // wavefront per 128 distinct bytes: a half-warp laid out 2 (rows) x 8 (columns) reads 32 B of the row operand and
    // 128 B of the column operand per load = 1 wavefront each (8 per warp k-step instead of 12 for a 1 x 16 layout).
    const int tx = ((warp & 1) << 3) | (lane & 7);
    const int ty = ((warp >> 1) << 2) | ((lane >> 4) << 1) | ((lane >> 3) & 1);

    const int item = blockIdx.x;
    const int pi = find_pair(item, prm.pairs, prm.npairs, prm.uniform_items, &PairDev::item0);
    const PairDev* pdp = prm.pairs + pi;
    const int pd_nsplit = pdp->nsplit, pd_nrt = pdp->nrt;
    const int pd_n2 = (MODE == 0 && prm.ncols) ? prm.ncols[pi] : pdp->n2;
    const int local = item - pdp->item0;
    const int ct = local / pd_nsplit, sp = local - ct * pd_nsplit;
    const int t_begin = (int)(((long long)sp * pd_nrt) / pd_nsplit);
    const int t_end = (int)(((long long)(sp + 1) * pd_nrt) / pd_nsplit);
    const int ntiles = t_end - t_begin;
    const int xcol = pdp->x_col0, ycol = pdp->y_col0 + ct * TILE;
    const int total_stages = ntiles * nkc;
    if (ct * TILE >= pd_n2) return;   // re-check pass: fewer compacted columns than the grid was sized for (uniform exit)
    if (MODE != 2 && prm.tileflag && prm.tileflag[pdp->ytile0 + ct] == 0) return;   // exact fallback: flagged tiles only

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
    if (tid == PRODUCER_TID) {
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
    float bs[8]; uint32_t bi[8];          // MODE 0: best distance / row; MODE 2: smallest clamped d2 / row
    float b2[MODE == 2 ? 8 : 1];          // MODE 2: second-smallest clamped d2
    u64 l5[MODE == 1 ? 8 : 1][5];         // MODE 1
    const float INF = __int_as_float(0x7f800000);
#pragma unroll
    for (int j = 0; j < 8; ++j) { bs[j] = INF; bi[j] = IDX_NONE; b2[MODE == 2 ? j : 0] = INF; }
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
            // The refilling thread must belong to the warp that runs AHEAD: the SM's warp arbiter favours the highest
            // warp id, so a low-id producer lags and the leading warps find their stage not even requested yet
            // (ncu: 8.4 % of samples parked on the full barrier with thread 0 as producer).
            if (tid == PRODUCER_TID && s >= 1) {
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
                if (MODE == 2) {
                    const float c0 = fmaxf(v0, 0.f), c1 = fmaxf(v1, 0.f);
                    const int ja = 2 * jp, jb = 2 * jp + 1;
                    b2[MODE == 2 ? ja : 0] = fminf(b2[MODE == 2 ? ja : 0], fmaxf(c0, bs[ja]));
                    b2[MODE == 2 ? jb : 0] = fminf(b2[MODE == 2 ? jb : 0], fmaxf(c1, bs[jb]));
                    const bool p0 = c0 < bs[ja], p1 = c1 < bs[jb];
                    bs[ja] = fminf(bs[ja], c0); bi[ja] = p0 ? row : bi[ja];
                    bs[jb] = fminf(bs[jb], c1); bi[jb] = p1 ? row : bi[jb];
                    continue;
                }
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

    // ---- cross-thread reduction: the 16 threads sharing a column are 4 lanes (xor 8, xor 16) x 4 warps --------
    __syncthreads();
    const int wrow = warp >> 1;   // 0..3: which 4-ty band this warp covers
    if (MODE == 2) {
        float* red2 = reinterpret_cast<float*>(red + 4 * TILE);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            u64 key = make_key(bs[j], bi[j]);
            float sec = b2[MODE == 2 ? j : 0];
#pragma unroll
            for (int m = 8; m <= 16; m <<= 1) {
                const u64 ok = shfl_xor_u64(key, m);
                const float os = __shfl_xor_sync(0xffffffffu, sec, m);
                // second smallest of the union: the smaller "second" or the larger of the two minima
                sec = fminf(fminf(sec, os), fmaxf(__uint_as_float((uint32_t)(key >> 32)), __uint_as_float((uint32_t)(ok >> 32))));
                key = ok < key ? ok : key;
            }
            const int col = (j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4));
            if (lane < 8) { red[wrow * TILE + col] = key; red2[wrow * TILE + col] = sec; }
        }
        __syncthreads();
        if (tid < TILE) {
            u64 key = red[tid];
            float sec = red2[tid];
#pragma unroll
            for (int w = 1; w < 4; ++w) {
                const u64 ok = red[w * TILE + tid];
                sec = fminf(fminf(sec, red2[w * TILE + tid]),
                            fmaxf(__uint_as_float((uint32_t)(key >> 32)), __uint_as_float((uint32_t)(ok >> 32))));
                key = ok < key ? ok : key;
            }
            if (ct * TILE + tid < pd_n2) {
                // cross-CTA merge: the loser of every atomicMin exchange (incoming or displaced) feeds `best2`
                const u64 old = atomicMin(prm.best + ycol + tid, key);
                uint32_t loser = (uint32_t)((old < key ? key : old) >> 32);   // old == ~0: 0xffffffff, ignored by min
                const uint32_t secb = __float_as_uint(sec);
                loser = loser < secb ? loser : secb;
                atomicMin(prm.best2 + ycol + tid, loser);
            }
        }
    } else if (MODE == 0) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            u64 key = make_key(bs[j], bi[j]);
            u64 o = shfl_xor_u64(key, 8);
            key = o < key ? o : key;
            o = shfl_xor_u64(key, 16);
            key = o < key ? o : key;
            const int col = (j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4));
            if (lane < 8) red[wrow * TILE + col] = key;
        }
        __syncthreads();
        if (tid < TILE) {
            u64 m = red[tid];
#pragma unroll
            for (int w = 1; w < 4; ++w) { const u64 o = red[w * TILE + tid]; m = o < m ? o : m; }
            if (ct * TILE + tid < pd_n2) atomicMin(prm.best + ycol + tid, m);
        }
    } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            u64 mine[5], other[5];
#pragma unroll
            for (int r = 0; r < 5; ++r) mine[r] = l5[MODE == 1 ? j : 0][r];
#pragma unroll
            for (int r = 0; r < 5; ++r) other[r] = shfl_xor_u64(mine[r], 8);
#pragma unroll
            for (int r = 0; r < 5; ++r) insert5(mine, other[r]);
#pragma unroll
            for (int r = 0; r < 5; ++r) other[r] = shfl_xor_u64(mine[r], 16);
#pragma unroll
            for (int r = 0; r < 5; ++r) insert5(mine, other[r]);
            const int col = (j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4));
            if (lane < 8) {
#pragma unroll
                for (int r = 0; r < 5; ++r) red[((size_t)wrow * TILE + col) * 5 + r] = mine[r];
            }
        }
        __syncthreads();
        if (tid < TILE) {
            u64 m[5];
#pragma unroll
            for (int r = 0; r < 5; ++r) m[r] = red[(size_t)tid * 5 + r];
            for (int w = 1; w < 4; ++w)
#pragma unroll
                for (int r = 0; r < 5; ++r) insert5(m, red[((size_t)w * TILE + tid) * 5 + r]);
#pragma unroll
            for (int r = 0; r < 5; ++r) prm.part5[((size_t)item * TILE + tid) * 5 + r] = m[r];
        }
    }
}


// ---------------------------------------------------------------------------------------------------------------
// Gated kernel (the fast path of both solvers): same operand staging and FFMA2 main loop as nn_kernel, but the
// epilogue does NOT track minima per element.  Every column keeps a threshold tau_j = raw clamped d2 of its current
// K-th best candidate (K = 1: arg-min, K = 5: top-5); an element costs one packed add pair and one compare, and only
// the rare elements below tau are appended to a small per-column candidate list in shared memory.  After every row
// tile the lists are cut back to the exact K best by (sqrt_rn(d2), row) and tau is lowered.
//
// Exactness: rows arrive in ascending tile order, so a later row can only enter a column's top K when its distance is
// STRICTLY smaller than the current K-th distance (equal distances lose to the lower index already held), hence
// sqrt(d2') < sqrt(d2_K)  =>  d2' < d2_K = tau: the strict raw-value gate never drops a winner, and the cut itself
// orders by the exact (sqrt_rn, row) key.  Square roots are taken for list entries only.
// A list that overflows (more than CAP - K appends to one column within one tile: rows arriving sorted by decreasing
// distance) flags its column tile; the exact kernel (nn_kernel<0/1>) then redoes flagged tiles, so the result is
// exact for every input and fast for every realistic one.
// ---------------------------------------------------------------------------------------------------------------
template <int K> struct GatedCfg { static constexpr int CAP = (K == 1) ? 1 : 32; };   // K == 1 keeps no lists

template <int K, int CAP>
__device__ __forceinline__ void gated_append(int col, float v, uint32_t row, float* s_val, uint32_t* s_row, int* s_cnt,
                                             int* s_ovf) {
    const int slot = atomicAdd(&s_cnt[col], 1);
    if (slot < CAP) { s_val[slot * TILE + col] = fmaxf(v, 0.f); s_row[slot * TILE + col] = row; }
    else *s_ovf = 1;
}

// K = 1 only: a passing candidate is merged straight into the column's exact (sqrt_rn(d2), row) key with a 64-bit
// shared-memory atomicMin (no list, no cut, nothing to overflow) and lowers the column's threshold immediately.  Rows of one tile are processed in no
// particular index order, so the new threshold must still admit every value whose sqrt_rn equals the candidate's
// (a lower-index row with the same distance wins the tie): tau' = up((s + 1 ulp)^2) with s = sqrt_rn(c) is above the
// whole rounding class of s, and any x >= tau' has sqrt_rn(x) > s.
__device__ __forceinline__ void gated_best1(int col, float v, uint32_t row, float* s_tau, u64* s_best) {
    const float sr = __fsqrt_rn(fmaxf(v, 0.f));
    atomicMin(s_best + col, make_key(sr, row));
    const float su = __uint_as_float(__float_as_uint(sr) + 1u);
    atomicMin(reinterpret_cast<unsigned*>(s_tau) + col, __float_as_uint(__fmul_ru(su, su)));
}

// one thread per column: cut the list back to the K best by (sqrt_rn(d2), row), lower tau
template <int K, int CAP>
__device__ __forceinline__ void gated_compact(int col, float* s_val, uint32_t* s_row, int* s_cnt, float* s_tau, int& kept) {
    int n = s_cnt[col];
    if (n == kept) return;                 // nothing appended since the last cut
    if (n > CAP) n = CAP;                  // overflow: already flagged by the appender
    // entries [0, kept) are the sorted survivors of the last cut: they seed the list as they are
    u64 key[K]; float raw[K];
#pragma unroll
    for (int r = 0; r < K; ++r) {
        const bool have = r < kept;
        const float rv = have ? s_val[r * TILE + col] : __int_as_float(0x7f800000);
        raw[r] = rv;
        key[r] = have ? make_key(__fsqrt_rn(rv), s_row[r * TILE + col]) : ~0ull;
    }
    for (int e = kept; e < n; ++e) {
        float rv = s_val[e * TILE + col];
        u64 kk = make_key(__fsqrt_rn(rv), s_row[e * TILE + col]);
#pragma unroll
        for (int r = 0; r < K; ++r) {
            const bool p = kk < key[r];
            const u64 tk = key[r]; const float tr = raw[r];
            key[r] = p ? kk : tk; raw[r] = p ? rv : tr;
            kk = p ? tk : kk; rv = p ? tr : rv;
        }
    }
    const int m = n < K ? n : K;
#pragma unroll
    for (int r = 0; r < K; ++r)
        if (r < m) { s_val[r * TILE + col] = raw[r]; s_row[r * TILE + col] = (uint32_t)(key[r] & 0xffffffffull); }
    s_cnt[col] = m;
    kept = m;
    // tau = top of the sqrt_rn rounding class of the K-th best, not its raw value: the first tile's bootstrap walks the
    // rows in interleaved order with a cut between steps, so a LOWER-index row with the same rounded distance as the
    // held K-th best can still arrive and must pass the gate (the (distance, index) order then prefers it)
    if (m == K) {
        const float sk = __fsqrt_rn(raw[K - 1]);
        const float su = __uint_as_float(__float_as_uint(sk) + 1u);
        s_tau[col] = __fmul_ru(su, su);
    }
}

// NXP: L1 prefetch of the row norms at tile start, load in the epilogue.  FUSE (K == 1): the last CTA of a column tile
// writes the int64 map itself (small problems, where a finalise launch is a tenth of the call); a separate instance so
// that the plain kernel's register allocation — tuned at the 128-register cap — stays exactly as it was.
template <int K, int KC, int NSTAGE, int NXP, int FUSE = 0>
__global__ void __launch_bounds__(NN_THREADS, 2)
nn_gated_kernel(const __grid_constant__ CUtensorMap mapX, const __grid_constant__ CUtensorMap mapY, const NNParams prm) {
    constexpr int CAP = GatedCfg<K>::CAP;
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    const int k = prm.k;
    const int nkc = (k + KC - 1) / KC;
    float* sB = reinterpret_cast<float*>(smem_raw);                 // [nkc*KC][TILE]
    float* sA = sB + (size_t)nkc * KC * TILE;                        // NSTAGE x [KC][TILE]
    uint64_t* bars = reinterpret_cast<uint64_t*>(sA + (size_t)NSTAGE * KC * TILE);
    uint64_t* fullB = bars;
    uint64_t* full = bars + 1;
    uint64_t* empty = bars + 1 + NSTAGE;
    float* s_tau = reinterpret_cast<float*>(bars + 16);              // [TILE]
    int* s_cnt = reinterpret_cast<int*>(s_tau + TILE);               // [TILE]
    float* s_val = reinterpret_cast<float*>(s_cnt + TILE);           // [CAP][TILE] raw clamped d2
    uint32_t* s_row = reinterpret_cast<uint32_t*>(s_val + CAP * TILE);   // [CAP][TILE]
    int* s_ovf = reinterpret_cast<int*>(s_row + CAP * TILE);
    u64* s_best = reinterpret_cast<u64*>(s_cnt);                     // K == 1: [TILE] exact keys (lists unused)

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // Lane layout: half-warp = 2 (row groups) x 8 (column groups).  Tried and measured slower on B200: 16 x 2 with
    // warp-private columns (8 instead of 12 shared-memory wavefronts per k-step, tools/lds_pattern.cu) — arg-min
    // 46.0 vs 44.7 ms at 100k x 100k x 128, top-5 3.9 vs 3.1 ms at 50k x 20k x 50.
    const int tx = ((warp & 1) << 3) | (lane & 7);
    const int ty = ((warp >> 1) << 2) | ((lane >> 4) << 1) | ((lane >> 3) & 1);

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
    const float INF = __int_as_float(0x7f800000);
    if (prm.ncols && prm.ncols[pdp->ytile0 + ct] == 0) return;   // decided by the tensor-pipe screen (uniform exit)

    if (tid == 0) {
        tma_prefetch_desc(&mapX);
        tma_prefetch_desc(&mapY);
        mbar_init(fullB, 1);
        for (int s = 0; s < NSTAGE; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], NN_THREADS / 32); }
        fence_barrier_init();
        *s_ovf = 0;
    }
    if (tid < TILE) {
        s_tau[tid] = INF;
        if (K == 1) s_best[tid] = ~0ull; else s_cnt[tid] = 0;
    }
    __syncthreads();
    // Small-problem instance: launched with programmatic stream serialization behind the preparation kernel, so the
    // launch latency and the prologue above overlap that kernel's tail; nothing it wrote is touched before this wait
    // (a no-op when the launch carries no such attribute).
    if (FUSE) asm volatile("griddepcontrol.wait;" ::: "memory");

    auto issue = [&](int s) {
        const int slot = s % NSTAGE, t = s / nkc, kc = s - t * nkc;
        mbar_arrive_expect_tx(&full[slot], KC * TILE * sizeof(float));
        tma_load_2d(sA + (size_t)slot * KC * TILE, &mapX, xcol + (t_begin + t) * TILE, kc * KC, &full[slot]);
    };
    if (tid == PRODUCER_TID) {
        mbar_arrive_expect_tx(fullB, (uint32_t)(nkc * KC * TILE * sizeof(float)));
        for (int kc = 0; kc < nkc; ++kc) tma_load_2d(sB + (size_t)kc * KC * TILE, &mapY, ycol, kc * KC, fullB);
        for (int s = 0; s < NSTAGE && s < total_stages; ++s) issue(s);
    }

    u64 ny2[4];
    {
        const float4 n0 = __ldg(reinterpret_cast<const float4*>(prm.ny + ycol + tx * 4));
        const float4 n1 = __ldg(reinterpret_cast<const float4*>(prm.ny + ycol + 64 + tx * 4));
        ny2[0] = pack2(n0.x, n0.y); ny2[1] = pack2(n0.z, n0.w);
        ny2[2] = pack2(n1.x, n1.y); ny2[3] = pack2(n1.z, n1.w);
    }
    int kept = 0;                          // threads < TILE: entries their column held after its last cut

    mbar_wait(fullB, 0);

    int s = 0;
    for (int t = 0; t < ntiles; ++t) {
        const int rowbase = (t_begin + t) * TILE;
        // row norms: pulled into L1 now, loaded in the epilogue (ptxas sinks a plain early load down to its use to save
        // registers, where it then costs an L2 round trip per tile)
        float4 nxa, nxb;
        if (NXP) {
            prefetch_l1(prm.nx + xcol + rowbase + ty * 4);
            prefetch_l1(prm.nx + xcol + rowbase + 64 + ty * 4);
        } else {
            nxa = __ldg(reinterpret_cast<const float4*>(prm.nx + xcol + rowbase + ty * 4));
            nxb = __ldg(reinterpret_cast<const float4*>(prm.nx + xcol + rowbase + 64 + ty * 4));
        }
        u64 acc[8][4];
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = 0ull;

        for (int kc = 0; kc < nkc; ++kc, ++s) {
            const int slot = s % NSTAGE;
            const uint32_t ph = (uint32_t)((s / NSTAGE) & 1);
            if (tid == PRODUCER_TID && s >= 1) {
                const int sn = s - 1 + NSTAGE;
                if (sn < total_stages) {
                    mbar_wait(&empty[(s - 1) % NSTAGE], (uint32_t)(((s - 1) / NSTAGE) & 1));
                    issue(sn);
                }
            }
            mbar_wait(&full[slot], ph);
            const float* A = sA + (size_t)slot * KC * TILE + ty * 4;
            const float* B = sB + (size_t)kc * KC * TILE + tx * 4;
            auto step = [&](int kk) {
                const float4 a0 = *reinterpret_cast<const float4*>(A + kk * TILE);
                const float4 a1 = *reinterpret_cast<const float4*>(A + kk * TILE + 64);
                const ulonglong2 b0 = *reinterpret_cast<const ulonglong2*>(B + kk * TILE);
                const ulonglong2 b1 = *reinterpret_cast<const ulonglong2*>(B + kk * TILE + 64);
                const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
                const u64 bp[4] = {b0.x, b0.y, b1.x, b1.y};
#pragma unroll
                for (int j = 0; j < 4; ++j)
#pragma unroll
                    for (int i = 0; i < 8; ++i) fma2v(acc[i][j], pack2(av[i], av[i]), bp[j]);
            };
            // Every stage runs the full unrolled KC steps: the k-rows past k in the last stage are TMA zero fill on BOTH
            // operands, and fma(0, 0, acc) leaves acc as it is (a -0 accumulator would become +0, which the following
            // "+ nx" with nx >= +0 maps to the same sum).  The rolled tail loop this replaces ran a third of the
            // k = 50 work (stages 17 + 17 + 16) at 47 instead of 39 instructions per k-step (ncu, 256 pairs).
#pragma unroll
            for (int kk = 0; kk < KC; ++kk) step(kk);
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[slot]);
        }

        if (NXP) {
            nxa = __ldg(reinterpret_cast<const float4*>(prm.nx + xcol + rowbase + ty * 4));
            nxb = __ldg(reinterpret_cast<const float4*>(prm.nx + xcol + rowbase + 64 + ty * 4));
        }
        const float nxv[8] = {nxa.x, nxa.y, nxa.z, nxa.w, nxb.x, nxb.y, nxb.z, nxb.w};
        // thresholds of this thread's 8 columns: read from shared memory per tile (K == 5; they only change at a cut) or
        // per row (K == 1) instead of living in registers across the main loop — at the 128-register cap eight more
        // live values there cost the top-5 instance one register move per 5 packed FMAs and a spill (ncu, 256 pairs)
        float tau[8];
        if (K != 1) {
            const float4 t0 = *reinterpret_cast<const float4*>(s_tau + tx * 4);
            const float4 t1 = *reinterpret_cast<const float4*>(s_tau + 64 + tx * 4);
            tau[0] = t0.x; tau[1] = t0.y; tau[2] = t0.z; tau[3] = t0.w;
            tau[4] = t1.x; tau[5] = t1.y; tau[6] = t1.z; tau[7] = t1.w;
        }
        // row i of this thread: values (acc + nx) + ny for its 8 columns, gate against tau, append the rare passes
        auto gate_row = [&](int i) {
            const uint32_t row = (uint32_t)(rowbase + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4)));
            const u64 nn = pack2(nxv[i], nxv[i]);
            float v[8];
#pragma unroll
            for (int jp = 0; jp < 4; ++jp) unpack2(add2(add2(acc[i][jp], nn), ny2[jp]), v[2 * jp], v[2 * jp + 1]);
            if (K == 1) {
                // arg-min: tau is re-read from shared memory for every row, because appenders lower it at once
                // (see gated_append_lower) — a stale threshold lets (n_new / n_seen)^m columns collect m appends
                const float4 t0 = *reinterpret_cast<const float4*>(s_tau + tx * 4);
                const float4 t1 = *reinterpret_cast<const float4*>(s_tau + 64 + tx * 4);
                tau[0] = t0.x; tau[1] = t0.y; tau[2] = t0.z; tau[3] = t0.w;
                tau[4] = t1.x; tau[5] = t1.y; tau[6] = t1.z; tau[7] = t1.w;
            }
            bool any = false;
#pragma unroll
            for (int j = 0; j < 8; ++j) any |= (v[j] < tau[j]);
            if (any) {
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    if (v[j] < tau[j]) {
                        const int col = j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4);
                        if (K == 1) gated_best1(col, v[j], row, s_tau, s_best);
                        else gated_append<K, CAP>(col, v[j], row, s_val, s_row, s_cnt, s_ovf);
                    }
            }
        };
        auto cut_and_reload = [&]() {
            __syncthreads();
            if (tid < TILE) gated_compact<K, CAP>(tid, s_val, s_row, s_cnt, s_tau, kept);
            __syncthreads();
            const float4 t0 = *reinterpret_cast<const float4*>(s_tau + tx * 4);
            const float4 t1 = *reinterpret_cast<const float4*>(s_tau + 64 + tx * 4);
            tau[0] = t0.x; tau[1] = t0.y; tau[2] = t0.z; tau[3] = t0.w;
            tau[4] = t1.x; tau[5] = t1.y; tau[6] = t1.z; tau[7] = t1.w;
        };

        if (t == 0) {
            // bootstrap (tau = +inf): bounded appends per cut so the first tile cannot overflow
            if (K == 1 && FUSE) {
                // Small problems run 2-3 row tiles per CTA, so this first tile is a third of the work: one square root
                // per column instead of one per element.  cm = the column's raw minimum over the 32 rows of the 4 lanes
                // that share it, s = sqrt_rn(cm), tau' = up((s + 1ulp)^2): a row ties or beats s exactly when its raw
                // value is below tau' (then sqrt_rn(c) <= s, and c >= cm gives sqrt_rn(c) >= s), and the lowest such
                // row index is the reference's winner.  (Tried in the plain instance at 100k x 100k x 128: the extra
                // live values cost the main loop registers there; this instance only serves small problems.)
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    float c[8];
                    float cm = INF;
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        float v0, v1;
                        unpack2(add2(add2(acc[i][j >> 1], pack2(nxv[i], nxv[i])), ny2[j >> 1]), v0, v1);
                        c[i] = fmaxf((j & 1) ? v1 : v0, 0.f);
                        cm = fminf(cm, c[i]);
                    }
                    cm = fminf(cm, __shfl_xor_sync(0xffffffffu, cm, 8));
                    cm = fminf(cm, __shfl_xor_sync(0xffffffffu, cm, 16));
                    const float sr = __fsqrt_rn(cm);
                    const float su = __uint_as_float(__float_as_uint(sr) + 1u);
                    const float tn = __fmul_ru(su, su);
                    uint32_t bi = IDX_NONE;
#pragma unroll
                    for (int i = 7; i >= 0; --i)
                        if (c[i] < tn) bi = (uint32_t)(rowbase + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4)));
                    bi = min(bi, __shfl_xor_sync(0xffffffffu, bi, 8));
                    bi = min(bi, __shfl_xor_sync(0xffffffffu, bi, 16));
                    if ((lane & 24) == 0 && bi != IDX_NONE) {
                        const int col = j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4);
                        atomicMin(s_best + col, make_key(sr, bi));
                        atomicMin(reinterpret_cast<unsigned*>(s_tau) + col, __float_as_uint(tn));
                    }
                }
            } else if (K == 1) {
                // exact local arg-min over this thread's 8 rows, merged across the 4 lanes sharing a column
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    float bd = INF, braw = INF; uint32_t bi = IDX_NONE;
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        float v0, v1;
                        unpack2(add2(add2(acc[i][j >> 1], pack2(nxv[i], nxv[i])), ny2[j >> 1]), v0, v1);
                        const float c = fmaxf((j & 1) ? v1 : v0, 0.f);
                        const float d = __fsqrt_rn(c);
                        const bool p = d < bd;
                        bd = p ? d : bd; braw = p ? c : braw;
                        bi = p ? (uint32_t)(rowbase + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4))) : bi;
                    }
                    u64 key = make_key(bd, bi);
#pragma unroll
                    for (int m = 8; m <= 16; m <<= 1) {
                        const u64 ok = shfl_xor_u64(key, m);
                        const float oraw = __shfl_xor_sync(0xffffffffu, braw, m);
                        const bool p = ok < key;
                        key = p ? ok : key; braw = p ? oraw : braw;
                    }
                    const uint32_t wi = (uint32_t)(key & 0xffffffffull);
                    if ((lane & 24) == 0 && wi != IDX_NONE)
                        gated_best1(j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4), braw, wi, s_tau, s_best);
                }
            } else {
                // one row of every lane per cut: at most 16 appends per column on top of the K kept
                for (int step_i = 0; step_i < 8; ++step_i) {
#pragma unroll
                    for (int i = 0; i < 8; ++i)
                        if (i == step_i) gate_row(i);
                    cut_and_reload();
                }
            }
        } else if (K != 1 && t == 1) {
            // the threshold is still loose (K-th best of 128 rows): two cuts keep the (n_new / n_seen)^m tail of
            // the append count away from CAP
#pragma unroll
            for (int i = 0; i < 4; ++i) gate_row(i);
            cut_and_reload();
#pragma unroll
            for (int i = 4; i < 8; ++i) gate_row(i);
            cut_and_reload();
        } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) gate_row(i);
            // K == 1 runs without any CTA-wide synchronisation.  K == 5 cuts (two CTA barriers) after every tile while
            // the threshold is loose, then after every 2nd / 4th tile: the appends of an interval of n rows after r
            // seen rows exceed m with probability ~ C(m+4,4) (n/(r+n))^m, < 1e-9 for m = CAP - K at n/(r+n) <= 1/3.
            if (K != 1 && (t < 4 || t == ntiles - 1 || (t < 8 ? (t & 1) == 1 : (t & 3) == 3))) cut_and_reload();
        }
    }

    // ---- results: the lists hold the exact K best of this CTA's rows, sorted by (distance, row) ------------------
    if (K == 1) {
        __syncthreads();
        if (tid < TILE && ct * TILE + tid < pd_n2) atomicMin(prm.best + ycol + tid, s_best[tid]);
        if (FUSE) {
            // fused finalise (saves a launch): the last of this column tile's row-split CTAs to arrive turns the merged
            // keys into the int64 map.  Release: every thread's atomicMin, fence, barrier, then one counter increment;
            // acquire: the CTA that reads nsplit-1 fences and reads the keys from L2, where the atomics were performed.
            // The pair fields are looked up again here rather than kept live across the main loop (128-register cap).
            __threadfence();
            __syncthreads();
            const int it2 = blockIdx.x;
            const PairDev* q = prm.pairs + find_pair(it2, prm.pairs, prm.npairs, prm.uniform_items, &PairDev::item0);
            const int ns2 = q->nsplit, ct2 = (it2 - q->item0) / ns2;
            if (tid == 0) *s_ovf = (atomicAdd(prm.tilecnt + q->ytile0 + ct2, 1) == ns2 - 1);
            __syncthreads();
            const int j = ct2 * TILE + tid, n2 = q->n2;
            if (*s_ovf && tid < TILE && j < n2) {
                __threadfence();
                uint32_t idx = (uint32_t)(__ldcg(prm.best + q->y_col0 + j) & 0xffffffffull);
                if (idx >= IDX_NONE) idx = 0;  // all distances were +inf/NaN: out of contract, mirror argmin's first index
                int64_t* o = prm.out + q->out_off;
                o[j] = (int64_t)idx;
                if (!(prm.flags & F2P_ROW0_ONLY)) o[(int64_t)n2 + j] = row1_value(j, prm.flags);
            }
        }
    } else {
        if (tid < TILE) {
#pragma unroll
            for (int r = 0; r < K; ++r)
                prm.part5[((size_t)item * TILE + tid) * 5 + r] =
                    r < kept ? make_key(__fsqrt_rn(s_val[r * TILE + tid]), s_row[r * TILE + tid]) : ~0ull;
        }
        if (tid == 0 && *s_ovf) prm.tileflag[pdp->ytile0 + ct] = 1;
    }
}

// ---- finalize: keys -> int64 index maps ---------------------------------------------------------------
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
    if (idx >= IDX_NONE) idx = 0;  // all distances were +inf/NaN: out of contract, mirror argmin's first index
    int64_t* o = out + pd->out_off;
    o[j] = (int64_t)idx;
    if (!(flags & F2P_ROW0_ONLY)) o[(int64_t)n2 + j] = row1_value(j, flags);
}

// Merges the nsplit partial top-5 lists of every column and writes nn_query's [2, 5*n2] candidate layout
// (spacial_filtering.py:36-38): column 5j+r = (r-th nearest CAD row of scene point j, j).
__global__ void __launch_bounds__(TILE)
finalize_top5_kernel(const u64* __restrict__ part5, const PairDev* __restrict__ pairs, int npairs, int uniform_tiles,
                     unsigned flags, int64_t* __restrict__ cand, int cand_is_out_layout, const int* __restrict__ tc_only) {
    if (tc_only && tc_only[blockIdx.x] == 0) return;   // this tile's candidates were written by the tensor-pipe path
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

// ---- two-phase arg-min support kernels ----------------------------------------------------------------
// Initialises the per-column merge state of an arg-min pass (and the per-pair re-check counters).
__global__ void argmin_init_kernel(u64* best, uint32_t* best2, u64* best_exact, size_t n, int* nflag, int npairs,
                                   int* tileflag) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        best[i] = ~0ull;
        if (best2) best2[i] = 0x7f800000u;
        if (best_exact) best_exact[i] = ~0ull;
        if (tileflag && i < n / TILE) tileflag[i] = 0;
    }
    if (nflag && i < (size_t)npairs) nflag[i] = 0;
}

// After the raw pass: a column is decided when its two smallest clamped d2 round to different sqrt_rn values
// (sqrt is monotone, so every other row is farther).  Otherwise exact ties in distance are possible and the lowest
// index among them may not be the raw minimum: the column joins the pair's re-check list.  Also writes row 1.
__global__ void __launch_bounds__(TILE)
finalize_raw_kernel(const u64* __restrict__ best, const uint32_t* __restrict__ best2, const PairDev* __restrict__ pairs,
                    int npairs, int uniform_tiles, unsigned flags, int64_t* __restrict__ out, int* __restrict__ nflag,
                    int* __restrict__ flagged) {
    const int pi = find_pair(blockIdx.x, pairs, npairs, uniform_tiles, &PairDev::ytile0);
    const PairDev* pd = pairs + pi;
    const int j = (blockIdx.x - pd->ytile0) * TILE + threadIdx.x;
    const int n2 = pd->n2;
    if (j >= n2) return;
    const u64 key = best[pd->y_col0 + j];
    uint32_t idx = (uint32_t)(key & 0xffffffffull);
    const float m1 = __uint_as_float((uint32_t)(key >> 32));
    const float m2 = __uint_as_float(best2[pd->y_col0 + j]);
    int64_t* o = out + pd->out_off;
    if (!(flags & F2P_ROW0_ONLY)) o[(int64_t)n2 + j] = row1_value(j, flags);
    if (idx == IDX_NONE) { o[j] = 0; return; }              // all distances +inf/NaN: out of contract
    if (__fsqrt_rn(m2) == __fsqrt_rn(m1)) {                  // possible exact tie after the square root
        const int f = atomicAdd(nflag + pi, 1);
        flagged[pd->y_col0 + f] = j;
    } else {
        o[j] = (int64_t)idx;
    }
}

// Compacts the flagged columns of every pair into the re-check operand (same k-major layout as yT).
__global__ void __launch_bounds__(TILE)
gather_flagged_kernel(const float* __restrict__ yT, const float* __restrict__ ny, const PairDev* __restrict__ pairs,
                      int npairs, int uniform_tiles, int k, int64_t ld2_total, const int* __restrict__ nflag,
                      const int* __restrict__ flagged, float* __restrict__ yTf, float* __restrict__ nyf) {
    const int pi = find_pair(blockIdx.x, pairs, npairs, uniform_tiles, &PairDev::ytile0);
    const PairDev* pd = pairs + pi;
    const int f = (blockIdx.x - pd->ytile0) * TILE + threadIdx.x;
    const int nf = nflag[pi];
    if ((blockIdx.x - pd->ytile0) * TILE >= nf) return;
    const int64_t dst = (int64_t)pd->y_col0 + f;
    if (f < nf) {
        const int64_t src = (int64_t)pd->y_col0 + flagged[pd->y_col0 + f];
        for (int c = 0; c < k; ++c) yTf[(int64_t)c * ld2_total + dst] = yT[(int64_t)c * ld2_total + src];
        nyf[dst] = ny[src];
    } else {   // tail of the last compacted tile: defined values, results are never read back
        for (int c = 0; c < k; ++c) yTf[(int64_t)c * ld2_total + dst] = 0.f;
        nyf[dst] = 0.f;
    }
}

__global__ void __launch_bounds__(TILE)
finalize_recheck_kernel(const u64* __restrict__ best_exact, const PairDev* __restrict__ pairs, int npairs,
                        int uniform_tiles, const int* __restrict__ nflag, const int* __restrict__ flagged,
                        int64_t* __restrict__ out) {
    const int pi = find_pair(blockIdx.x, pairs, npairs, uniform_tiles, &PairDev::ytile0);
    const PairDev* pd = pairs + pi;
    const int f = (blockIdx.x - pd->ytile0) * TILE + threadIdx.x;
    if (f >= nflag[pi]) return;
    const int j = flagged[pd->y_col0 + f];
    out[pd->out_off + j] = (int64_t)(uint32_t)(best_exact[pd->y_col0 + f] & 0xffffffffull);
}

// ---- host launchers -----------------------------------------------------------------------------------
static size_t nn_smem_bytes_v(int k, int mode, int kc, int nst) {
    const int nkc = (k + kc - 1) / kc;
    size_t b = ((size_t)nkc * kc * TILE + (size_t)nst * kc * TILE) * sizeof(float) + 16 * sizeof(uint64_t);
    b += (mode != 1 ? (size_t)8 * TILE : (size_t)4 * TILE * 5) * sizeof(u64);
    return b;
}
size_t nn_smem_bytes(int k, int mode) { return nn_smem_bytes_v(k, mode, KC, NSTAGE); }
size_t top5_ws_bytes(int64_t total_items, int64_t total_ytiles) {
    return sizeof(u64) * (size_t)total_items * TILE * 5 + sizeof(int) * (size_t)total_ytiles;
}

template <int MODE>
static int launch_mode(const float* xT, int64_t ld1_total, const float* yT, int64_t ld2_total, int k, const NNParams& prm,
                       int total_items, cudaStream_t stream) {
    CUtensorMap mapX, mapY;
    int rc = make_tensor_map_2d(&mapX, xT, (uint64_t)ld1_total, (uint64_t)k, (uint64_t)ld1_total, TILE, KC);
    if (rc) return rc;
    rc = make_tensor_map_2d(&mapY, yT, (uint64_t)ld2_total, (uint64_t)k, (uint64_t)ld2_total, TILE, KC);
    if (rc) return rc;
    const size_t smem = nn_smem_bytes_v(k, MODE, KC, NSTAGE);
    F2P_CUDA(cudaFuncSetAttribute(nn_kernel<MODE, KC, NSTAGE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    nn_kernel<MODE, KC, NSTAGE><<<total_items, NN_THREADS, smem, stream>>>(mapX, mapY, prm);
    F2P_CUDA(cudaGetLastError());
    return 0;
}

template <int K>
static size_t gated_smem_bytes(int k, int kc, int nstage) {
    const int nkc = (k + kc - 1) / kc;
    return ((size_t)nkc * kc * TILE + (size_t)nstage * kc * TILE) * sizeof(float) + 16 * sizeof(uint64_t) +
           2 * TILE * sizeof(int) + (size_t)GatedCfg<K>::CAP * TILE * 8 + 16;
}

// k-rows per TMA stage of the gated kernels: 15, 16 or 17, whichever pads k least (then fewest stages), so that every
// stage is full and runs the unrolled loop — k = 50 -> 3 x 17 instead of 3 x 16 + a 2-step tail stage that costs a
// whole barrier round trip and a rolled loop; k = 30 -> 2 x 15; k = 64 / 128 -> 16.  The padded k-rows are TMA
// zero-fill on both operands (fma(0, 0, acc) == acc: acc is never -0).  F2P_GATED_KC overrides (tools/nn_time.py).
static int choose_gated_kc(int k, int total_items) {
    static const int forced = [] { const char* e = getenv("F2P_GATED_KC"); return e ? atoi(e) : 0; }();
    if (forced == 15 || forced == 16 || forced == 17 || forced == 25 || (forced == 50 && k <= 50) || forced == 32) return forced;
    int best = 16, best_pad = 1 << 30, best_n = 1 << 30;
    // 25 / 50 (k <= 50 only, where two top-5 CTAs still fit one SM's shared memory): k = 50 runs as ONE 50-row stage per
    // tile on a 2-deep ring (same shared memory as 4 x 12.5-row-equivalents): no padded step, one barrier round trip per
    // tile.  Measured per 256 pairs, top-5: 3 x 17 6.21 ms, 2 x 25 5.98 ms, 1 x 50 5.92 ms.
    // 32: 32-row stages on a 2-deep ring, the shared memory of 16-row stages on the 4-deep ring — k = 128 runs 4 barrier
    // round trips per tile instead of 8.  Measured: 100k x 100k x 128 (12.5k work items) 44.40 -> 43.89 ms, but
    // 50k x 20k x 128 (3.8k items) 4.79 -> 5.00 ms and 50k x 20k x 64 3.01 -> 3.46 ms, so it is taken for k >= 128 and
    // long grids only (tools/kc32.sh, tools/kc64.sh).
    const bool deep = k >= 128 && total_items >= 8192;
    const int cand[5] = {16, 17, 15, k <= 50 ? 25 : 32, 50};
    for (int c = 0; c < (k <= 50 ? 5 : (deep ? 4 : 3)); ++c) {
        const int n = (k + cand[c] - 1) / cand[c], pad = n * cand[c] - k;
        if (pad < best_pad || (pad == best_pad && n < best_n)) { best = cand[c]; best_pad = pad; best_n = n; }
    }
    return best;
}

template <int K, int KCV, int NST = NSTAGE>
static int launch_gated_kc(const float* xT, int64_t ld1_total, const float* yT, int64_t ld2_total, int k, const NNParams& prm,
                           int total_items, cudaStream_t stream) {
    CUtensorMap mapX, mapY;
    int rc = make_tensor_map_2d(&mapX, xT, (uint64_t)ld1_total, (uint64_t)k, (uint64_t)ld1_total, TILE, KCV);
    if (rc) return rc;
    rc = make_tensor_map_2d(&mapY, yT, (uint64_t)ld2_total, (uint64_t)k, (uint64_t)ld2_total, TILE, KCV);
    if (rc) return rc;
    // Row-norm L1 prefetch: +2 % for the arg-min kernel at 100k x 100k x 128, slightly negative for top-5 (registers).
    // F2P_GATED_NXP overrides (tools/nn_time.py).
    const char* ep = getenv("F2P_GATED_NXP");
    const int nxp = ep ? atoi(ep) : (K == 1 ? 1 : 0);
    const size_t smem = gated_smem_bytes<K>(k, KCV, NST);
    if (K == 1 && prm.tilecnt) {
        F2P_CUDA(cudaFuncSetAttribute(nn_gated_kernel<1, KCV, NST, 1, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)total_items); cfg.blockDim = dim3(NN_THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = stream;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        F2P_CUDA(cudaLaunchKernelEx(&cfg, nn_gated_kernel<1, KCV, NST, 1, 1>, mapX, mapY, prm));
    } else if (nxp) {
        F2P_CUDA(cudaFuncSetAttribute(nn_gated_kernel<K, KCV, NST, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        nn_gated_kernel<K, KCV, NST, 1><<<total_items, NN_THREADS, smem, stream>>>(mapX, mapY, prm);
    } else {
        F2P_CUDA(cudaFuncSetAttribute(nn_gated_kernel<K, KCV, NST, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        nn_gated_kernel<K, KCV, NST, 0><<<total_items, NN_THREADS, smem, stream>>>(mapX, mapY, prm);
    }
    F2P_CUDA(cudaGetLastError());
    return 0;
}

template <int K>
static int launch_gated(const float* xT, int64_t ld1_total, const float* yT, int64_t ld2_total, int k, const NNParams& prm,
                        int total_items, cudaStream_t stream) {
    switch (choose_gated_kc(k, total_items)) {
        case 15: return launch_gated_kc<K, 15>(xT, ld1_total, yT, ld2_total, k, prm, total_items, stream);
        case 17: return launch_gated_kc<K, 17>(xT, ld1_total, yT, ld2_total, k, prm, total_items, stream);
        case 25: return launch_gated_kc<K, 25>(xT, ld1_total, yT, ld2_total, k, prm, total_items, stream);
        case 50: return launch_gated_kc<K, 50, 2>(xT, ld1_total, yT, ld2_total, k, prm, total_items, stream);
        case 32: return launch_gated_kc<K, 32, 2>(xT, ld1_total, yT, ld2_total, k, prm, total_items, stream);
        default: return launch_gated_kc<K, 16>(xT, ld1_total, yT, ld2_total, k, prm, total_items, stream);
    }
}

size_t argmin_ws_bytes(int64_t ld2_total, int k, int npairs) {
    const size_t n = (size_t)ld2_total;
    size_t b = 0;
    b += n * sizeof(u64);                       // best
    b += n * sizeof(uint32_t);                  // best2
    b += n * sizeof(u64);                       // best_exact
    b += n * sizeof(int);                       // flagged
    b += n * sizeof(float);                     // nyf
    b += ((size_t)(npairs > 0 ? npairs : 1) * sizeof(int) + 255) & ~(size_t)255;   // nflag
    b += ((n / TILE + 1) * sizeof(int) + 255) & ~(size_t)255;                        // tilecnt
    b += (size_t)k * n * sizeof(float);         // yTf
    return (b + 255) & ~(size_t)255;
}

void argmin_ws_carve(ArgminWs& w, unsigned char* base, int64_t ld2_total, int k, int npairs) {
    const size_t n = (size_t)ld2_total;
    unsigned char* p = base;
    w.best = reinterpret_cast<u64*>(p); p += n * sizeof(u64);
    w.best_exact = reinterpret_cast<u64*>(p); p += n * sizeof(u64);
    w.yTf = reinterpret_cast<float*>(p); p += (size_t)k * n * sizeof(float);
    w.nyf = reinterpret_cast<float*>(p); p += n * sizeof(float);
    w.best2 = reinterpret_cast<uint32_t*>(p); p += n * sizeof(uint32_t);
    w.flagged = reinterpret_cast<int*>(p); p += n * sizeof(int);
    w.nflag = reinterpret_cast<int*>(p); p += ((size_t)(npairs > 0 ? npairs : 1) * sizeof(int) + 255) & ~(size_t)255;
    w.tilecnt = reinterpret_cast<int*>(p);
}

// Path switch (tests): F2P_NN_PATH = "exact" (nn_kernel<0/1> alone), "two_phase" (arg-min: raw pass + re-check of
// sqrt-ambiguous columns), "gated" (default: threshold-gated kernel + exact fallback for overflowed column tiles).
enum NNPath { PATH_GATED = 0, PATH_EXACT = 1, PATH_TWO_PHASE = 2 };
static NNPath nn_path() {
    const char* e = getenv("F2P_NN_PATH");
    if (e && e[0] == 'e') return PATH_EXACT;
    if (e && e[0] == 't' && e[1] == 'w') return PATH_TWO_PHASE;
    return PATH_GATED;
}
// "tc": tensor-pipe screen in front of the gated kernels regardless of size; "gated" / "exact" / "two_phase": never;
// unset: by size.  Every choice gives the same (exact) result.
bool tc_enabled(double flops, int k, int top5) {
    if (tc_kp(k) < 32 || !tc_fits(k, top5)) return false;   // the packed rows must fill one 128-byte swizzle row
    const char* e = getenv("F2P_NN_PATH");
    if (e && e[0] == 't' && e[1] == 'c') return true;
    if (e && e[0]) return false;
    return flops >= TC_MIN_FLOPS;
}

constexpr int FUSE_MAX_ITEMS = 16 * 148;   // ~8 waves of 2 CTAs per SM on a B200

int launch_argmin(const float* xT, int64_t ld1_total, const float* nx, const float* yT, int64_t ld2_total,
                  const float* ny, int k, const PairDev* pairs_dev, int npairs, int uniform_items, int total_items,
                  int uniform_ytiles, int total_ytiles, double flops, unsigned flags, int64_t* out, const ArgminWs& w,
                  cudaStream_t stream, bool pre_init, const TCPlan* tc) {
    if (total_items <= 0 || total_ytiles <= 0) return 0;
    NNParams prm;
    prm.pairs = pairs_dev; prm.npairs = npairs; prm.uniform_items = uniform_items; prm.k = k;
    prm.nx = nx; prm.ny = ny; prm.best = w.best; prm.part5 = nullptr; prm.best2 = nullptr; prm.ncols = nullptr;
    prm.tileflag = nullptr; prm.tilecnt = nullptr; prm.out = out; prm.flags = flags;
    const NNPath path = nn_path();
    (void)flops;
    const unsigned iblocks = (unsigned)((ld2_total + 255) / 256);
    int rc;
    if (tc) {
        // tensor-pipe screen + exact verification merge their keys into `best`; the gated kernel then only runs the
        // column tiles the screen flagged (its CTAs exit at once everywhere else)
        if (!pre_init)
            argmin_init_kernel<<<iblocks, 256, 0, stream>>>(w.best, nullptr, nullptr, (size_t)ld2_total, nullptr, 0, nullptr);
        if ((rc = launch_tc(0, xT, ld1_total, nx, yT, ld2_total, ny, k, pairs_dev, npairs, tc->uni_items, tc->total_items,
                            tc->uni_xtiles, tc->total_xtiles, uniform_ytiles, total_ytiles, tc->ld1_tc_total, flags, w.best,
                            nullptr, 0, tc->ws, stream))) return rc;
        prm.ncols = tc->ws.tcflag;
        if ((rc = launch_gated<1>(xT, ld1_total, yT, ld2_total, k, prm, total_items, stream))) return rc;
        finalize_argmin_kernel<<<total_ytiles, TILE, 0, stream>>>(w.best, pairs_dev, npairs, uniform_ytiles, flags, out);
        F2P_CUDA(cudaGetLastError());
        return 0;
    }
    if (path == PATH_GATED && pre_init && total_items <= FUSE_MAX_ITEMS) {
        // small problem (a few waves): keys and arrival counters were reset by launch_prep, the kernel finalises its own
        // column tiles — one launch instead of three
        prm.tilecnt = w.tilecnt;
        return launch_gated<1>(xT, ld1_total, yT, ld2_total, k, prm, total_items, stream);
    }
    if (path == PATH_GATED) {
        if (!pre_init)
            argmin_init_kernel<<<iblocks, 256, 0, stream>>>(w.best, nullptr, nullptr, (size_t)ld2_total, nullptr, 0, nullptr);
        if ((rc = launch_gated<1>(xT, ld1_total, yT, ld2_total, k, prm, total_items, stream))) return rc;   // exact, no fallback
        finalize_argmin_kernel<<<total_ytiles, TILE, 0, stream>>>(w.best, pairs_dev, npairs, uniform_ytiles, flags, out);
        F2P_CUDA(cudaGetLastError());
        return 0;
    }
    if (path == PATH_EXACT) {
        argmin_init_kernel<<<iblocks, 256, 0, stream>>>(w.best, nullptr, nullptr, (size_t)ld2_total, nullptr, 0, nullptr);
        if ((rc = launch_mode<0>(xT, ld1_total, yT, ld2_total, k, prm, total_items, stream))) return rc;
        finalize_argmin_kernel<<<total_ytiles, TILE, 0, stream>>>(w.best, pairs_dev, npairs, uniform_ytiles, flags, out);
        F2P_CUDA(cudaGetLastError());
        return 0;
    }
    argmin_init_kernel<<<iblocks, 256, 0, stream>>>(w.best, w.best2, w.best_exact, (size_t)ld2_total, w.nflag, npairs, nullptr);
    prm.best2 = w.best2;
    if ((rc = launch_mode<2>(xT, ld1_total, yT, ld2_total, k, prm, total_items, stream))) return rc;
    finalize_raw_kernel<<<total_ytiles, TILE, 0, stream>>>(w.best, w.best2, pairs_dev, npairs, uniform_ytiles, flags, out,
                                                          w.nflag, w.flagged);
    gather_flagged_kernel<<<total_ytiles, TILE, 0, stream>>>(yT, ny, pairs_dev, npairs, uniform_ytiles, k, ld2_total,
                                                            w.nflag, w.flagged, w.yTf, w.nyf);
    // exact re-check over the compacted columns: the grid is sized for the worst case, surplus CTAs exit at once
    prm.best = w.best_exact; prm.best2 = nullptr; prm.ny = w.nyf; prm.ncols = w.nflag;
    if ((rc = launch_mode<0>(xT, ld1_total, w.yTf, ld2_total, k, prm, total_items, stream))) return rc;
    finalize_recheck_kernel<<<total_ytiles, TILE, 0, stream>>>(w.best_exact, pairs_dev, npairs, uniform_ytiles, w.nflag,
                                                              w.flagged, out);
    F2P_CUDA(cudaGetLastError());
    return 0;
}

// ---- blocked arg-min: one (row block, column block) of a pair, merged into persistent keys -----------------------
// keys[j] = min over every block processed so far of (bits(d) << 32 | global row): order = (distance, row), so exact
// ties still go to the lowest CAD row of the whole pair, whatever order the row blocks arrive in.
__global__ void __launch_bounds__(256)
merge_keys_kernel(const u64* __restrict__ best, int n2, uint32_t row0, u64* __restrict__ keys) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n2) return;
    const u64 kx = best[j];
    if ((uint32_t)(kx & 0xffffffffull) >= IDX_NONE) return;       // no finite distance in this block
    atomicMin(keys + j, kx + (u64)row0);
}
__global__ void __launch_bounds__(256)
keys_to_map_kernel(const u64* __restrict__ keys, int n2, unsigned flags, int64_t* __restrict__ out) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n2) return;
    uint32_t idx = (uint32_t)(keys[j] & 0xffffffffull);
    if (idx >= IDX_NONE) idx = 0;      // all distances +inf/NaN: out of contract, mirror argmin's first index
    out[j] = (int64_t)idx;
    out[(int64_t)n2 + j] = row1_value(j, flags);
}

int launch_argmin_keys(const float* xT, int64_t ld1, const float* nx, const float* yT, int64_t ld2, const float* ny, int k,
                       const PairDev* pair_dev, int items, int ytiles, int n2, uint32_t row0, u64* keys, const ArgminWs& w,
                       cudaStream_t stream, const TCPlan* tc) {
    if (items <= 0 || ytiles <= 0) return 0;
    NNParams prm;
    prm.pairs = pair_dev; prm.npairs = 1; prm.uniform_items = items; prm.k = k;
    prm.nx = nx; prm.ny = ny; prm.best = w.best; prm.part5 = nullptr; prm.best2 = nullptr; prm.ncols = nullptr;
    prm.tileflag = nullptr; prm.tilecnt = nullptr; prm.out = nullptr; prm.flags = F2P_ROW0_ONLY;
    int rc;
    argmin_init_kernel<<<(unsigned)((ld2 + 255) / 256), 256, 0, stream>>>(w.best, nullptr, nullptr, (size_t)ld2, nullptr, 0, nullptr);
    if (tc) {
        if ((rc = launch_tc(0, xT, ld1, nx, yT, ld2, ny, k, pair_dev, 1, tc->uni_items, tc->total_items, tc->uni_xtiles,
                            tc->total_xtiles, ytiles, ytiles, tc->ld1_tc_total, F2P_ROW0_ONLY, w.best, nullptr, 0, tc->ws, stream,
                            keys))) return rc;          // the keys held so far seed the screen's thresholds
        prm.ncols = tc->ws.tcflag;
    }
    if ((rc = launch_gated<1>(xT, ld1, yT, ld2, k, prm, items, stream))) return rc;
    merge_keys_kernel<<<(unsigned)((n2 + 255) / 256), 256, 0, stream>>>(w.best, n2, row0, keys);
    F2P_CUDA(cudaGetLastError());
    return 0;
}

int launch_keys_to_map(const u64* keys, int n2, unsigned flags, int64_t* out, cudaStream_t stream) {
    if (n2 <= 0) return 0;
    keys_to_map_kernel<<<(unsigned)((n2 + 255) / 256), 256, 0, stream>>>(keys, n2, flags, out);
    F2P_CUDA(cudaGetLastError());
    return 0;
}

int launch_top5(const float* xT, int64_t ld1_total, const float* nx, const float* yT, int64_t ld2_total, const float* ny,
                int k, const PairDev* pairs_dev, int npairs, int uniform_items, int total_items, int uniform_ytiles,
                int total_ytiles, unsigned flags, u64* part5, int64_t* cand, int cand_is_out_layout, cudaStream_t stream,
                const TCPlan* tc) {
    if (total_items <= 0 || total_ytiles <= 0) return 0;
    NNParams prm;
    prm.pairs = pairs_dev; prm.npairs = npairs; prm.uniform_items = uniform_items; prm.k = k;
    prm.nx = nx; prm.ny = ny; prm.best = nullptr; prm.part5 = part5; prm.best2 = nullptr; prm.ncols = nullptr;
    prm.tileflag = nullptr; prm.tilecnt = nullptr; prm.out = nullptr; prm.flags = flags;
    int rc;
    if (tc) {
        // the screen + verification write nn_query's candidates for every column tile they can decide; the gated
        // kernel, its exact fallback and the merge below then only touch the flagged tiles
        if ((rc = launch_tc(1, xT, ld1_total, nx, yT, ld2_total, ny, k, pairs_dev, npairs, tc->uni_items, tc->total_items,
                            tc->uni_xtiles, tc->total_xtiles, uniform_ytiles, total_ytiles, tc->ld1_tc_total, flags, nullptr,
                            cand, cand_is_out_layout, tc->ws, stream))) return rc;
    }
    const int* tc_only = tc ? tc->ws.tcflag : nullptr;
    if (tc || nn_path() == PATH_GATED) {
        // tile flags follow the partial lists in the workspace (top5_ws_bytes)
        prm.tileflag = reinterpret_cast<int*>(part5 + (size_t)total_items * TILE * 5);
        F2P_CUDA(cudaMemsetAsync(prm.tileflag, 0, sizeof(int) * (size_t)total_ytiles, stream));
        prm.ncols = tc_only;
        if ((rc = launch_gated<5>(xT, ld1_total, yT, ld2_total, k, prm, total_items, stream))) return rc;
        prm.ncols = nullptr;
    }
    rc = launch_mode<1>(xT, ld1_total, yT, ld2_total, k, prm, total_items, stream);   // gated path: flagged tiles only
    if (rc) return rc;
    finalize_top5_kernel<<<total_ytiles, TILE, 0, stream>>>(part5, pairs_dev, npairs, uniform_ytiles, flags, cand,
                                                            cand_is_out_layout, tc_only);
    F2P_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace f2p

/*
 * f2p_oracle.c — CPU restatement of the reference's fmap->point-map path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library; the product (fmap2pointmap_solvers + libf2p_sm100.so) never does.
 *
 * The reference path is pure Python over PyTorch ATen ops (CPU: MKL sgemm + ATen reduce kernels);
 * its arithmetic therefore lives in third-party code (torch 2.x ATen, Intel MKL) that is not under
 * /root/reference.  This file restates, operation by operation and rounding by rounding, what those
 * ops compute on the CPU path, anchored on the reference call sites:
 *
 *   fmap2pointmap_solvers/naive.py:17-21,33-34            (projection, cdist, argmin, row-1 packing)
 *   fmap2pointmap_solvers/spacial_filtering.py:16-21      (entry), :32-38 (cdist, sort, top-5 packing)
 *   fmap2pointmap_solvers/spacial_filtering.py:42-49      (euclidean_distance)
 *   fmap2pointmap_solvers/spacial_filtering.py:51-75      (three voting rounds + branch)
 *
 * Parity pin: tests/test_oracle_golden.py checks this file bit-for-bit against golden vectors that
 * oracle/make_golden.py produced by importing and running the unmodified reference in the build
 * container (torch 2.11 CPU), and, when /root/reference is present, against the live reference.
 *
 * Build: see oracle/Makefile  (gcc -O3 -mavx2 -mfma -ffp-contract=off -fopenmp).
 * -ffp-contract=off is REQUIRED: every fma below is explicit, every other op rounds separately.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>

#define F2PO_API __attribute__((visibility("default")))

/* ------------------------------------------------------------------------------------------------
 * ATen cascade sum (aten/src/ATen/native/cpu/SumKernel.cpp: multi_row_sum / row_sum), used by
 * `.sum(-1)` inside cdist's row norms (naive.py:33 -> _euclidean_dist) and by `.mean(0)`
 * (spacial_filtering.py:54,58,68-73).  `get(ctx,i)` returns the i-th addend.
 * ---------------------------------------------------------------------------------------------- */
static int ceil_log2_i64(int64_t x) {
    if (x <= 2) return 1;
    int l = 0; uint64_t v = (uint64_t)(x - 1);
    while (v) { v >>= 1; ++l; }
    return l;
}

typedef float (*f2po_get)(const void* ctx, int64_t i);

/* one accumulator chain of multi_row_sum<acc_t, nrows=1>: items i = 0..size-1 */
static float cascade_sum(f2po_get get, const void* ctx, int64_t size) {
    const int num_levels = 4;
    int level_power = ceil_log2_i64(size) / num_levels;
    if (level_power < 4) level_power = 4;
    const int64_t level_step = (int64_t)1 << level_power;
    const int64_t level_mask = level_step - 1;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    int64_t i = 0;
    for (; i + level_step <= size;) {
        for (int64_t j = 0; j < level_step; ++j, ++i) acc[0] += get(ctx, i);
        for (int j = 1; j < num_levels; ++j) {
            acc[j] += acc[j - 1];
            acc[j - 1] = 0.f;
            const int64_t mask = level_mask << (j * level_power);
            if ((i & mask) != 0) break;
        }
    }
    for (; i < size; ++i) acc[0] += get(ctx, i);
    for (int j = 1; j < num_levels; ++j) acc[0] += acc[j];
    return acc[0];
}

/* row_sum<acc_t>: 4 interleaved cascade chains over size/4 items each, leftovers into chain 0 */
struct ilp_ctx { f2po_get get; const void* ctx; int64_t lane; };
static float ilp_get(const void* c, int64_t r) {
    const struct ilp_ctx* q = (const struct ilp_ctx*)c;
    return q->get(q->ctx, 4 * r + q->lane);
}
static float row_sum_ilp4(f2po_get get, const void* ctx, int64_t size) {
    const int64_t size_ilp = size / 4;
    float p[4];
    for (int l = 0; l < 4; ++l) {
        struct ilp_ctx q = { get, ctx, l };
        p[l] = cascade_sum(ilp_get, &q, size_ilp);
    }
    for (int64_t i = size_ilp * 4; i < size; ++i) p[0] += get(ctx, i);
    for (int l = 1; l < 4; ++l) p[0] += p[l];
    return p[0];
}

/* ------------------------------------------------------------------------------------------------
 * x.pow(2).sum(-1) as ATen computes it (vectorized_inner_sum, Vec = 8 floats):
 * squares are rounded to FP32 first (pow is its own kernel), then per-lane row_sum over the k/8
 * vectors, scalar tail, then lanes 0..7 folded into the tail accumulator.
 * ---------------------------------------------------------------------------------------------- */
struct sq_ctx { const float* x; int lane; };
static float sq_vec_get(const void* c, int64_t v) {
    const struct sq_ctx* q = (const struct sq_ctx*)c;
    float t = q->x[8 * v + q->lane];
    return t * t;
}
static float sq_scalar_get(const void* c, int64_t t) {
    const float* x = (const float*)c;
    return x[t] * x[t];
}
static float aten_sumsq_row(const float* x, int k) {
    /* k < Vec::size(): ATen takes scalar_inner_sum = row_sum with 4 interleaved scalar chains */
    if (k < 8) return row_sum_ilp4(sq_scalar_get, x, k);
    const int64_t nv = k / 8;
    float lanes[8];
    for (int l = 0; l < 8; ++l) {
        struct sq_ctx q = { x, l };
        lanes[l] = row_sum_ilp4(sq_vec_get, &q, nv);
    }
    float f = 0.f;
    for (int t = (int)(8 * nv); t < k; ++t) { float s = x[t] * x[t]; f += s; }
    for (int l = 0; l < 8; ++l) f += lanes[l];
    return f;
}

F2PO_API void f2po_rownorm(const float* x, int64_t n, int k, int64_t ld, float* out) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i) out[i] = aten_sumsq_row(x + i * ld, k);
}

/* ------------------------------------------------------------------------------------------------
 * Projection  feat_x = evecs_x @ C12.t()      (naive.py:20, spacial_filtering.py:19)
 * MKL sgemm for k <= 256: one FP32 accumulator per output, b ascending, fused multiply-add.
 * ---------------------------------------------------------------------------------------------- */
F2PO_API void f2po_project(const float* evecs, int64_t n, int k, int64_t ld,
                           const float* C, float* feat) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i) {
        const float* e = evecs + i * ld;
        for (int a = 0; a < k; ++a) {
            const float* c = C + (int64_t)a * k;
            float acc = 0.f;
            for (int b = 0; b < k; ++b) acc = __builtin_fmaf(e[b], c[b], acc);
            feat[i * (int64_t)k + a] = acc;
        }
    }
}

/* ------------------------------------------------------------------------------------------------
 * cdist (naive.py:33, spacial_filtering.py:32) for p=2 and max(V1,V2) > 25 is ATen _euclidean_dist:
 *   x_ = cat[-2*x, |x|^2, 1],  y_ = cat[y, 1, |y|^2],  r = x_ @ y_^T,  r.clamp_min_(0).sqrt_()
 * and MKL's sgemm (K = k+2 <= 258) is one FP32 accumulator, c ascending, fused multiply-add:
 *   acc = 0; acc = fma(-2*x[i,c], y[j,c], acc) for c<k; acc = fma(nx[i],1,acc); acc = fma(1,ny[j],acc)
 * d2row[j] receives the UNclamped accumulator for one row i over columns [j0,j1).
 * yT is the k-major copy of y ([k][ldy]) so the inner loop vectorises over j without changing any
 * per-element operation order.
 * ---------------------------------------------------------------------------------------------- */
static void dist2_row(const float* xm2_i, float nx_i, const float* yT, int64_t ldy, const float* ny,
                      int k, int64_t j0, int64_t j1, float* d2row) {
    enum { JB = 64 };
    for (int64_t jb = j0; jb < j1; jb += JB) {
        const int64_t w = (j1 - jb < JB) ? (j1 - jb) : JB;
        float acc[JB];
        for (int64_t t = 0; t < JB; ++t) acc[t] = 0.f;
        if (w == JB) {
            for (int c = 0; c < k; ++c) {
                const float a = xm2_i[c];
                const float* yr = yT + (int64_t)c * ldy + jb;
                for (int t = 0; t < JB; ++t) acc[t] = __builtin_fmaf(a, yr[t], acc[t]);
            }
        } else {
            for (int c = 0; c < k; ++c) {
                const float a = xm2_i[c];
                const float* yr = yT + (int64_t)c * ldy + jb;
                for (int64_t t = 0; t < w; ++t) acc[t] = __builtin_fmaf(a, yr[t], acc[t]);
            }
        }
        for (int64_t t = 0; t < w; ++t) {
            float v = acc[t] + nx_i;      /* fma(nx,1,acc) == plain add */
            v = v + ny[jb + t];           /* fma(1,ny,acc) == plain add */
            d2row[jb - j0 + t] = v;
        }
    }
}

static inline float clamp_sqrt(float d2) {
    float c = d2 < 0.f ? 0.f : d2;     /* clamp_min_(0); NaN is out of contract */
    return sqrtf(c);                   /* IEEE correctly rounded */
}

/* Prepares -2*x (exact), nx, yT, ny.  Caller frees. */
struct nn_prep { float* xm2; float* nx; float* yT; float* ny; int64_t ldy; };
static int nn_prepare(const float* feat_x, int64_t n1, const float* feat_y, int64_t n2, int k,
                      struct nn_prep* p) {
    p->ldy = (n2 + 15) & ~(int64_t)15;
    p->xm2 = (float*)malloc(sizeof(float) * (size_t)(n1 > 0 ? n1 : 1) * k);
    p->nx  = (float*)malloc(sizeof(float) * (size_t)(n1 > 0 ? n1 : 1));
    p->yT  = (float*)malloc(sizeof(float) * (size_t)(p->ldy > 0 ? p->ldy : 16) * k);
    p->ny  = (float*)malloc(sizeof(float) * (size_t)(n2 > 0 ? n2 : 1));
    if (!p->xm2 || !p->nx || !p->yT || !p->ny) return -1;
    f2po_rownorm(feat_x, n1, k, k, p->nx);
    f2po_rownorm(feat_y, n2, k, k, p->ny);
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n1; ++i)
        for (int c = 0; c < k; ++c) p->xm2[i * k + c] = feat_x[i * k + c] * -2.0f;
#pragma omp parallel for schedule(static)
    for (int64_t j = 0; j < n2; ++j)
        for (int c = 0; c < k; ++c) p->yT[(int64_t)c * p->ldy + j] = feat_y[j * k + c];
    return 0;
}
static void nn_release(struct nn_prep* p) { free(p->xm2); free(p->nx); free(p->yT); free(p->ny); }

/* nn_query of naive.py:23-35: argmin over dim=-2 of sqrt(clamp(d2)); first (lowest) index wins ties.
 * idx[j] for j in [0,n2).  Optional dmin[j] receives the winning (sqrt'ed) distance. */
F2PO_API int f2po_nn_argmin(const float* feat_x, int64_t n1, const float* feat_y, int64_t n2, int k,
                            int64_t* idx, float* dmin) {
    struct nn_prep p;
    if (nn_prepare(feat_x, n1, feat_y, n2, k, &p)) return -1;
    enum { CB = 512 };
    const int64_t nblk = (n2 + CB - 1) / CB;
#pragma omp parallel for schedule(dynamic, 1)
    for (int64_t b = 0; b < nblk; ++b) {
        const int64_t j0 = b * CB, j1 = (j0 + CB < n2) ? j0 + CB : n2;
        float best[CB]; int64_t bi[CB]; float row[CB];
        for (int t = 0; t < CB; ++t) { best[t] = INFINITY; bi[t] = 0; }
        for (int64_t i = 0; i < n1; ++i) {
            dist2_row(p.xm2 + i * k, p.nx[i], p.yT, p.ldy, p.ny, k, j0, j1, row);
            for (int64_t t = 0; t < j1 - j0; ++t) {
                const float d = clamp_sqrt(row[t]);
                if (d < best[t] || i == 0) { best[t] = d; bi[t] = i; }
            }
        }
        for (int64_t t = 0; t < j1 - j0; ++t) { idx[j0 + t] = bi[t]; if (dmin) dmin[j0 + t] = best[t]; }
    }
    nn_release(&p);
    return 0;
}

/* nn_query of spacial_filtering.py:22-40: full sort along V1, first K=5 per column.  torch.sort is
 * not stable there, so the order of EXACT distance ties is not defined by the reference; the build
 * rule is (distance, index) ascending.  idx5[j*5+r] = r-th nearest row of column j.
 * If n1 < 5 only min(n1,5) entries per column exist; kk returns that count. */
F2PO_API int f2po_nn_top5(const float* feat_x, int64_t n1, const float* feat_y, int64_t n2, int k,
                          int64_t* idx5, float* d5, int* kk_out) {
    struct nn_prep p;
    if (nn_prepare(feat_x, n1, feat_y, n2, k, &p)) return -1;
    const int K = n1 < 5 ? (int)n1 : 5;
    if (kk_out) *kk_out = K;
    enum { CB = 256 };
    const int64_t nblk = (n2 + CB - 1) / CB;
#pragma omp parallel for schedule(dynamic, 1)
    for (int64_t b = 0; b < nblk; ++b) {
        const int64_t j0 = b * CB, j1 = (j0 + CB < n2) ? j0 + CB : n2;
        float bd[CB][5]; int64_t bi[CB][5]; int cnt[CB]; float row[CB];
        for (int t = 0; t < CB; ++t) cnt[t] = 0;
        for (int64_t i = 0; i < n1; ++i) {
            dist2_row(p.xm2 + i * k, p.nx[i], p.yT, p.ldy, p.ny, k, j0, j1, row);
            for (int64_t t = 0; t < j1 - j0; ++t) {
                const float d = clamp_sqrt(row[t]);
                int c = cnt[t];
                if (c == K && !(d < bd[t][K - 1])) continue;   /* i ascending: ties keep earlier index */
                int pos = c < K ? c : K - 1;
                while (pos > 0 && d < bd[t][pos - 1]) { bd[t][pos] = bd[t][pos - 1]; bi[t][pos] = bi[t][pos - 1]; --pos; }
                bd[t][pos] = d; bi[t][pos] = i;
                if (c < K) cnt[t] = c + 1;
            }
        }
        for (int64_t t = 0; t < j1 - j0; ++t)
            for (int r = 0; r < K; ++r) {
                idx5[(j0 + t) * K + r] = bi[t][r];
                if (d5) d5[(j0 + t) * K + r] = bd[t][r];
            }
    }
    nn_release(&p);
    return 0;
}

/* Unclamped d2 and sqrt'ed distance for explicit (i,j) pairs — used by tests to adjudicate ties. */
F2PO_API int f2po_pair_dist(const float* feat_x, int64_t n1, const float* feat_y, int64_t n2, int k,
                            const int64_t* ii, const int64_t* jj, int64_t npairs, float* d2, float* d) {
    float* nx = (float*)malloc(sizeof(float) * (size_t)(n1 > 0 ? n1 : 1));
    float* ny = (float*)malloc(sizeof(float) * (size_t)(n2 > 0 ? n2 : 1));
    if (!nx || !ny) return -1;
    f2po_rownorm(feat_x, n1, k, k, nx);
    f2po_rownorm(feat_y, n2, k, k, ny);
    for (int64_t q = 0; q < npairs; ++q) {
        const float* x = feat_x + ii[q] * k; const float* y = feat_y + jj[q] * k;
        float acc = 0.f;
        for (int c = 0; c < k; ++c) acc = __builtin_fmaf(x[c] * -2.0f, y[c], acc);
        acc = acc + nx[ii[q]];
        acc = acc + ny[jj[q]];
        if (d2) d2[q] = acc;
        if (d) d[q] = clamp_sqrt(acc);
    }
    free(nx); free(ny);
    return 0;
}

/* ------------------------------------------------------------------------------------------------
 * euclidean_distance (spacial_filtering.py:42-49): for points t[a], t[b] (3 floats each)
 *   d_c = fl(t[a,c]-t[b,c]);  q_c = fl(d_c*d_c);  s = fl(fl(q0+q1)+q2);  sqrt_rn(s)
 * (separate ATen kernels: no contraction anywhere).
 * ---------------------------------------------------------------------------------------------- */
static inline float ed3(const float* ta, const float* tb) {
    float d0 = ta[0] - tb[0], d1 = ta[1] - tb[1], d2 = ta[2] - tb[2];
    float q0 = d0 * d0, q1 = d1 * d1, q2 = d2 * d2;
    float s = q0 + q1; s = s + q2;
    return sqrtf(s);
}

struct vote_ctx { const float* X; const float* Y; int64_t b; };
static float vote_get(const void* c, int64_t a) {
    const struct vote_ctx* q = (const struct vote_ctx*)c;
    float A = ed3(q->X + 3 * a, q->X + 3 * q->b);
    float B = ed3(q->Y + 3 * a, q->Y + 3 * q->b);
    return fabsf(A - B);
}

/* torch.absolute(A-B).mean(0) (spacial_filtering.py:54,58,68-73) for gathered points
 * X[a] = CAD[p[0][a]], Y[a] = PC[p[1][a]], a in [0,M).  ATen: sum over rows for column b with
 * vectorized_outer_sum (columns < 32*floor(M/32): single cascade chain; the rest: 4-way interleaved
 * row_sum), then true division by fl(M). */
F2PO_API void f2po_vote_scores(const float* X, const float* Y, int64_t M, float* s) {
    const int64_t mfull = (M / 32) * 32;
    const float fM = (float)M;
#pragma omp parallel for schedule(dynamic, 16)
    for (int64_t b = 0; b < M; ++b) {
        struct vote_ctx q = { X, Y, b };
        float sum = (b < mfull) ? cascade_sum(vote_get, &q, M) : row_sum_ilp4(vote_get, &q, M);
        s[b] = sum / fM;
    }
}

/* Python-style gather index (negative wraps once), as `PC[p_pred[1]]` does after the int16 wrap. */
static inline int64_t pyidx(int64_t i, int64_t n) { return i < 0 ? i + n : i; }

/* spacial_filtering (spacial_filtering.py:51-75).  p is [2][M] int64 (row 0 CAD index, row 1 PC index),
 * thr[0..3] = fl32(0.3*diam), fl32(0.15*diam), fl32(0.055*diam), fl32(0.065*diam) formed by the caller
 * exactly as torch does.  Writes survivors to out [2][M'] (row stride M_out_ld) and
 * returns M'.  round_counts[0..2] = survivors entering rounds 1..3; *skipped = 1 when the reference
 * would print "skipped".  Returns <0 on error. */
F2PO_API int64_t f2po_spatial_filter(const float* CAD, int64_t ncad, const float* PC, int64_t npc,
                                     const int64_t* p, int64_t M, const float* thr,
                                     int64_t* out, int64_t out_ld, int64_t* round_counts, int* skipped) {
    int64_t* cur0 = (int64_t*)malloc(sizeof(int64_t) * (size_t)(M > 0 ? M : 1));
    int64_t* cur1 = (int64_t*)malloc(sizeof(int64_t) * (size_t)(M > 0 ? M : 1));
    float* X = (float*)malloc(sizeof(float) * 3 * (size_t)(M > 0 ? M : 1));
    float* Y = (float*)malloc(sizeof(float) * 3 * (size_t)(M > 0 ? M : 1));
    float* s = (float*)malloc(sizeof(float) * (size_t)(M > 0 ? M : 1));
    if (!cur0 || !cur1 || !X || !Y || !s) return -1;
    memcpy(cur0, p, sizeof(int64_t) * (size_t)M);
    memcpy(cur1, p + M, sizeof(int64_t) * (size_t)M);
    int64_t m = M;
    if (skipped) *skipped = 0;
    for (int round = 0; round < 3; ++round) {
        if (round_counts) round_counts[round] = m;
        for (int64_t a = 0; a < m; ++a) {
            const int64_t ci = pyidx(cur0[a], ncad), pi = pyidx(cur1[a], npc);
            if (ci < 0 || ci >= ncad || pi < 0 || pi >= npc) { m = -2; goto done; }
            memcpy(X + 3 * a, CAD + 3 * ci, 3 * sizeof(float));
            memcpy(Y + 3 * a, PC + 3 * pi, 3 * sizeof(float));
        }
        f2po_vote_scores(X, Y, m, s);
        float t = thr[round];
        if (round == 2) {
            int64_t c055 = 0, c065 = 0;
            for (int64_t a = 0; a < m; ++a) { c055 += s[a] < thr[2]; c065 += s[a] < thr[3]; }
            if (c055 == 0) { t = thr[3]; if (c065 == 0 && skipped) *skipped = 1; }
            else t = thr[2];
        }
        int64_t w = 0;
        for (int64_t a = 0; a < m; ++a)
            if (s[a] < t) { cur0[w] = cur0[a]; cur1[w] = cur1[a]; ++w; }
        m = w;
    }
    for (int64_t a = 0; a < m; ++a) { out[a] = cur0[a]; out[out_ld + a] = cur1[a]; }
done:
    free(cur0); free(cur1); free(X); free(Y); free(s);
    return m;
}

/* linspace(0,V2-1,V2).type(int16) promoted to int64 (naive.py:21, spacial_filtering.py:36):
 * exact below 2^24, wraps modulo 2^16 above 32767 on the reference's x86 CPU path. */
F2PO_API int64_t f2po_row1_value(int64_t j, int int16_compat) {
    if (!int16_compat) return j;
    return (int64_t)(int16_t)(uint16_t)(j & 0xFFFF);
}

/* naive_fmap2pointmap (naive.py:5-21).  out is [2][n2] int64. */
F2PO_API int f2po_naive(const float* C, const float* evecs_x, int64_t n1, const float* evecs_y,
                        int64_t n2, int k, int int16_compat, int64_t* out) {
    float* feat = (float*)malloc(sizeof(float) * (size_t)(n1 > 0 ? n1 : 1) * k);
    if (!feat) return -1;
    f2po_project(evecs_x, n1, k, k, C, feat);
    int rc = f2po_nn_argmin(feat, n1, evecs_y, n2, k, out, NULL);
    for (int64_t j = 0; j < n2; ++j) out[n2 + j] = f2po_row1_value(j, int16_compat);
    free(feat);
    return rc;
}

/* spacial_filtering_fmap2pointmap (spacial_filtering.py:4-21).  out must hold [2][5*n2] int64 with row
 * stride 5*n2; returns M' (columns 0..M'-1 of each row valid) or <0 on error.
 * cand (optional, [2][5*n2]) receives the pre-filter candidates of nn_query. */
F2PO_API int64_t f2po_spatial(const float* C, const float* evecs_x, int64_t n1, const float* evecs_y,
                              int64_t n2, int k, const float* CAD, int64_t ncad, const float* PC,
                              int64_t npc, const float* thr, int int16_compat, int64_t* out,
                              int64_t* cand, int64_t* round_counts, int* skipped) {
    float* feat = (float*)malloc(sizeof(float) * (size_t)(n1 > 0 ? n1 : 1) * k);
    int K = 0;
    int64_t* idx5 = (int64_t*)malloc(sizeof(int64_t) * 5 * (size_t)(n2 > 0 ? n2 : 1));
    int64_t* p = (int64_t*)malloc(sizeof(int64_t) * 10 * (size_t)(n2 > 0 ? n2 : 1));
    if (!feat || !idx5 || !p) return -1;
    f2po_project(evecs_x, n1, k, k, C, feat);
    if (f2po_nn_top5(feat, n1, evecs_y, n2, k, idx5, NULL, &K)) return -1;
    const int64_t M = (int64_t)K * n2;
    for (int64_t j = 0; j < n2; ++j)
        for (int r = 0; r < K; ++r) {
            p[j * K + r] = idx5[j * K + r];
            p[M + j * K + r] = f2po_row1_value(j, int16_compat);
        }
    if (cand) memcpy(cand, p, sizeof(int64_t) * 2 * (size_t)M);
    int64_t m = f2po_spatial_filter(CAD, ncad, PC, npc, p, M, thr, out, 5 * n2, round_counts, skipped);
    free(feat); free(idx5); free(p);
    return m;
}

/* ---- inlier ratio ------------------------------------------------------------------------------------
 * compute_inlier_ratio (utils/utils.py:81-105): CAD[pred[:,0]], PC_aligned[pred[:,1]] (torch indexing: negative
 * indices wrap once), sq = square(a-b).sum(-1) [((q0+q1)+q2), FP32], dist = sq ** 0.5 [sqrt; IEEE here, see the
 * header note on MKL's vsSqrt], inliers = (dist < thr).sum().  Returns the count, or -1 on an out-of-range index. */
F2PO_API int64_t f2po_inlier_count(const float* CAD, int64_t ncad, const float* PC, int64_t npc,
                                   const int64_t* p0, const int64_t* p1, int64_t M, float thr) {
    int64_t n = 0;
    for (int64_t c = 0; c < M; ++c) {
        int64_t i = p0[c], j = p1[c];
        if (i < 0) i += ncad;
        if (j < 0) j += npc;
        if (i < 0 || i >= ncad || j < 0 || j >= npc) return -1;
        const float dx = CAD[3 * i] - PC[3 * j], dy = CAD[3 * i + 1] - PC[3 * j + 1], dz = CAD[3 * i + 2] - PC[3 * j + 2];
        const float qx = dx * dx, qy = dy * dy, qz = dz * dz;
        const float s = (qx + qy) + qz;
        if (sqrtf(s) < thr) ++n;
    }
    return n;
}

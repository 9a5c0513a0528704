// lstsq.cu — ground-truth functional map of a sample: C_gt = lstsq(evecs2[P[:,1]], evecs1[P[:,0]]).
//
// Replaces `C_from_sparse_P` (utils/utils.py:67-80: two row gathers + torch.linalg.lstsq, called per sample at
// scripts/eval.py:74 and scripts/train.py:101).  A = evecs2[P[:,1]] and B = evecs1[P[:,0]] are [npos,k] with k ~ 30 and
// npos in the hundreds to tens of thousands, so the problem is a tall-skinny reduction followed by a tiny dense solve:
//   gram_kernel   G = A^T A, H = A^T B per 512-row slab, FP64 accumulation of the FP32 products (exact products:
//                 24+24 bits fit the 53-bit significand), gathers fused; slab partials go to the workspace
//   solve_kernel  one CTA per problem: slab partials folded in slab order (deterministic), FP64 Cholesky solve; if a
//                 pivot is small relative to the largest diagonal entry the CTA switches to a cyclic Jacobi
//                 eigen-decomposition of G and applies the pseudo-inverse with torch's rank rule
//                 (sigma_i > rcond * sigma_max, rcond = eps32 * max(npos, k)), i.e. the minimum-norm solution gelsy
//                 returns for rank-deficient and underdetermined systems (npos < k, duplicated positives, npos = 0)
// Floating-point contract: the result is the FP64 normal-equation solution rounded once to FP32; its distance to the
// exact least-squares solution is ~cond(A)^2 * 2^-53, below the FP32 QR of the reference (~cond(A) * 2^-24) for
// cond(A) < 2^29.  Parity is by tolerance (tests/test_gpu_parity.py).
#include <string.h>
#include <vector>
#include "common.cuh"
#include "internal.h"

namespace f2p {

constexpr int LS_T = 256;          // threads per CTA (both kernels)
constexpr int LS_SLAB = 512;       // rows of A / B per gram CTA
constexpr int LS_TILE = 32;        // rows staged in shared memory at a time

struct LsqPair {
    const int64_t* P; int64_t npos;
    const float* e1; const float* e2;
    int n1, n2;
    double* part;                  // [nslab][2][k][k] slab partials (G then H)
    int* sflag;                    // [nslab] bit1 = index out of range in that slab
    int slab0, nslab;              // first gram work item of this pair / number of slabs
    float* C; int32_t* status;
};

__device__ __forceinline__ int find_lpair(int item, const LsqPair* lp, int npairs) {
    int lo = 0, hi = npairs - 1;
    while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (lp[mid].slab0 <= item) lo = mid; else hi = mid - 1; }
    return lo;
}

// EPT = entries of [G | H] per thread = ceil(2*KMAX^2 / LS_T)
template <int KMAX>
__global__ void __launch_bounds__(LS_T)
gram_kernel(const LsqPair* __restrict__ lps, int npairs, int k, int64_t ld1, int64_t ld2) {
    constexpr int EPT = (2 * KMAX * KMAX + LS_T - 1) / LS_T;
    __shared__ float sT[2][LS_TILE][KMAX + 1];           // [0] = rows of A, [1] = rows of B
    __shared__ int s_bad;
    const int pi = find_lpair(blockIdx.x, lps, npairs);
    const LsqPair& lp = lps[pi];
    const int slab = blockIdx.x - lp.slab0;
    const int64_t row0 = (int64_t)slab * LS_SLAB;
    const int nrows = (int)min((int64_t)LS_SLAB, lp.npos - row0);
    const int tid = threadIdx.x;
    if (tid == 0) s_bad = 0;
    // this thread's entries: e < k*k -> G[i][j] = sum a_i a_j; else H[i][j] = sum a_i b_j
    int oi[EPT], oj[EPT];
    double acc[EPT];
    const int kk = k * k;
#pragma unroll
    for (int q = 0; q < EPT; ++q) {
        const int e = tid + q * LS_T;
        const int ee = e < 2 * kk ? e : 0;               // surplus threads redo entry 0 and never store
        const int h = ee >= kk, r = ee - h * kk;
        oi[q] = r / k;                                   // column of A
        oj[q] = (r % k) + h * (LS_TILE * (KMAX + 1));    // column of A (G) or of B (H)
        acc[q] = 0.0;
    }
    const float* sAB = &sT[0][0][0];
    for (int t0 = 0; t0 < nrows; t0 += LS_TILE) {
        const int nt = min(LS_TILE, nrows - t0);
        __syncthreads();
        for (int x = tid; x < nt * k; x += LS_T) {
            const int r = x / k, c = x % k;
            const int64_t* pr = lp.P + 2 * (row0 + t0 + r);
            int64_t i = pr[0], j = pr[1];
            if (i < 0) i += lp.n1;
            if (j < 0) j += lp.n2;
            float a = 0.f, b = 0.f;
            if (i < 0 || i >= lp.n1 || j < 0 || j >= lp.n2) s_bad = 2;
            else { a = lp.e2[j * ld2 + c]; b = lp.e1[i * ld1 + c]; }
            sT[0][r][c] = a; sT[1][r][c] = b;
        }
        __syncthreads();
        for (int r = 0; r < nt; ++r) {
            const float* row = sAB + r * (KMAX + 1);
#pragma unroll
            for (int q = 0; q < EPT; ++q) acc[q] = fma((double)row[oi[q]], (double)row[oj[q]], acc[q]);
        }
    }
    __syncthreads();
    double* out = lp.part + (size_t)slab * 2 * kk;
#pragma unroll
    for (int q = 0; q < EPT; ++q) { const int e = tid + q * LS_T; if (e < 2 * kk) out[e] = acc[q]; }
    if (tid == 0) lp.sflag[slab] = s_bad;
}

constexpr double LS_EPS32 = 1.1920928955078125e-07;    // torch's default rcond factor for float32 inputs

// Sum of v over the CTA (all threads get the result).  scratch: >= LS_T/32 doubles of shared memory.
__device__ __forceinline__ double cta_sum(double v, double* scratch) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) scratch[threadIdx.x >> 5] = v;
    __syncthreads();
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < LS_T / 32; ++w) t += scratch[w];
    return t;
}

// One CTA per problem.  Dynamic shared memory: G, H, W (k*k doubles each), rot (2*k doubles), scratch.
__global__ void __launch_bounds__(LS_T)
solve_kernel(const LsqPair* __restrict__ lps, int k) {
    extern __shared__ __align__(16) double sm[];
    const LsqPair& lp = lps[blockIdx.x];
    const int tid = threadIdx.x, kk = k * k;
    double* G = sm; double* H = G + kk; double* W = H + kk; double* rot = W + kk; double* scratch = rot + 2 * k;
    __shared__ int s_flag, s_fail;
    if (tid == 0) { s_flag = 0; s_fail = 0; }
    __syncthreads();
    // fold the slab partials in slab order
    for (int e = tid; e < 2 * kk; e += LS_T) {
        double a = 0.0;
        for (int sl = 0; sl < lp.nslab; ++sl) a += lp.part[(size_t)sl * 2 * kk + e];
        G[e] = a;                                        // H follows G
    }
    int fl = 0;
    for (int sl = tid; sl < lp.nslab; sl += LS_T) fl |= lp.sflag[sl];
    if (fl) s_flag = fl;
    __syncthreads();
    if (s_flag & 2) {                                    // torch raises IndexError here; the host side does too
        for (int e = tid; e < kk; e += LS_T) lp.C[e] = __int_as_float(0x7fc00000);
        if (tid == 0) *lp.status = 2;
        return;
    }
    double maxd = 0.0;
    for (int i = 0; i < k; ++i) maxd = fmax(maxd, G[i * k + i]);
    const double rcond = LS_EPS32 * (double)(lp.npos > k ? lp.npos : (int64_t)k);
    // Cholesky only when every pivot clears the rank threshold with a wide margin (a pivot bounds the smallest
    // eigenvalue from above only); everything else goes through the eigen-decomposition, which applies the rule itself
    const double gate = fmax(1e-7, 1e3 * rcond * rcond) * maxd;
    for (int e = tid; e < kk; e += LS_T) W[e] = G[e];
    bool chol = maxd > 0.0;
    for (int j = 0; j < k && chol; ++j) {
        __syncthreads();
        const double d = W[j * k + j];
        if (!(d > gate)) { chol = false; break; }        // uniform: every thread reads the same word
        const double sd = sqrt(d);
        __syncthreads();
        for (int i = j + tid; i < k; i += LS_T) W[i * k + j] = (i == j) ? sd : W[i * k + j] / sd;
        __syncthreads();
        const int n = k - 1 - j;
        for (int x = tid; x < n * n; x += LS_T) {
            const int li = x / n, ll = x % n;
            if (ll <= li) { const int i = j + 1 + li, l = j + 1 + ll; W[i * k + l] -= W[i * k + j] * W[l * k + j]; }
        }
    }
    __syncthreads();
    int rank = k;
    if (chol) {
        if (tid < k) {                                   // one right-hand side per thread: L y = h, L^T x = y
            const int c = tid;
            for (int i = 0; i < k; ++i) {
                double v = H[i * k + c];
                for (int l = 0; l < i; ++l) v -= W[i * k + l] * H[l * k + c];
                H[i * k + c] = v / W[i * k + i];
            }
            for (int i = k - 1; i >= 0; --i) {
                double v = H[i * k + c];
                for (int l = i + 1; l < k; ++l) v -= W[l * k + i] * H[l * k + c];
                H[i * k + c] = v / W[i * k + i];
            }
        }
        __syncthreads();
        for (int e = tid; e < kk; e += LS_T) lp.C[e] = (float)H[e];
    } else {
        // cyclic Jacobi, round-robin ordering: n/2 disjoint rotations per step, n-1 steps per sweep
        double* V = W;
        for (int e = tid; e < kk; e += LS_T) V[e] = (e / k == e % k) ? 1.0 : 0.0;
        const int n = (k + 1) & ~1, half = n / 2;
        double diag2 = 0.0;
        for (int i = tid; i < k; i += LS_T) diag2 += G[i * k + i] * G[i * k + i];
        const double scale2 = cta_sum(diag2, scratch);
        for (int sweep = 0; sweep < 40; ++sweep) {
            double off = 0.0;
            for (int e = tid; e < kk; e += LS_T) if (e / k != e % k) off += G[e] * G[e];
            off = cta_sum(off, scratch);
            if (!(off > 1e-30 * scale2)) break;          // also leaves on NaN
            for (int step = 0; step < n - 1; ++step) {
                if (tid < half) {
                    int p = tid == 0 ? n - 1 : (step + tid) % (n - 1);
                    int q = tid == 0 ? step : (step - tid + n - 1) % (n - 1);
                    if (p > q) { const int t = p; p = q; q = t; }
                    double c = 1.0, sn = 0.0;
                    if (q < k) {
                        const double gpq = G[p * k + q];
                        if (gpq != 0.0) {
                            const double tau = (G[q * k + q] - G[p * k + p]) / (2.0 * gpq);
                            const double t = (tau >= 0.0 ? 1.0 : -1.0) / (fabs(tau) + sqrt(1.0 + tau * tau));
                            c = 1.0 / sqrt(1.0 + t * t); sn = t * c;
                        }
                    }
                    rot[2 * tid] = c; rot[2 * tid + 1] = sn;
                    reinterpret_cast<int*>(scratch + LS_T / 32)[2 * tid] = p;
                    reinterpret_cast<int*>(scratch + LS_T / 32)[2 * tid + 1] = q;
                }
                __syncthreads();
                const int* pq = reinterpret_cast<const int*>(scratch + LS_T / 32);
                for (int x = tid; x < half * k; x += LS_T) {             // G <- G J, V <- V J (columns p, q)
                    const int m = x / k, i = x % k, p = pq[2 * m], q = pq[2 * m + 1];
                    if (q >= k) continue;
                    const double c = rot[2 * m], sn = rot[2 * m + 1];
                    const double gp = G[i * k + p], gq = G[i * k + q];
                    G[i * k + p] = c * gp - sn * gq; G[i * k + q] = sn * gp + c * gq;
                    const double vp = V[i * k + p], vq = V[i * k + q];
                    V[i * k + p] = c * vp - sn * vq; V[i * k + q] = sn * vp + c * vq;
                }
                __syncthreads();
                for (int x = tid; x < half * k; x += LS_T) {             // G <- J^T G (rows p, q)
                    const int m = x / k, j = x % k, p = pq[2 * m], q = pq[2 * m + 1];
                    if (q >= k) continue;
                    const double c = rot[2 * m], sn = rot[2 * m + 1];
                    const double gp = G[p * k + j], gq = G[q * k + j];
                    G[p * k + j] = c * gp - sn * gq; G[q * k + j] = sn * gp + c * gq;
                }
                __syncthreads();
            }
        }
        // pseudo-inverse: X = V diag(1/lambda_i, lambda_i > (rcond sigma_max)^2) V^T H
        double lmax = 0.0;
        for (int i = 0; i < k; ++i) lmax = fmax(lmax, G[i * k + i]);
        const double cut = rcond * rcond * lmax;
        rank = 0;
        for (int i = 0; i < k; ++i) rank += (G[i * k + i] > cut && G[i * k + i] > 0.0) ? 1 : 0;
        double* T = rot + 2 * k + LS_T;                  // k*k doubles behind the scratch area
        for (int e = tid; e < kk; e += LS_T) {
            const int i = e / k, c = e % k;
            const double lam = G[i * k + i];
            double v = 0.0;
            if (lam > cut && lam > 0.0) {
                for (int r = 0; r < k; ++r) v += V[r * k + i] * H[r * k + c];
                v /= lam;
            }
            T[e] = v;
        }
        __syncthreads();
        for (int e = tid; e < kk; e += LS_T) {
            const int r = e / k, c = e % k;
            double v = 0.0;
            for (int i = 0; i < k; ++i) v += V[r * k + i] * T[i * k + c];
            lp.C[e] = (float)v;
        }
    }
    if (tid == 0) *lp.status = rank < k ? 1 : 0;
}

static size_t solve_smem_bytes(int k) {
    // G, H, W, T (k*k each), rot (2k), scratch + pair table (LS_T doubles: 8 warp sums + k ints, generously)
    return ((size_t)4 * k * k + 2 * k + LS_T) * sizeof(double);
}

static inline size_t al256(size_t b) { return (b + 255) & ~(size_t)255; }
static inline int nslabs(int64_t npos) { return (int)((npos + LS_SLAB - 1) / LS_SLAB); }

}  // namespace f2p

using namespace f2p;

extern "C" {

size_t f2p_lstsq_workspace_bytes(const F2PLstsqDesc* d, int npairs, int k) {
    if (!d || npairs <= 0 || k <= 0) return 256;
    size_t b = al256(sizeof(LsqPair) * (size_t)npairs);
    for (int i = 0; i < npairs; ++i) {
        const size_t ns = (size_t)nslabs(d[i].npos > 0 ? d[i].npos : 0);
        b += al256(ns * 2 * (size_t)k * k * sizeof(double)) + al256(ns * sizeof(int));
    }
    return b;
}

int f2p_c_from_sparse_p_batched(const int64_t* P, const float* evecs1, int64_t ld1, const float* evecs2, int64_t ld2,
                                const F2PLstsqDesc* d, int npairs, int k, float* C, int32_t* status, void* ws_v,
                                size_t ws_bytes, void* stream) {
    if (npairs == 0) return 0;
    if (!evecs1 || !evecs2 || !d || !C || !status || !ws_v) { set_error("null pointer"); return F2P_EINVAL; }
    if (npairs < 0 || k < 1 || ld1 < k || ld2 < k) { set_error("npairs=%d k=%d ld1=%lld ld2=%lld", npairs, k, (long long)ld1, (long long)ld2); return F2P_EINVAL; }
    if (k > F2P_LSTSQ_MAX_K) { set_error("k=%d > F2P_LSTSQ_MAX_K=%d", k, F2P_LSTSQ_MAX_K); return F2P_ENOTSUP; }
    if (ws_bytes < f2p_lstsq_workspace_bytes(d, npairs, k)) { set_error("workspace too small"); return F2P_ENOSPACE; }
    unsigned char* ws = static_cast<unsigned char*>(ws_v);
    std::vector<LsqPair> lps((size_t)npairs);
    size_t off = al256(sizeof(LsqPair) * (size_t)npairs);
    int total = 0;
    for (int i = 0; i < npairs; ++i) {
        if (d[i].npos < 0 || d[i].n1 < 1 || d[i].n2 < 1 || (d[i].npos > 0 && !P)) {
            set_error("pair %d: npos=%lld n1=%d n2=%d", i, (long long)d[i].npos, d[i].n1, d[i].n2); return F2P_EINVAL;
        }
        LsqPair& lp = lps[i];
        lp.P = P + d[i].p_off; lp.npos = d[i].npos;
        lp.e1 = evecs1 + d[i].evecs1_off; lp.e2 = evecs2 + d[i].evecs2_off;
        lp.n1 = d[i].n1; lp.n2 = d[i].n2;
        lp.nslab = nslabs(d[i].npos); lp.slab0 = total; total += lp.nslab;
        lp.part = reinterpret_cast<double*>(ws + off); off += al256((size_t)lp.nslab * 2 * k * k * sizeof(double));
        lp.sflag = reinterpret_cast<int*>(ws + off); off += al256((size_t)lp.nslab * sizeof(int));
        lp.C = C + (size_t)i * k * k; lp.status = status + i;
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    F2P_CUDA(cudaMemcpyAsync(ws, lps.data(), sizeof(LsqPair) * (size_t)npairs, cudaMemcpyHostToDevice, st));
    const LsqPair* lpd = reinterpret_cast<const LsqPair*>(ws);
    if (total > 0) {
        if (k <= 32) gram_kernel<32><<<total, LS_T, 0, st>>>(lpd, npairs, k, ld1, ld2);
        else gram_kernel<64><<<total, LS_T, 0, st>>>(lpd, npairs, k, ld1, ld2);
    }
    const size_t smem = solve_smem_bytes(k);
    static bool attr_set = false;
    if (!attr_set) {
        F2P_CUDA(cudaFuncSetAttribute(solve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)solve_smem_bytes(F2P_LSTSQ_MAX_K)));
        attr_set = true;
    }
    solve_kernel<<<npairs, LS_T, smem, st>>>(lpd, k);
    F2P_CUDA(cudaGetLastError());
    return 0;
}

}  // extern "C"

/*
 * f2p.h — C-ABI of libf2p_sm100.so: B200 (sm_100a) functional-map -> point-map conversion.
 *
 * The reference has no FFI: its path is Python over ATen.  Every entry point below replaces a group of ATen
 * call sites of /root/reference/fmap2pointmap_solvers (cited per function); the drop-in Python package
 * `fmap2pointmap_solvers` binds them with ctypes (INTEGRATION.md shows the stub).
 *
 * Conventions
 *   - plain pointers and sizes only; `stream` is a cudaStream_t passed as void* (NULL = legacy default stream);
 *   - *_dev pointers are device pointers on the current CUDA device, *_host pointers are host pointers;
 *   - all floating data is FP32 row-major, all returned indices are int64 (the reference's dtype);
 *   - every function returns 0 on success, a negative F2P_E* code on argument errors, or a positive
 *     cudaError_t; f2p_last_error() returns a thread-local description;
 *   - functions are asynchronous w.r.t. the host unless stated; the workspace must stay alive and must not
 *     be shared by concurrent calls on different streams.
 */
#ifndef F2P_H_
#define F2P_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define F2P_VERSION 100

#define F2P_EINVAL   (-1)  /* bad argument (null pointer, negative size, k out of range)            */
#define F2P_ENOSPACE (-2)  /* workspace too small                                                    */
#define F2P_ENOTSUP  (-3)  /* shape outside the supported envelope (k > F2P_MAX_K, n > 2^31-129 ...) */
#define F2P_EDEVICE  (-4)  /* not an sm_100 device / driver entry point missing                      */

#define F2P_MAX_K 256      /* the reference's single-accumulator sgemm contract holds for k <= 256   */
#define F2P_TOPK 5         /* K of spacial_filtering.py:33                                           */

/* flags */
#define F2P_INT16_COMPAT 1u /* row 1 = int16-wrapped point index, as naive.py:21 / spacial_filtering.py:36 */

/* One object/scene pair of a batch.  Offsets are in ELEMENTS from the batch base pointers.
 * (scripts/eval.py:83-87 is the per-sample loop this replaces.) */
typedef struct F2PPairDesc {
    int64_t evecs_x_off;  /* into evecs_x base: [n1,k] with row stride ldx */
    int64_t evecs_y_off;  /* into evecs_y base: [n2,k] with row stride ldy */
    int64_t c_off;        /* into C base: [k,k] contiguous                 */
    int64_t cad_off;      /* into CAD base: [ncad,3] contiguous (spatial only) */
    int64_t pc_off;       /* into PC base:  [npc,3] contiguous (spatial only)  */
    int64_t out_off;      /* into out base: naive [2,n2]; spatial: 2*5*n2 slots holding the dense [2,M'] result */
    int32_t n1, n2;       /* rows of evecs_x / evecs_y                     */
    int32_t ldx, ldy;     /* row strides of evecs_x / evecs_y in elements  */
    int32_t ncad, npc;    /* rows of CAD / PC (>= n1 / n2; zero-padded rows allowed) */
    float thr[4];         /* fl32(c*diam) for c = 0.3, 0.15, 0.055, 0.065 (spacial_filtering.py:54,58,68-73) */
} F2PPairDesc;

int f2p_version(void);
const char* f2p_last_error(void);

/* Device properties the host side needs for sizing (SM count, etc.). Returns 0 or an error. */
int f2p_device_info(int* sm_count, int* cc_major, int* cc_minor, size_t* smem_optin);

/* Workspace bytes for a batch with the given totals (sum over pairs of padded sizes is computed inside):
 * spatial != 0 additionally reserves the top-5 and filter scratch. */
size_t f2p_workspace_bytes(const F2PPairDesc* descs_host, int npairs, int k, int spatial);

/* ---- whole-path entry points (device buffers) ------------------------------------------------------- */

/* naive_fmap2pointmap (naive.py:5-21) for a batch of pairs sharing k.
 * out[pair]: int64 [2,n2]: row 0 = argmin CAD row (lowest index on exact distance ties), row 1 = point index. */
int f2p_naive_batched(const float* evecs_x_dev, const float* evecs_y_dev, const float* C_dev,
                      const F2PPairDesc* descs_host, int npairs, int k, unsigned flags,
                      int64_t* out_dev, void* ws_dev, size_t ws_bytes, void* stream);

/* spacial_filtering_fmap2pointmap (spacial_filtering.py:4-75) for a batch of pairs sharing k.
 * out[pair]: 2*5*n2 int64 slots; the result is written DENSE as int64 [2,M'] (row stride M') at out_off, so the
 * caller's tensor is a view of the first 2*M' slots; counts_dev[pair] = M' (int32);
 * status_dev[pair] bit0 = the reference would have printed "skipped" (spacial_filtering.py:70-71).
 * cand_dev (optional, 2*5*n2 slots per pair at out_off, row stride 5*n2) receives nn_query's [2,5*n2] candidates
 * (spacial_filtering.py:22-40). */
int f2p_spatial_batched(const float* evecs_x_dev, const float* evecs_y_dev, const float* C_dev,
                        const float* cad_dev, const float* pc_dev,
                        const F2PPairDesc* descs_host, int npairs, int k, unsigned flags,
                        int64_t* out_dev, int32_t* counts_dev, int32_t* status_dev, int64_t* cand_dev,
                        void* ws_dev, size_t ws_bytes, void* stream);

/* ---- single-kernel entry points (device buffers; used by tests, bench roofline and the Python host) -- */

/* Kernel 1: spectral projection (naive.py:20 / spacial_filtering.py:19) into the k-major operand layout:
 * xT[c][i] = -2 * sum_b evecs[i,b]*C[c,b]  (sequential-b FMA, exact scaling), ld = leading dim (multiple of 128),
 * nrm[i] = |feat_x[i]|^2 in ATen's summation order; columns i >= n are zero / nrm = +inf.
 * C_dev == NULL transposes only (xT[c][i] = evecs[i,c], nrm = |row|^2, padding nrm = 0): the evecs_y operand. */
int f2p_project(const float* evecs_dev, int n, int k, int ld_in, const float* C_dev,
                float* xT_dev, float* nrm_dev, int ld, void* ws_dev, size_t ws_bytes, void* stream);

/* Kernel 2: fused all-pairs squared distance + running arg-min (naive.py:33-34) over k-major operands.
 * idx_out[j] (int64) = argmin_i sqrt(max(0, xT[:,i].yT[:,j] + nx[i] + ny[j])), lowest i on exact ties. */
int f2p_nn_argmin(const float* xT_dev, const float* nx_dev, int n1, int ld1,
                  const float* yT_dev, const float* ny_dev, int n2, int ld2, int k,
                  int64_t* idx_out_dev, void* ws_dev, size_t ws_bytes, void* stream);

/* Kernel 2b: same distances, 5 nearest rows per column ordered by (distance, index) (spacial_filtering.py:32-35).
 * idx5_out[j*5+r] (int64). Requires n1 >= 5. */
int f2p_nn_top5(const float* xT_dev, const float* nx_dev, int n1, int ld1,
                const float* yT_dev, const float* ny_dev, int n2, int ld2, int k,
                int64_t* idx5_out_dev, void* ws_dev, size_t ws_bytes, void* stream);

/* Kernel 2 on ONE (row block, column block) of a pair, for host pipelines that overlap the upload of a huge pair with its
 * computation (naive.py:33-34 evaluated block by block; fmap2pointmap_solvers/pipeline.py: naive_host_blocked).
 * keys_dev[j] (u64, j < n2; set every byte to 0xFF before the first block) = min over the blocks processed so far of
 * (bits(d_ij) << 32 | row0 + i): merging by (distance, global row) keeps the reference's lowest-index rule for exact ties
 * across blocks, in whatever order they arrive; the keys held on entry also seed the thresholds of the block's scan, so
 * later row blocks of a scene block record only rows that can still win.  xT / nx describe the block's n1 rows (global rows row0 .. row0 + n1 - 1),
 * yT / ny its n2 scene points; workspace as for f2p_nn_argmin (f2p_nn_workspace_bytes(n1, n2, k, 0)). */
int f2p_nn_argmin_keys(const float* xT_dev, const float* nx_dev, int n1, int ld1, int row0,
                       const float* yT_dev, const float* ny_dev, int n2, int ld2, int k,
                       unsigned long long* keys_dev, void* ws_dev, size_t ws_bytes, void* stream);
/* keys -> the naive solver's int64 [2,n2] map (naive.py:21): row 0 = CAD row, row 1 = scene point index. */
int f2p_keys_to_map(const unsigned long long* keys_dev, int n2, unsigned flags, int64_t* out_dev, void* stream);

/* Kernel 3: spatial-consistency filter (spacial_filtering.py:51-75) on candidate pairs p[2][M] (row stride M).
 * Writes the survivors dense as int64 [2,M'] (row stride M') into out (2*M slots), *count_dev = M',
 * *status_dev bit0 = "skipped";
 * mask_dev (optional, uint8 [M]) = 1 for surviving original columns; scores_dev (optional float [3][M]) = the
 * per-round mean |A-B| of the columns alive in that round, in original column positions (others untouched). */
int f2p_spatial_filter(const float* cad_dev, int ncad, const float* pc_dev, int npc,
                       const int64_t* p_dev, int M, const float thr[4],
                       int64_t* out_dev, int32_t* count_dev, int32_t* status_dev,
                       uint8_t* mask_dev, float* scores_dev,
                       void* ws_dev, size_t ws_bytes, void* stream);

/* Kernel 2 without a workspace argument and without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same result as f2p_nn_argmin (the FP32 kernels; idx_out[j] int64, lowest i on exact ties).
 * The scratch is a static device array inside the library (2.25 MB held by every GPU that loads it, until exit), which
 * limits the shape: n2 <= F2P_STATIC_MAX_N2, k <= F2P_STATIC_MAX_K (n1 is not limited); beyond that the call returns
 * F2P_ENOTSUP.  ld2 must equal round_up(n2, 128).  The function calls no cudaMalloc / cudaFree or allocating routine.
 * ALL calls share the one array, which keeps the last call's contents (every call re-initialises what it reads): calls
 * that overlap on the GPU — different streams or threads — race; serialise them. */
#define F2P_STATIC_MAX_N2 4096
#define F2P_STATIC_MAX_K  128
int f2p_nn_argmin_static(const float* xT_dev, const float* nx_dev, int n1, int ld1,
                         const float* yT_dev, const float* ny_dev, int n2, int ld2, int k,
                         int64_t* idx_out_dev, void* stream);

size_t f2p_project_workspace_bytes(void);
size_t f2p_nn_workspace_bytes(int n1, int n2, int k, int top5);
size_t f2p_filter_workspace_bytes(int M);

/* ---- inlier ratio of the predicted maps (utils/utils.py:81-105, called per sample at scripts/eval.py:88) ---- */

/* For pair i and every column c < M_i of its dense int64 [2,M_i] map at maps_dev + descs[i].out_off (M_i = n2 for
 * naive maps; counts_dev[i] when spatial != 0): inlier iff sqrt(|CAD[p0] - PC_aligned[p1]|^2) < thr_host[i], with
 * CAD at cad_dev + cad_off ([ncad,3]) and the ground-truth-aligned cloud at pc_aligned_dev + pc_off ([npc,3]).
 * inliers_dev[i] = count (int32; -1 when an index is out of range), ratio_dev[i] (optional) = fp32(count)/fp32(M_i),
 * 0 when M_i == 0 (the reference returns 0 for an empty map).  Negative indices wrap once, as torch indexing does. */
int f2p_inlier_ratio_batched(const float* cad_dev, const float* pc_aligned_dev, const int64_t* maps_dev,
                             const F2PPairDesc* descs_host, int npairs, int spatial, const int32_t* counts_dev,
                             const float* thr_host, int32_t* inliers_dev, float* ratio_dev,
                             void* ws_dev, size_t ws_bytes, void* stream);
size_t f2p_inlier_workspace_bytes(int npairs);

/* ---- ground-truth correspondences of a sample (dataset/object.py:281-288 find_positives, :311-317 get_overlap) ---- */

/* All (i,j) with sqrt(((d0^2+d1^2)+d2^2)) <= r between pc1 [n1,3] and pc2 [n2,3] (both float64 when is_f64 != 0, else
 * float32; the arithmetic runs in that type, r is cast to it), in row-major order.
 * Pass 1 `_count`: *total_dev = N (int64), overlap12_dev[i] / overlap21_dev[j] (optional int8) = 1 when row i /
 * column j has a positive; the per-row offsets stay in the workspace.  Pass 2 `_write` (same arguments and workspace,
 * after the caller sized pairs_dev from N): pairs_dev int64 [N,2], at most `capacity` pairs are written. */
int f2p_find_positives_count(const void* pc1_dev, int n1, const void* pc2_dev, int n2, int is_f64, double r,
                             int64_t* total_dev, int8_t* overlap12_dev, int8_t* overlap21_dev,
                             void* ws_dev, size_t ws_bytes, void* stream);
int f2p_find_positives_write(const void* pc1_dev, int n1, const void* pc2_dev, int n2, int is_f64, double r,
                             int64_t* pairs_dev, int64_t capacity, void* ws_dev, size_t ws_bytes, void* stream);
size_t f2p_positives_workspace_bytes(int n1);

/* ---- ground-truth functional map of a sample (utils/utils.py:67-80 C_from_sparse_P; scripts/eval.py:74, train.py:101) ---- */

/* One least-squares problem of a batch: positives P (int64 [npos,2], rows (i in evecs1, j in evecs2)) at
 * P_dev + p_off, bases at evecs1_dev + evecs1_off ([n1, >=k], row stride ld1) and evecs2_dev + evecs2_off. */
typedef struct F2PLstsqDesc {
    int64_t p_off;        /* in int64 elements (2 per positive) */
    int64_t npos;
    int64_t evecs1_off, evecs2_off;   /* in floats */
    int32_t n1, n2;
} F2PLstsqDesc;

#define F2P_LSTSQ_MAX_K 64

/* C[pair] (fp32 [k,k] at C_dev + pair*k*k) = argmin_X || evecs2[P[:,1]] X - evecs1[P[:,0]] ||_F, the minimum-norm
 * solution when the system is rank deficient — what torch.linalg.lstsq(...)[0][:k] returns (driver gelsy, rcond =
 * eps32*max(npos,k)).  Floating-point contract: the Gram matrices A^T A, A^T B are accumulated in FP64 in a fixed order
 * (deterministic), solved in FP64 (Cholesky; Jacobi eigen-decomposition + pseudo-inverse when a pivot is small) and
 * rounded to FP32 once — parity with the reference is by tolerance (tests/test_gpu_parity.py states it).
 * status_dev[pair]: bit0 = rank deficient (minimum-norm branch), bit1 = an index of P out of range (C = NaN).
 * Negative indices wrap once, as torch indexing does.  k <= F2P_LSTSQ_MAX_K. */
int f2p_c_from_sparse_p_batched(const int64_t* P_dev, const float* evecs1_dev, int64_t ld1, const float* evecs2_dev,
                                int64_t ld2, const F2PLstsqDesc* descs_host, int npairs, int k, float* C_dev,
                                int32_t* status_dev, void* ws_dev, size_t ws_bytes, void* stream);
size_t f2p_lstsq_workspace_bytes(const F2PLstsqDesc* descs_host, int npairs, int k);

/* ---- pairwise-consistency front half of the TEASER++ solve (TEASER-plusplus/teaser/src/registration.cc) ----
 * For the configuration scripts/test_teaser.py:363-370 uses (estimate_scaling = false).  src / dst are the N matched
 * points as float64 [N,3] (the memory layout of TEASER's column-major Eigen 3xN matrices).
 * Parity: the inlier mask is pinned to the reference's own fixture (test/teaser/data/registration_test/
 * fixed_scale_inliers.csv, tests/golden/teaser_golden.npz); the reference itself cannot be compiled here (Eigen is not in
 * the image), so the order of the 3-term norm sum stays an assumption that the fixture does not exercise (csrc/tims.cu). */

/* computeTIMs (registration.cc:481-519) fused with ScaleInliersSelector::solveForScale (:427-445) and the inlier-graph
 * edge loop (:582-590): pair (i,j), i < j, is an edge iff | |src_j-src_i| - |dst_j-dst_i| | <= beta
 * (beta = 2*noise_bound*sqrt(cbar2)).  Pass 1 `_count`: *total_dev = E (int64); degree_dev (optional int32 [N]) =
 * vertex degrees; mask_dev (optional uint8 [N(N-1)/2]) = scale_inliers_mask_ in TIM order.  Pass 2 `_write` (same
 * arguments and workspace): edges_dev int32 [E,2] = (i,j) in the reference's addEdge order (i ascending, then j). */
int f2p_tim_graph_count(const double* src_dev, const double* dst_dev, int n, double beta, int64_t* total_dev,
                        int32_t* degree_dev, uint8_t* mask_dev, void* ws_dev, size_t ws_bytes, void* stream);
int f2p_tim_graph_write(const double* src_dev, const double* dst_dev, int n, double beta, int32_t* edges_dev,
                        int64_t capacity, void* ws_dev, size_t ws_bytes, void* stream);
size_t f2p_tim_workspace_bytes(int n);

/* computeTIMs alone (registration.cc:481-519): vtilde_dev float64 [N(N-1)/2, 3] (= column-major 3 x T),
 * map_dev (optional) int32 [N(N-1)/2, 2] = (i, j). */
int f2p_compute_tims(const double* v_dev, int n, double* vtilde_dev, int32_t* map_dev, void* stream);

/* ---- host-buffer convenience (what a non-torch caller binds; copies are inside the call) ------------- */

/* Single pair, HOST pointers in and out; synchronous (one submission + flush on a per-thread pair stream of depth 1,
 * so repeated calls allocate nothing).  out_host: naive [2,n2]; spatial: room for 2*5*n2 slots, the first 2*(*count)
 * receive the dense [2,M'] result. */
int f2p_naive_host(const float* C_host, const float* evecs_x_host, int n1, const float* evecs_y_host, int n2,
                   int k, unsigned flags, int64_t* out_host);
int f2p_spatial_host(const float* C_host, const float* evecs_x_host, int n1, const float* evecs_y_host, int n2,
                     int k, const float* cad_host, int ncad, const float* pc_host, int npc, const float thr[4],
                     unsigned flags, int64_t* out_host, int32_t* count_host, int32_t* status_host);

/* A stream of single-pair conversions with host buffers — the per-sample loop of scripts/eval.py:83-91 with `depth`
 * pairs in flight.  The object owns an upload stream, a download stream and `depth` staging slots, each with its own
 * kernel stream and workspace (device + pinned memory, grow-only: a steady stream of pairs allocates nothing per
 * submission; the kernels of small pairs in flight overlap on the GPU) on the device that is current at creation.  submit_* enqueues
 * the upload of the pair, its kernels and the download of its result and returns at once; when all slots are in flight
 * it first completes the OLDEST submission.  A submission is complete — its out_host (and count_host / status_host)
 * written — once *completed_out, the number of completed submissions counted in submission order, has passed it;
 * flush completes everything in flight.  Input pointers must stay valid and unchanged until the submission is complete
 * (pinned inputs keep the uploads asynchronous; pageable ones work).  Results are bit-identical to f2p_naive_host /
 * f2p_spatial_host.  One object per host thread; out_host sizes as for the *_host calls below. */
typedef struct F2PPairStream F2PPairStream;
int f2p_pair_stream_create(int depth, F2PPairStream** stream_out);             /* depth 1..16 */
int f2p_pair_stream_destroy(F2PPairStream* s);                                 /* flushes, then frees everything */
int f2p_pair_stream_submit_naive(F2PPairStream* s, const float* C_host, const float* evecs_x_host, int n1,
                                 const float* evecs_y_host, int n2, int k, unsigned flags, int64_t* out_host,
                                 long long* completed_out);
int f2p_pair_stream_submit_spatial(F2PPairStream* s, const float* C_host, const float* evecs_x_host, int n1,
                                   const float* evecs_y_host, int n2, int k, const float* cad_host, int ncad,
                                   const float* pc_host, int npc, const float thr[4], unsigned flags, int64_t* out_host,
                                   int32_t* count_host, int32_t* status_host, long long* completed_out);
int f2p_pair_stream_flush(F2PPairStream* s, long long* completed_out);

/* Measured FP32 FMA throughput of this device with the kernel-2 instruction mix (packed fma.rn.f32x2):
 * runs ~ms of pure FMA work and returns TFLOP/s (roofline denominator next to the nominal figure). */
int f2p_measure_fp32_peak(double* tflops_packed, double* tflops_scalar, void* stream);

/* Measured TF32 tensor-pipe throughput of this device with the instruction shape of the kernel-2 screen (tcgen05.mma
 * kind::tf32, M128 N256 K8, operands in shared memory): TFLOP/s, the roofline denominator of csrc/tc.cu. */
int f2p_measure_tf32_peak(double* tflops, void* stream);

/* Measured TMEM -> register bandwidth (tcgen05.ld 32x32b.x32, 8 warps per SM, no MMA running): bytes per clock and SM at
 * the nominal SM clock, and GB/s for the device.  Bounds the screen's epilogue (csrc/tc.cu); measurement aid. */
int f2p_measure_tmem_ld(double* bytes_per_clk_sm, double* gbs, void* stream);

/* Counters of the most recent tensor-pipe screen (synchronous read; measurement aid, no reference counterpart):
 * out[0] candidates recorded by the screen, out[1] candidate lists that overflowed (their column tiles were redone by the
 * FP32 kernels), out[2] candidates re-evaluated exactly, out[3] re-evaluated candidates whose tensor-pipe lower bound
 * exceeded the exact value (must be 0: the run-time check of the error-band assumptions), out[4..7] reserved. */
int f2p_tc_last_stats(unsigned long long out[8]);

/* Stage timing of f2p_naive_batched / f2p_spatial_batched for the per-kernel roofline of bench.py (measurement aid; no
 * reference counterpart).  Off by default; when on, the two entry points record CUDA events on the caller's stream at
 * their stage boundaries (one timed call at a time).  f2p_stage_times waits for the last timed call and returns
 *   ms[0] projection (both sides, norms fused)   ms[1] kernel 2 (key fill + gated arg-min / top-5 + finalise)
 *   ms[2] filter init   ms[3], ms[5], ms[7] vote of round 0, 1, 2   ms[4], ms[6], ms[8] combine of round 0, 1, 2
 * (zeros for stages the call did not run) and *filter_evals = sum over pairs and rounds of M_r^2, the (a,b) distance
 * pair evaluations the reference performs for that batch (spacial_filtering.py:52-64; 0 for the naive solver), and
 * *ambiguous_cols = columns (summed over pairs and rounds) whose screened score fell inside the error band of a threshold
 * and were decided by the ordered arithmetic instead (csrc/filter.cu).  Either pointer may be NULL. */
/* Test hook: worst relative error of the screened vote's square root (sqrt.approx.ftz.f32) over the `count` floats
 * with bit patterns first_bits, first_bits + stride, ...; the error band of csrc/filter.cu assumes <= 2^-21. */
int f2p_test_sqrt_approx_error(uint32_t first_bits, uint32_t count, uint32_t stride, double* max_rel_err, void* stream);

/* Test hook (negative control of tests/test_gpu_parity.py): overrides the screened vote's error band (a fraction of the
 * pair's extent D) for later calls of this process; band < 0 restores the proven value.  A band of 0 lets the cheap
 * scores decide alone and CAN change the survivor set — nothing else in the library can, and no environment variable
 * reaches it. */
int f2p_test_vote_band(double band);

/* Environment switches read by the library — all NON-semantic (every setting returns the same, exact result; they select
 * between equivalent kernels or tilings for tests and tools/):
 *   F2P_NN_PATH = tc | gated | exact | two_phase   kernel-2 path (default: tensor-pipe screen for large problems, gated FP32 else)
 *   F2P_GATED_KC, F2P_GATED_NXP, F2P_NSPLIT, F2P_TOP5_MIN_TILES   tiling of the FP32 kernel-2 instances
 *   F2P_VOTE_SCREEN = 0   ordered vote for every column;   F2P_VOTE_RB = 1 | 2 | 4   rows per basic block of the vote tile */

#define F2P_STAGE_SLOTS 9
int f2p_stage_timing(int enable);
int f2p_stage_times(float ms[F2P_STAGE_SLOTS], double* filter_evals, double* ambiguous_cols);
/* Work the vote kernels of the last timed call actually EXECUTED: sum over pairs and over the rounds that ran of
 * M_r (M_r + 1) / 2 evaluations (each unordered pair once; a round whose predecessor pruned nothing is skipped), and the
 * number of (pair, round) votes skipped.  Synchronous; call after f2p_stage_times. */
int f2p_stage_filter_executed(double* executed_evals, double* skipped_rounds);

#ifdef __cplusplus
}
#endif
#endif /* F2P_H_ */

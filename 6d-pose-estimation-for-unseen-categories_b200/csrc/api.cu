// api.cu — the extern "C" surface of libf2p_sm100.so (declared in include/f2p.h): argument checks,
// workspace carving, work decomposition and kernel sequencing.  No torch types, no CPU compute path.
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>
#include <mutex>
#include <vector>
#include "common.cuh"
#include "internal.h"

namespace f2p {

// Row tiles an item of the top-5 kernel streams at least (big batches): its bootstrap costs 12 list cuts, so a CTA
// should amortise it over more tiles than the arg-min kernel needs (F2P_TOP5_MIN_TILES overrides, tools/).
static int top5_min_tiles() {
    static const int v = [] { const char* e = getenv("F2P_TOP5_MIN_TILES"); const int x = e ? atoi(e) : 40; return x > 0 ? x : 40; }();
    return v;
}


static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
    va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof(g_err), fmt, ap); va_end(ap);
}
const char* get_error() { return g_err; }

// ---- TMA descriptor encoding through the driver entry point (no link-time libcuda dependency) ---------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    });
    return fn;
}

int make_tensor_map_2d(CUtensorMap* map, const float* base, uint64_t cols, uint64_t rows, uint64_t row_stride_elems,
                       uint32_t box_cols, uint32_t box_rows, int swizzle128) {
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) { set_error("cuTensorMapEncodeTiled entry point unavailable"); return F2P_EDEVICE; }
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {row_stride_elems * sizeof(float)};
    cuuint32_t box[2] = {box_cols, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed (%d) cols=%llu rows=%llu", (int)r,
                                       (unsigned long long)cols, (unsigned long long)rows); return F2P_EDEVICE; }
    return 0;
}

// ---- stage timing (bench.py's live per-kernel breakdown; off by default) -------------------------------
struct StageState {
    bool on = false, created = false;
    cudaEvent_t ev[STAGE_MARKS];
    bool have[STAGE_MARKS] = {};
    std::vector<int*> cnt;
};
static StageState g_stage;
static std::mutex g_stage_mu;
bool stage_timing_on() { return g_stage.on; }
static void stage_begin() {
    if (!g_stage.on) return;
    std::lock_guard<std::mutex> lk(g_stage_mu);
    for (int i = 0; i < STAGE_MARKS; ++i) g_stage.have[i] = false;
    g_stage.cnt.clear();
}
void stage_mark(int mark, cudaStream_t st) {
    if (!g_stage.on || mark < 0 || mark >= STAGE_MARKS) return;
    std::lock_guard<std::mutex> lk(g_stage_mu);
    if (!g_stage.created) return;
    if (cudaEventRecord(g_stage.ev[mark], st) == cudaSuccess) g_stage.have[mark] = true;
}
void stage_filter_counts(int* const* cnt_dev, int npairs) {
    if (!g_stage.on) return;
    std::lock_guard<std::mutex> lk(g_stage_mu);
    g_stage.cnt.assign(cnt_dev, cnt_dev + npairs);
}

// SM count of the CURRENT device, cached per device id (a process may drive several GPUs)
static int sm_count() {
    static int cache[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cache[dev] == 0) {
        int n = 0;
        cache[dev] = (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0) ? n : 148;
    }
    return cache[dev];
}

// ---- batch plan: host-side decomposition into tiles / work items and the workspace layout --------------
struct Plan {
    std::vector<PairDev> pairs;
    int64_t ld1_total = 0, ld2_total = 0;
    int total_xtiles = 0, total_ytiles = 0, total_items = 0;
    int uni_xtiles = 0, uni_ytiles = 0, uni_items = 0;   // >0 when every pair has the same count
    int64_t cand_total = 0;                              // sum of 5*n2
    int maxM = 0;
    double flops = 0;                                    // 2*n1*n2*k summed over pairs
    // workspace offsets (bytes)
    size_t off_pairs = 0, off_fpairs = 0, off_xT = 0, off_nx = 0, off_yT = 0, off_ny = 0, off_keys = 0, off_cand = 0,
           off_filter = 0, off_tc = 0, total_bytes = 0;
    std::vector<size_t> filter_off;
    // tensor-pipe screen (tc.cu)
    bool use_tc = false;
    int spatial = 0;
    TCPlan tc;
};

static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

static int build_plan(const F2PPairDesc* d, int npairs, int k, int spatial, bool own_cand, Plan& P) {
    if (npairs < 0 || (npairs > 0 && !d)) { set_error("null pair descriptors"); return F2P_EINVAL; }
    if (k < 1) { set_error("k must be >= 1 (got %d)", k); return F2P_EINVAL; }
    if (k > F2P_MAX_K) { set_error("k=%d exceeds F2P_MAX_K=%d (the reference's sequential-k contract)", k, F2P_MAX_K); return F2P_ENOTSUP; }
    P.pairs.resize(npairs);
    int64_t base_items = 0;
    for (int i = 0; i < npairs; ++i) {
        if (d[i].n1 < 1 || d[i].n2 < 0) { set_error("pair %d: n1=%d n2=%d", i, d[i].n1, d[i].n2); return F2P_EINVAL; }
        if (spatial && d[i].n1 < F2P_TOPK) { set_error("pair %d: spatial solver needs n1 >= 5 (got %d)", i, d[i].n1); return F2P_ENOTSUP; }
        if (spatial && 5ll * d[i].n2 > 524288) { set_error("pair %d: 5*n2 > 524288 candidates unsupported", i); return F2P_ENOTSUP; }
        base_items += (d[i].n2 + TILE - 1) / TILE;
    }
    // the tensor-pipe screen decides large calls; the FP32 kernels then only see the tiles it flags
    double flops_all = 0.0;
    for (int i = 0; i < npairs; ++i) flops_all += 2.0 * d[i].n1 * (double)d[i].n2 * k;
    const bool tc_call = npairs > 0 && tc_enabled(flops_all, k, spatial);
    for (int i = 0; i < npairs; ++i) {
        PairDev& p = P.pairs[i];
        memset(&p, 0, sizeof(p));
        p.ex_off = d[i].evecs_x_off; p.ey_off = d[i].evecs_y_off; p.c_off = d[i].c_off;
        p.cad_off = d[i].cad_off; p.pc_off = d[i].pc_off; p.out_off = d[i].out_off;
        p.n1 = d[i].n1; p.n2 = d[i].n2; p.ldx = d[i].ldx; p.ldy = d[i].ldy; p.ncad = d[i].ncad; p.npc = d[i].npc;
        memcpy(p.thr, d[i].thr, sizeof(p.thr));
        p.nrt = (p.n1 + TILE - 1) / TILE; p.nct = (p.n2 + TILE - 1) / TILE;
        int ns = choose_nsplit(base_items, p.nrt, sm_count(), spatial ? top5_min_tiles() : 16);
        if (tc_call) ns = fallback_nsplit(ns, base_items, sm_count());
        p.nsplit = ns;
        if (P.ld1_total + p.nrt * (int64_t)TILE > 0x7fffff00ll || P.ld2_total + p.nct * (int64_t)TILE > 0x7fffff00ll ||
            (int64_t)P.total_items + (int64_t)p.nct * ns > 0x7fffff00ll) {
            set_error("batch too large for 32-bit tile indexing; split the batch"); return F2P_ENOTSUP;
        }
        p.x_col0 = (int)P.ld1_total; p.y_col0 = (int)P.ld2_total;
        p.xtile0 = P.total_xtiles; p.ytile0 = P.total_ytiles; p.item0 = P.total_items;
        p.cand0 = (int)0;
        P.ld1_total += (int64_t)p.nrt * TILE; P.ld2_total += (int64_t)p.nct * TILE;
        P.total_xtiles += p.nrt; P.total_ytiles += p.nct; P.total_items += p.nct * ns;
        if (P.cand_total > 0x3fffffffll) { set_error("too many candidates in one batch; split the batch"); return F2P_ENOTSUP; }
        p.cand0 = (int)P.cand_total;
        P.cand_total += 5ll * p.n2;
        if (5 * p.n2 > P.maxM) P.maxM = 5 * p.n2;
        P.flops += 2.0 * p.n1 * (double)p.n2 * k;
    }
    // tensor-pipe screen: its own row tiling (256-row tiles, one CTA per SM)
    P.spatial = spatial;
    P.use_tc = tc_call;
    P.tc = TCPlan();
    if (P.use_tc) {
        int64_t rows = 0, items = 0, xt = 0;
        for (int i = 0; i < npairs; ++i) {
            PairDev& p = P.pairs[i];
            p.tc_nrt = (p.n1 + TC_ROWS - 1) / TC_ROWS;
            p.tc_nsplit = choose_tc_nsplit(base_items, p.tc_nrt, sm_count());
            if (rows + (int64_t)p.tc_nrt * TC_ROWS > 0x7fffff00ll || (items + (int64_t)p.nct * p.tc_nsplit) * TC_CG > 0x7fffff00ll) {
                set_error("batch too large for 32-bit tile indexing; split the batch"); return F2P_ENOTSUP;
            }
            p.xa_row0 = (int)rows; p.tc_item0 = (int)items; p.tc_xtile0 = (int)xt;
            rows += (int64_t)p.tc_nrt * TC_ROWS; items += (int64_t)((p.nct + TC_CG - 1) / TC_CG) * p.tc_nsplit; xt += 2 * p.tc_nrt;
        }
        P.tc.ld1_tc_total = rows; P.tc.total_items = (int)items; P.tc.total_xtiles = (int)xt;
    }
    auto uniform = [&](auto getter) {
        if (npairs == 0) return 0;
        const int v = getter(P.pairs[0]);
        for (int i = 1; i < npairs; ++i) if (getter(P.pairs[i]) != v) return 0;
        return v;
    };
    P.uni_xtiles = uniform([](const PairDev& p) { return p.nrt; });
    P.uni_ytiles = uniform([](const PairDev& p) { return p.nct; });
    P.uni_items = uniform([](const PairDev& p) { return p.nct * p.nsplit; });
    if (P.use_tc) {
        P.tc.uni_items = uniform([](const PairDev& p) { return ((p.nct + TC_CG - 1) / TC_CG) * p.tc_nsplit; });
        P.tc.uni_xtiles = uniform([](const PairDev& p) { return 2 * p.tc_nrt; });
    }

    size_t o = 0;
    P.off_pairs = o; o = align_up(o + sizeof(PairDev) * (size_t)(npairs > 0 ? npairs : 1), 256);
    P.off_fpairs = o; o = align_up(o + (spatial ? filter_table_bytes(npairs) : 0), 256);
    P.off_xT = o; o = align_up(o + sizeof(float) * (size_t)k * (size_t)P.ld1_total, 256);
    P.off_nx = o; o = align_up(o + sizeof(float) * (size_t)P.ld1_total, 256);
    P.off_yT = o; o = align_up(o + sizeof(float) * (size_t)k * (size_t)P.ld2_total, 256);
    P.off_ny = o; o = align_up(o + sizeof(float) * (size_t)P.ld2_total, 256);
    P.off_keys = o;
    o = align_up(o + (spatial ? top5_ws_bytes(P.total_items, P.total_ytiles) : argmin_ws_bytes(P.ld2_total, k, npairs)), 256);
    if (spatial) {
        P.off_cand = o;
        if (own_cand) o = align_up(o + sizeof(int64_t) * 2 * (size_t)P.cand_total, 256);
        P.off_filter = o;
        P.filter_off.resize(npairs);
        for (int i = 0; i < npairs; ++i) { P.filter_off[i] = o; o += filter_pair_bytes(5 * P.pairs[i].n2); }
    }
    if (P.use_tc) {
        P.off_tc = o;
        o = align_up(o + tc_ws_bytes(P.tc.ld1_tc_total, P.ld2_total, k, P.tc.total_items, P.total_ytiles, npairs, spatial), 256);
    }
    P.total_bytes = o;
    return 0;
}

// carves the tensor-pipe workspace of a call; returns the plan to hand to launch_argmin / launch_top5 (or null)
static const TCPlan* tc_plan(Plan& P, unsigned char* ws, int k, int npairs) {
    if (!P.use_tc) return nullptr;
    tc_ws_carve(P.tc.ws, ws + P.off_tc, P.tc.ld1_tc_total, P.ld2_total, k, P.tc.total_items, P.total_ytiles, npairs, P.spatial);
    return &P.tc;
}

static int check_ws(const Plan& P, void* ws, size_t ws_bytes) {
    if (!ws && P.total_bytes > 0) { set_error("workspace pointer is null"); return F2P_EINVAL; }
    if ((reinterpret_cast<uintptr_t>(ws) & 255) != 0) { set_error("workspace must be 256-byte aligned"); return F2P_EINVAL; }
    if (ws_bytes < P.total_bytes) { set_error("workspace too small: need %zu bytes, got %zu", P.total_bytes, ws_bytes); return F2P_ENOSPACE; }
    return 0;
}

// projection of both operands + norms for the whole batch
// (one launch; `aw` non-null: also resets the arg-min merge keys and arrival counters of every column tile)
static int run_prep(const Plan& P, const float* ex, const float* ey, const float* C, int npairs, int k,
                    unsigned char* ws, cudaStream_t st, const ArgminWs* aw = nullptr) {
    const PairDev* pd = reinterpret_cast<const PairDev*>(ws + P.off_pairs);
    return launch_prep(ex, C, ey, pd, npairs, P.uni_xtiles, P.total_xtiles, P.uni_ytiles, P.total_ytiles, k,
                       reinterpret_cast<float*>(ws + P.off_xT), P.ld1_total, reinterpret_cast<float*>(ws + P.off_nx),
                       reinterpret_cast<float*>(ws + P.off_yT), P.ld2_total, reinterpret_cast<float*>(ws + P.off_ny),
                       aw ? aw->best : nullptr, aw ? aw->tilecnt : nullptr, st);
}

}  // namespace f2p

using namespace f2p;

extern "C" {

int f2p_version(void) { return F2P_VERSION; }
const char* f2p_last_error(void) { return get_error(); }

int f2p_device_info(int* sm, int* major, int* minor, size_t* smem_optin) {
    int dev = 0;
    F2P_CUDA(cudaGetDevice(&dev));
    cudaDeviceProp p;
    F2P_CUDA(cudaGetDeviceProperties(&p, dev));
    if (sm) *sm = p.multiProcessorCount;
    if (major) *major = p.major;
    if (minor) *minor = p.minor;
    if (smem_optin) *smem_optin = p.sharedMemPerBlockOptin;
    return 0;
}

size_t f2p_workspace_bytes(const F2PPairDesc* descs, int npairs, int k, int spatial) {
    Plan P;
    if (build_plan(descs, npairs, k, spatial, true, P)) return 0;
    return P.total_bytes;
}

int f2p_naive_batched(const float* ex, const float* ey, const float* C, const F2PPairDesc* descs, int npairs, int k,
                      unsigned flags, int64_t* out, void* ws_v, size_t ws_bytes, void* stream) {
    if (npairs == 0) return 0;
    if (!ex || !ey || !C || !out) { set_error("null device pointer"); return F2P_EINVAL; }
    Plan P;
    int rc = build_plan(descs, npairs, k, 0, false, P);
    if (rc) return rc;
    if ((rc = check_ws(P, ws_v, ws_bytes))) return rc;
    unsigned char* ws = static_cast<unsigned char*>(ws_v);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    F2P_CUDA(cudaMemcpyAsync(ws + P.off_pairs, P.pairs.data(), sizeof(PairDev) * (size_t)npairs, cudaMemcpyHostToDevice, st));
    stage_begin();
    stage_mark(0, st);
    ArgminWs aw;
    argmin_ws_carve(aw, ws + P.off_keys, P.ld2_total, k, npairs);
    if ((rc = run_prep(P, ex, ey, C, npairs, k, ws, st, &aw))) return rc;
    stage_mark(1, st);
    const PairDev* pd = reinterpret_cast<const PairDev*>(ws + P.off_pairs);
    rc = launch_argmin(reinterpret_cast<float*>(ws + P.off_xT), P.ld1_total, reinterpret_cast<float*>(ws + P.off_nx),
                         reinterpret_cast<float*>(ws + P.off_yT), P.ld2_total, reinterpret_cast<float*>(ws + P.off_ny), k, pd,
                         npairs, P.uni_items, P.total_items, P.uni_ytiles, P.total_ytiles, P.flops, flags, out, aw, st, true,
                         tc_plan(P, ws, k, npairs));
    stage_mark(2, st);
    return rc;
}

int f2p_spatial_batched(const float* ex, const float* ey, const float* C, const float* cad, const float* pc,
                        const F2PPairDesc* descs, int npairs, int k, unsigned flags, int64_t* out, int32_t* counts,
                        int32_t* status, int64_t* cand_dev, void* ws_v, size_t ws_bytes, void* stream) {
    if (npairs == 0) return 0;
    if (!ex || !ey || !C || !cad || !pc || !out || !counts || !status) { set_error("null device pointer"); return F2P_EINVAL; }
    Plan P;
    int rc = build_plan(descs, npairs, k, 1, cand_dev == nullptr, P);
    if (rc) return rc;
    if ((rc = check_ws(P, ws_v, ws_bytes))) return rc;
    unsigned char* ws = static_cast<unsigned char*>(ws_v);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    F2P_CUDA(cudaMemcpyAsync(ws + P.off_pairs, P.pairs.data(), sizeof(PairDev) * (size_t)npairs, cudaMemcpyHostToDevice, st));
    stage_begin();
    stage_mark(0, st);
    if ((rc = run_prep(P, ex, ey, C, npairs, k, ws, st))) return rc;
    stage_mark(1, st);
    const PairDev* pd = reinterpret_cast<const PairDev*>(ws + P.off_pairs);
    u64* part5 = reinterpret_cast<u64*>(ws + P.off_keys);
    // candidates: caller's buffer uses the `out` layout (out_off, [2][5*n2]); the internal one is packed by cand0
    int64_t* cand = cand_dev ? cand_dev : reinterpret_cast<int64_t*>(ws + P.off_cand);
    rc = launch_top5(reinterpret_cast<float*>(ws + P.off_xT), P.ld1_total, reinterpret_cast<float*>(ws + P.off_nx),
                     reinterpret_cast<float*>(ws + P.off_yT), P.ld2_total, reinterpret_cast<float*>(ws + P.off_ny), k, pd,
                     npairs, P.uni_items, P.total_items, P.uni_ytiles, P.total_ytiles, flags, part5, cand,
                     cand_dev != nullptr, st, tc_plan(P, ws, k, npairs));
    if (rc) return rc;
    stage_mark(2, st);
    return run_filter_batch(P.pairs.data(), npairs, cad, pc, cand, cand_dev != nullptr, out, counts, status,
                            ws + P.off_fpairs, ws, P.filter_off.data(), st);
}

int f2p_stage_timing(int enable) {
    std::lock_guard<std::mutex> lk(g_stage_mu);
    if (enable && !g_stage.created) {
        for (int i = 0; i < STAGE_MARKS; ++i) F2P_CUDA(cudaEventCreate(&g_stage.ev[i]));
        g_stage.created = true;
    }
    g_stage.on = enable != 0;
    for (int i = 0; i < STAGE_MARKS; ++i) g_stage.have[i] = false;
    g_stage.cnt.clear();
    return 0;
}

int f2p_stage_filter_executed(double* executed_evals, double* skipped_rounds) {
    if (!executed_evals) { set_error("null pointer"); return F2P_EINVAL; }
    std::lock_guard<std::mutex> lk(g_stage_mu);
    double s = 0.0, sk = 0.0;
    for (int* p : g_stage.cnt) {
        int c[9];
        F2P_CUDA(cudaMemcpy(c, p, sizeof(c), cudaMemcpyDeviceToHost));
        for (int r = 0; r < 3; ++r) {
            if (r > 0 && c[r] == c[r - 1]) { sk += 1.0; continue; }     // predecessor pruned nothing: the vote is skipped
            s += 0.5 * (double)c[r] * ((double)c[r] + 1.0);              // each unordered pair (and the diagonal) once
        }
    }
    *executed_evals = s;
    if (skipped_rounds) *skipped_rounds = sk;
    return 0;
}

int f2p_stage_times(float* ms, double* filter_evals, double* ambiguous_cols) {
    if (!ms) { set_error("null pointer"); return F2P_EINVAL; }
    std::lock_guard<std::mutex> lk(g_stage_mu);
    for (int i = 0; i < STAGE_MARKS - 1; ++i) ms[i] = 0.f;
    if (filter_evals) *filter_evals = 0.0;
    if (ambiguous_cols) *ambiguous_cols = 0.0;
    if (!g_stage.created || !g_stage.have[0]) { set_error("no timed call recorded (f2p_stage_timing(1) first)"); return F2P_EINVAL; }
    int last = 0;
    for (int i = 0; i < STAGE_MARKS; ++i) if (g_stage.have[i]) last = i;
    F2P_CUDA(cudaEventSynchronize(g_stage.ev[last]));
    for (int i = 0; i + 1 < STAGE_MARKS; ++i)
        if (g_stage.have[i] && g_stage.have[i + 1]) F2P_CUDA(cudaEventElapsedTime(&ms[i], g_stage.ev[i], g_stage.ev[i + 1]));
    if (filter_evals || ambiguous_cols) {
        double s = 0.0, amb = 0.0;
        for (int* p : g_stage.cnt) {
            int c[9];
            F2P_CUDA(cudaMemcpy(c, p, sizeof(c), cudaMemcpyDeviceToHost));
            for (int r = 0; r < 3; ++r) { s += (double)c[r] * (double)c[r]; amb += (double)c[6 + r]; }
        }
        if (filter_evals) *filter_evals = s;
        if (ambiguous_cols) *ambiguous_cols = amb;
    }
    return 0;
}

}  // extern "C"

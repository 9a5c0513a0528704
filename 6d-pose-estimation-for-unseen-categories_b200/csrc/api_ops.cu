// api_ops.cu — single-kernel C-ABI entry points (tests, roofline bench, host-side composition) and the
// host-buffer convenience calls.  Declared in include/f2p.h.
#include <string.h>
#include <mutex>
#include <vector>
#include "common.cuh"
#include "internal.h"

using namespace f2p;

namespace {

int single_pair(PairDev& p, int n1, int n2, int ld1, int ld2, int nsplit_hint) {
    memset(&p, 0, sizeof(p));
    p.n1 = n1; p.n2 = n2;
    p.nrt = (n1 + TILE - 1) / TILE; p.nct = (n2 + TILE - 1) / TILE;
    if (ld1 < p.nrt * TILE || ld2 < p.nct * TILE || (ld1 % TILE) || (ld2 % TILE)) {
        set_error("leading dimensions must be multiples of %d covering the padded extents (ld1=%d ld2=%d)", TILE, ld1, ld2);
        return F2P_EINVAL;
    }
    p.nsplit = nsplit_hint;
    return 0;
}

int sm_count_now() {
    int dev = 0, sm = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, dev);
    return sm > 0 ? sm : 148;
}

int pick_nsplit(int nrt, int nct) { return choose_nsplit(nct, nrt, sm_count_now()); }

// tensor-pipe screen geometry of a single pair (tc.cu); returns false when the screen is not used for this shape
bool single_pair_tc(PairDev& p, int k, int top5, TCPlan& tc) {
    if (!tc_enabled(2.0 * p.n1 * (double)p.n2 * k, k, top5)) return false;
    p.tc_nrt = (p.n1 + TC_ROWS - 1) / TC_ROWS;
    p.tc_nsplit = choose_tc_nsplit(p.nct, p.tc_nrt, sm_count_now());
    p.tc_item0 = 0; p.tc_xtile0 = 0; p.xa_row0 = 0;
    tc.total_items = tc.uni_items = ((p.nct + TC_CG - 1) / TC_CG) * p.tc_nsplit;
    tc.total_xtiles = tc.uni_xtiles = 2 * p.tc_nrt;
    tc.ld1_tc_total = (int64_t)p.tc_nrt * TC_ROWS;
    return true;
}

}  // namespace

extern "C" {

size_t f2p_project_workspace_bytes(void) { return 256; }

int f2p_project(const float* evecs, int n, int k, int ld_in, const float* C, float* xT, float* nrm, int ld,
                void* ws, size_t ws_bytes, void* stream) {
    if (!evecs || !xT || !nrm || !ws) { set_error("null pointer"); return F2P_EINVAL; }
    if (k < 1 || k > F2P_MAX_K || n < 0 || ld_in < k) { set_error("bad shape n=%d k=%d ld_in=%d", n, k, ld_in); return k > F2P_MAX_K ? F2P_ENOTSUP : F2P_EINVAL; }
    if (ws_bytes < 256) { set_error("workspace too small"); return F2P_ENOSPACE; }
    PairDev p;
    const bool is_x = (C != nullptr);
    int rc = single_pair(p, is_x ? n : 1, is_x ? 0 : n, is_x ? ld : TILE, is_x ? 0 : ld, 1);
    if (rc) return rc;
    p.ldx = ld_in; p.ldy = ld_in;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    F2P_CUDA(cudaMemcpyAsync(ws, &p, sizeof(p), cudaMemcpyHostToDevice, st));
    const int tiles = is_x ? p.nrt : p.nct;
    return launch_project(is_x, evecs, C, static_cast<const PairDev*>(ws), 1, tiles, tiles, k, xT, ld, nrm, st);
}

size_t f2p_nn_workspace_bytes(int n1, int n2, int k, int top5) {
    const int nrt = (n1 + TILE - 1) / TILE, nct = (n2 + TILE - 1) / TILE;
    const int ns = pick_nsplit(nrt, nct);
    size_t b = 256;
    if (top5) b += top5_ws_bytes((int64_t)nct * ns, nct);
    else b += argmin_ws_bytes((int64_t)nct * TILE, k, 1);
    b = (b + 255) & ~(size_t)255;
    PairDev p; memset(&p, 0, sizeof(p));
    p.n1 = n1; p.n2 = n2; p.nct = nct;
    TCPlan tc;
    if (n1 >= 1 && n2 >= 1 && single_pair_tc(p, k, top5, tc))
        b += tc_ws_bytes(tc.ld1_tc_total, (int64_t)nct * TILE, k, tc.total_items, nct, 1, top5);
    return (b + 255) & ~(size_t)255;
}

static int nn_common(int mode, const float* xT, const float* nx, int n1, int ld1, const float* yT, const float* ny,
                     int n2, int ld2, int k, int64_t* out, void* ws_v, size_t ws_bytes, void* stream) {
    if (!xT || !nx || !yT || !ny || !out || !ws_v) { set_error("null pointer"); return F2P_EINVAL; }
    if (k < 1 || n1 < 1 || n2 < 0) { set_error("bad shape n1=%d n2=%d k=%d", n1, n2, k); return F2P_EINVAL; }
    if (k > F2P_MAX_K) { set_error("k=%d exceeds F2P_MAX_K", k); return F2P_ENOTSUP; }
    if (mode == 1 && n1 < F2P_TOPK) { set_error("top-5 needs n1 >= 5"); return F2P_ENOTSUP; }
    if (n2 == 0) return 0;
    if (ws_bytes < f2p_nn_workspace_bytes(n1, n2, k, mode)) { set_error("workspace too small"); return F2P_ENOSPACE; }
    PairDev p;
    const int nrt = (n1 + TILE - 1) / TILE, nct = (n2 + TILE - 1) / TILE;
    int ns = pick_nsplit(nrt, nct);
    // arg-min decided by the tensor-pipe screen: the FP32 launch only visits flagged tiles (its workspace does not depend on ns)
    if (mode == 0 && ld2 == nct * TILE && tc_enabled(2.0 * n1 * (double)n2 * k, k, 0)) ns = fallback_nsplit(ns, nct, sm_count_now());
    int rc = single_pair(p, n1, n2, ld1, ld2, ns);
    if (rc) return rc;
    if (mode == 0 && ld2 != nct * TILE) {
        set_error("f2p_nn_argmin needs ld2 == round_up(n2, %d) (the re-check operand shares the layout)", TILE);
        return F2P_EINVAL;
    }
    unsigned char* ws = static_cast<unsigned char*>(ws_v);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int items = p.nct * p.nsplit;
    TCPlan tc;
    const bool use_tc = ld2 == nct * TILE && single_pair_tc(p, k, mode, tc);
    if (use_tc) {
        size_t off = 256 + (mode ? top5_ws_bytes(items, p.nct) : argmin_ws_bytes(ld2, k, 1));
        off = (off + 255) & ~(size_t)255;
        tc_ws_carve(tc.ws, ws + off, tc.ld1_tc_total, ld2, k, tc.total_items, p.nct, 1, mode);
    }
    F2P_CUDA(cudaMemcpyAsync(ws, &p, sizeof(p), cudaMemcpyHostToDevice, st));
    const PairDev* pd = reinterpret_cast<const PairDev*>(ws);
    if (mode == 0) {
        ArgminWs aw;
        argmin_ws_carve(aw, ws + 256, ld2, k, 1);
        return launch_argmin(xT, ld1, nx, yT, ld2, ny, k, pd, 1, items, items, p.nct, p.nct, 2.0 * n1 * (double)n2 * k,
                             F2P_ROW0_ONLY, out, aw, st, false, use_tc ? &tc : nullptr);
    }
    return launch_top5(xT, ld1, nx, yT, ld2, ny, k, pd, 1, items, items, p.nct, p.nct, F2P_ROW0_ONLY,
                       reinterpret_cast<unsigned long long*>(ws + 256), out, 1, st, use_tc ? &tc : nullptr);
}

int f2p_nn_argmin(const float* xT, const float* nx, int n1, int ld1, const float* yT, const float* ny, int n2, int ld2,
                  int k, int64_t* idx_out, void* ws, size_t ws_bytes, void* stream) {
    return nn_common(0, xT, nx, n1, ld1, yT, ny, n2, ld2, k, idx_out, ws, ws_bytes, stream);
}
int f2p_nn_top5(const float* xT, const float* nx, int n1, int ld1, const float* yT, const float* ny, int n2, int ld2,
                int k, int64_t* idx5_out, void* ws, size_t ws_bytes, void* stream) {
    return nn_common(1, xT, nx, n1, ld1, yT, ny, n2, ld2, k, idx5_out, ws, ws_bytes, stream);
}

// ---- allocation-free arg-min (no counterpart in the reference) -----------
// For callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc): the
// scratch of kernel 2 (merge keys, re-check buffers, the pair record) lives in ONE file-scope device array instead of a
// caller-provided workspace.  Every GPU that loads the library holds these 2.25 MB until exit.  All calls share the
// array and it keeps the last call's contents: the launch sequence re-initialises everything it reads (argmin_init_kernel
// resets the merge keys / counters of the n2 columns in use), and calls that overlap on the GPU (two streams, two
// threads) RACE on it — serialise them.
constexpr size_t F2P_STATIC_SCRATCH_BYTES = 2304 * 1024;      // >= 256 + argmin_ws_bytes(4096, 128, 1) = 2 212 608
static __device__ __align__(256) unsigned char g_static_scratch[F2P_STATIC_SCRATCH_BYTES];

// In host code `g_static_scratch` is only a stand-in for the symbol, NOT a device address: the address is looked up
// once per device (with that device current) and cached.
static int static_scratch_ptr(unsigned char** out) {
    static unsigned char* cache[64] = {nullptr};
    static std::mutex mu;
    int dev = 0;
    F2P_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) { set_error("device index %d out of range", dev); return F2P_EDEVICE; }
    std::lock_guard<std::mutex> lk(mu);
    if (!cache[dev]) {
        void* p = nullptr;
        F2P_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));
        if (!p) { set_error("cudaGetSymbolAddress returned a null address"); return F2P_EDEVICE; }
        cache[dev] = static_cast<unsigned char*>(p);
    }
    *out = cache[dev];
    return 0;
}

int f2p_nn_argmin_static(const float* xT, const float* nx, int n1, int ld1, const float* yT, const float* ny, int n2,
                         int ld2, int k, int64_t* idx_out, void* stream) {
    if (!xT || !nx || !yT || !ny || !idx_out) { set_error("null pointer"); return F2P_EINVAL; }
    if (k < 1 || n1 < 1 || n2 < 0) { set_error("bad shape n1=%d n2=%d k=%d", n1, n2, k); return F2P_EINVAL; }
    if (n2 > F2P_STATIC_MAX_N2 || k > F2P_STATIC_MAX_K) {
        set_error("f2p_nn_argmin_static: n2=%d k=%d exceed the static scratch (n2 <= %d, k <= %d); use f2p_nn_argmin "
                  "with a caller-provided workspace", n2, k, F2P_STATIC_MAX_N2, F2P_STATIC_MAX_K);
        return F2P_ENOTSUP;
    }
    if (n2 == 0) return 0;
    PairDev p;
    const int nrt = (n1 + TILE - 1) / TILE, nct = (n2 + TILE - 1) / TILE;
    int rc = single_pair(p, n1, n2, ld1, ld2, pick_nsplit(nrt, nct));
    if (rc) return rc;
    if (ld2 != nct * TILE) { set_error("f2p_nn_argmin_static needs ld2 == round_up(n2, %d)", TILE); return F2P_EINVAL; }
    if (256 + argmin_ws_bytes(ld2, k, 1) > F2P_STATIC_SCRATCH_BYTES) {      // cannot happen within the limits above
        set_error("f2p_nn_argmin_static: scratch of %zu bytes exceeds the static array", 256 + argmin_ws_bytes(ld2, k, 1));
        return F2P_ENOSPACE;
    }
    unsigned char* ws = nullptr;
    if ((rc = static_scratch_ptr(&ws))) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    F2P_CUDA(cudaMemcpyAsync(ws, &p, sizeof(p), cudaMemcpyHostToDevice, st));
    const int items = p.nct * p.nsplit;
    ArgminWs aw;
    argmin_ws_carve(aw, ws + 256, ld2, k, 1);
    // always the FP32 kernels (the tensor-pipe screen needs a workspace that grows with n1); pre_init = false: the
    // sequence starts with argmin_init_kernel, so nothing of an earlier call is read
    return launch_argmin(xT, ld1, nx, yT, ld2, ny, k, reinterpret_cast<const PairDev*>(ws), 1, items, items, p.nct, p.nct,
                         2.0 * n1 * (double)n2 * k, F2P_ROW0_ONLY, idx_out, aw, st, false, nullptr);
}

// ---- blocked arg-min (host pipelines that overlap the upload of a huge pair with its computation) ------------------
int f2p_nn_argmin_keys(const float* xT, const float* nx, int n1, int ld1, int row0, const float* yT, const float* ny, int n2,
                       int ld2, int k, unsigned long long* keys, void* ws_v, size_t ws_bytes, void* stream) {
    if (!xT || !nx || !yT || !ny || !keys || !ws_v) { set_error("null pointer"); return F2P_EINVAL; }
    if (k < 1 || n1 < 1 || n2 < 0 || row0 < 0) { set_error("bad shape n1=%d n2=%d k=%d row0=%d", n1, n2, k, row0); return F2P_EINVAL; }
    if (k > F2P_MAX_K) { set_error("k=%d exceeds F2P_MAX_K", k); return F2P_ENOTSUP; }
    if ((long long)row0 + n1 > 0x7ffffff0ll) { set_error("row0 + n1 exceeds the 31-bit row index"); return F2P_ENOTSUP; }
    if (n2 == 0) return 0;
    if (ws_bytes < f2p_nn_workspace_bytes(n1, n2, k, 0)) { set_error("workspace too small"); return F2P_ENOSPACE; }
    PairDev p;
    const int nrt = (n1 + TILE - 1) / TILE, nct = (n2 + TILE - 1) / TILE;
    int ns = pick_nsplit(nrt, nct);
    if (tc_enabled(2.0 * n1 * (double)n2 * k, k, 0)) ns = fallback_nsplit(ns, nct, sm_count_now());
    int rc = single_pair(p, n1, n2, ld1, ld2, ns);
    if (rc) return rc;
    if (ld2 != nct * TILE) { set_error("f2p_nn_argmin_keys needs ld2 == round_up(n2, %d)", TILE); return F2P_EINVAL; }
    unsigned char* ws = static_cast<unsigned char*>(ws_v);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int items = p.nct * p.nsplit;
    TCPlan tc;
    const bool use_tc = single_pair_tc(p, k, 0, tc);
    if (use_tc) {
        size_t off = (256 + argmin_ws_bytes(ld2, k, 1) + 255) & ~(size_t)255;
        tc_ws_carve(tc.ws, ws + off, tc.ld1_tc_total, ld2, k, tc.total_items, p.nct, 1, 0);
    }
    F2P_CUDA(cudaMemcpyAsync(ws, &p, sizeof(p), cudaMemcpyHostToDevice, st));
    ArgminWs aw;
    argmin_ws_carve(aw, ws + 256, ld2, k, 1);
    return launch_argmin_keys(xT, ld1, nx, yT, ld2, ny, k, reinterpret_cast<const PairDev*>(ws), items, p.nct, n2,
                              (uint32_t)row0, keys, aw, st, use_tc ? &tc : nullptr);
}

int f2p_keys_to_map(const unsigned long long* keys, int n2, unsigned flags, int64_t* out, void* stream) {
    if (!keys || !out) { set_error("null pointer"); return F2P_EINVAL; }
    if (n2 < 0) { set_error("bad n2=%d", n2); return F2P_EINVAL; }
    return launch_keys_to_map(keys, n2, flags, out, static_cast<cudaStream_t>(stream));
}

size_t f2p_filter_workspace_bytes(int M) { return 256 + filter_pair_bytes(M < 0 ? 0 : M); }

int f2p_spatial_filter(const float* cad, int ncad, const float* pc, int npc, const int64_t* p, int M,
                       const float thr[4], int64_t* out, int32_t* count, int32_t* status, uint8_t* mask, float* scores,
                       void* ws, size_t ws_bytes, void* stream) {
    if (!cad || !pc || !p || !out || !count || !status || !ws || !thr) { set_error("null pointer"); return F2P_EINVAL; }
    if (M < 0 || ncad < 1 || npc < 1) { set_error("bad shape M=%d ncad=%d npc=%d", M, ncad, npc); return F2P_EINVAL; }
    if (M > 524288) { set_error("M > 524288 candidates unsupported"); return F2P_ENOTSUP; }
    return run_filter_single(cad, ncad, pc, npc, p, M, thr, out, count, status, mask, scores,
                             static_cast<unsigned char*>(ws), ws_bytes, static_cast<cudaStream_t>(stream));
}

int f2p_test_vote_band(double band) { set_vote_band_override(band); return 0; }

int f2p_measure_tf32_peak(double* tflops, void* stream) { return measure_tf32_peak(tflops, static_cast<cudaStream_t>(stream)); }

int f2p_measure_tmem_ld(double* bytes_per_clk_sm, double* gbs, void* stream) {
    return measure_tmem_ld(bytes_per_clk_sm, gbs, static_cast<cudaStream_t>(stream));
}

int f2p_tc_last_stats(unsigned long long out[8]) {
    if (!out) { set_error("null pointer"); return F2P_EINVAL; }
    return tc_last_stats(out);
}

int f2p_measure_fp32_peak(double* packed, double* scalar, void* stream) {
    return measure_fp32_peak(packed, scalar, static_cast<cudaStream_t>(stream));
}

}  // extern "C"

// api_host.cu — host-buffer entry points: what a non-torch caller (C, C++, ctypes with numpy) binds.
//
// F2PPairStream is the per-sample loop of scripts/eval.py:83-91 as a small native executor: `depth` staging slots on the
// device, an upload stream, a download stream and one kernel stream per slot, and in-order completion into the caller's
// own buffers, so pair i+1 is uploaded and pair i-1 downloaded while pair i computes — and the kernels of small pairs in
// flight overlap on the GPU (one 5k x 2k pair keeps the SMs busy 70 % of its 50 us).  Device and pinned memory is grow-only and owned by
// the stream object: a steady stream of pairs allocates nothing per submission.  f2p_naive_host / f2p_spatial_host are
// one submission + flush on a per-thread stream object of depth 1.
#include <string.h>
#include <vector>
#include "common.cuh"
#include "internal.h"

using namespace f2p;

#define F2P_TRY(expr) do { int rc__ = (expr); if (rc__) return rc__; } while (0)

struct F2PPairStream {
    struct Slot {
        float* in = nullptr;        size_t in_cap = 0;       // device staging: C | evecs_x | evecs_y | cad | pc (floats)
        int64_t* out_d = nullptr;   size_t out_cap = 0;      // device result slots (+ 2 int32 words: count, status)
        int64_t* out_h = nullptr;   size_t outh_cap = 0;     // pinned mirror of out_d
        cudaStream_t comp = nullptr;                         // this slot's kernels: pairs in flight may overlap on the GPU
        void* ws = nullptr;         size_t ws_cap = 0;       // (a small pair leaves most SMs idle), so each has its own workspace
        cudaEvent_t up = nullptr, kdone = nullptr, done = nullptr;
        bool busy = false;
        int64_t* user_out = nullptr; int32_t* user_count = nullptr; int32_t* user_status = nullptr;
        size_t out_elems = 0;       // int64 elements the submission may produce (naive 2*n2, spatial 10*n2)
        bool spatial = false;
    };
    int device = 0, depth = 1;
    cudaStream_t h2d = nullptr, d2h = nullptr;
    std::vector<Slot> slots;
    long long submitted = 0, completed = 0;
};

namespace {

constexpr size_t ALIGN_F = 64;       // floats: every staged operand starts on a 256-byte boundary
inline size_t pad_f(size_t n) { return (n + ALIGN_F - 1) / ALIGN_F * ALIGN_F; }

int check_device(const F2PPairStream* s) {
    int dev = -1;
    F2P_CUDA(cudaGetDevice(&dev));
    if (dev != s->device) { set_error("pair stream belongs to device %d, current device is %d", s->device, dev); return F2P_EINVAL; }
    return 0;
}

// completes the submission held by `sl`: waits for its download and hands the result to the caller's buffers
int complete(F2PPairStream* s, F2PPairStream::Slot& sl) {
    if (!sl.busy) return 0;
    F2P_CUDA(cudaEventSynchronize(sl.done));
    if (sl.spatial) {
        const int32_t* meta = reinterpret_cast<const int32_t*>(sl.out_h + sl.out_elems);
        int32_t cnt = meta[0];
        if (cnt < 0 || (size_t)cnt * 2 > sl.out_elems) cnt = 0;
        *sl.user_count = cnt; *sl.user_status = meta[1];
        memcpy(sl.user_out, sl.out_h, sizeof(int64_t) * 2 * (size_t)cnt);       // the dense [2, M'] prefix
    } else if (sl.out_elems) {
        memcpy(sl.user_out, sl.out_h, sizeof(int64_t) * sl.out_elems);
    }
    sl.busy = false;
    ++s->completed;
    return 0;
}

int grow(F2PPairStream* s, F2PPairStream::Slot& sl, size_t in_floats, size_t out_elems, size_t ws_bytes) {
    if (in_floats > sl.in_cap) {
        if (sl.in) F2P_CUDA(cudaFree(sl.in));
        sl.in = nullptr; sl.in_cap = 0;
        F2P_CUDA(cudaMalloc(&sl.in, sizeof(float) * (in_floats + in_floats / 4)));
        sl.in_cap = in_floats + in_floats / 4;
    }
    const size_t need = out_elems + 2;                       // + (count, status) words
    if (need > sl.out_cap) {
        if (sl.out_d) F2P_CUDA(cudaFree(sl.out_d));
        if (sl.out_h) F2P_CUDA(cudaFreeHost(sl.out_h));
        sl.out_d = nullptr; sl.out_h = nullptr; sl.out_cap = 0;
        F2P_CUDA(cudaMalloc(&sl.out_d, sizeof(int64_t) * (need + need / 4)));
        F2P_CUDA(cudaMallocHost(&sl.out_h, sizeof(int64_t) * (need + need / 4)));
        sl.out_cap = need + need / 4;
    }
    if (ws_bytes > sl.ws_cap) {                              // the slot's previous submission is complete: nothing uses it
        if (sl.ws) F2P_CUDA(cudaFree(sl.ws));
        sl.ws = nullptr; sl.ws_cap = 0;
        F2P_CUDA(cudaMalloc(&sl.ws, ws_bytes + ws_bytes / 4));
        sl.ws_cap = ws_bytes + ws_bytes / 4;
    }
    (void)s;
    return 0;
}

int submit(F2PPairStream* s, bool spatial, const float* C, const float* ex, int n1, const float* ey, int n2, int k,
           const float* cad, int ncad, const float* pc, int npc, const float* thr, unsigned flags,
           int64_t* out, int32_t* count, int32_t* status, long long* completed_out) {
    if (!s) { set_error("null pair stream"); return F2P_EINVAL; }
    if (n1 < 0 || n2 < 0 || k < 1 || ncad < 0 || npc < 0) { set_error("negative size"); return F2P_EINVAL; }
    // an operand without elements may come as a null pointer (an empty array has no address)
    if (!C || (!ex && n1 > 0) || (n2 > 0 && (!ey || !out)) ||
        (spatial && ((!cad && ncad > 0) || (!pc && npc > 0) || !thr || !count || !status))) { set_error("null pointer"); return F2P_EINVAL; }
    F2P_TRY(check_device(s));
    F2PPairStream::Slot& sl = s->slots[(size_t)(s->submitted % s->depth)];
    F2P_TRY(complete(s, sl));
    const size_t out_elems = (spatial ? 10 : 2) * (size_t)n2;
    if (n2 == 0) {          // nothing to convert: an empty map (spatial: status 1, as f2p_spatial_batched reports it)
        F2P_TRY(grow(s, sl, 0, 0, 0));
        int32_t* meta = reinterpret_cast<int32_t*>(sl.out_h);
        meta[0] = 0; meta[1] = 1;
    } else {
        F2PPairDesc d; memset(&d, 0, sizeof(d));
        d.n1 = n1; d.n2 = n2; d.ldx = k; d.ldy = k; d.ncad = spatial ? ncad : n1; d.npc = spatial ? npc : n2;
        if (spatial) memcpy(d.thr, thr, sizeof(d.thr));
        const size_t wsb = f2p_workspace_bytes(&d, 1, k, spatial ? 1 : 0);
        if (wsb == 0) {
            if (k > F2P_MAX_K) { set_error("k = %d exceeds F2P_MAX_K = %d", k, F2P_MAX_K); return F2P_ENOTSUP; }
            if (n1 < 1) { set_error("pair 0: evecs_x has no rows (argmin of an empty dimension)"); return F2P_EINVAL; }
            return F2P_EINVAL;
        }
        const size_t oC = 0, ox = oC + pad_f((size_t)k * k), oy = ox + pad_f((size_t)n1 * k), ocad = oy + pad_f((size_t)n2 * k),
                     opc = ocad + (spatial ? pad_f(3 * (size_t)ncad) : 0), total = opc + (spatial ? pad_f(3 * (size_t)npc) : 0);
        F2P_TRY(grow(s, sl, total, out_elems, wsb));
        auto up = [&](size_t off, const float* src, size_t n) -> int {
            if (n) F2P_CUDA(cudaMemcpyAsync(sl.in + off, src, sizeof(float) * n, cudaMemcpyHostToDevice, s->h2d));
            return 0;
        };
        F2P_TRY(up(oC, C, (size_t)k * k)); F2P_TRY(up(ox, ex, (size_t)n1 * k)); F2P_TRY(up(oy, ey, (size_t)n2 * k));
        if (spatial) { F2P_TRY(up(ocad, cad, 3 * (size_t)ncad)); F2P_TRY(up(opc, pc, 3 * (size_t)npc)); }
        F2P_CUDA(cudaEventRecord(sl.up, s->h2d));
        F2P_CUDA(cudaStreamWaitEvent(sl.comp, sl.up, 0));
        int32_t* meta_d = reinterpret_cast<int32_t*>(sl.out_d + out_elems);
        if (spatial) {
            F2P_TRY(f2p_spatial_batched(sl.in + ox, sl.in + oy, sl.in + oC, sl.in + ocad, sl.in + opc, &d, 1, k, flags,
                                        sl.out_d, meta_d, meta_d + 1, nullptr, sl.ws, sl.ws_cap, sl.comp));
        } else {
            F2P_TRY(f2p_naive_batched(sl.in + ox, sl.in + oy, sl.in + oC, &d, 1, k, flags, sl.out_d, sl.ws, sl.ws_cap, sl.comp));
        }
        F2P_CUDA(cudaEventRecord(sl.kdone, sl.comp));
        F2P_CUDA(cudaStreamWaitEvent(s->d2h, sl.kdone, 0));
        F2P_CUDA(cudaMemcpyAsync(sl.out_h, sl.out_d, sizeof(int64_t) * (out_elems + (spatial ? 1 : 0)), cudaMemcpyDeviceToHost, s->d2h));
    }
    F2P_CUDA(cudaEventRecord(sl.done, s->d2h));
    sl.spatial = spatial; sl.user_out = out; sl.user_count = count; sl.user_status = status; sl.out_elems = out_elems;
    sl.busy = true;
    ++s->submitted;
    if (completed_out) *completed_out = s->completed;
    return 0;
}

int flush(F2PPairStream* s) {
    for (long long j = s->completed; j < s->submitted; ++j) F2P_TRY(complete(s, s->slots[(size_t)(j % s->depth)]));
    return 0;
}

void release(F2PPairStream* s) {
    for (auto& sl : s->slots) {
        if (sl.in) cudaFree(sl.in);
        if (sl.out_d) cudaFree(sl.out_d);
        if (sl.out_h) cudaFreeHost(sl.out_h);
        if (sl.up) cudaEventDestroy(sl.up);
        if (sl.kdone) cudaEventDestroy(sl.kdone);
        if (sl.done) cudaEventDestroy(sl.done);
        if (sl.ws) cudaFree(sl.ws);
        if (sl.comp) cudaStreamDestroy(sl.comp);
    }
    if (s->h2d) cudaStreamDestroy(s->h2d);
    if (s->d2h) cudaStreamDestroy(s->d2h);
    delete s;
}

int create(int depth, F2PPairStream** out) {
    if (!out || depth < 1 || depth > 16) { set_error("pair stream depth must be 1..16"); return F2P_EINVAL; }
    *out = nullptr;
    F2PPairStream* s = new F2PPairStream();
    s->depth = depth; s->slots.resize((size_t)depth);
    auto fail = [&](cudaError_t e, const char* what) { set_error("%s failed: %s", what, cudaGetErrorString(e)); release(s); return (int)e; };
    cudaError_t e;
    if ((e = cudaGetDevice(&s->device)) != cudaSuccess) return fail(e, "cudaGetDevice");
    for (cudaStream_t* st : {&s->h2d, &s->d2h})
        if ((e = cudaStreamCreateWithFlags(st, cudaStreamNonBlocking)) != cudaSuccess) return fail(e, "cudaStreamCreateWithFlags");
    for (auto& sl : s->slots) {
        if ((e = cudaStreamCreateWithFlags(&sl.comp, cudaStreamNonBlocking)) != cudaSuccess) return fail(e, "cudaStreamCreateWithFlags");
        for (cudaEvent_t* ev : {&sl.up, &sl.kdone, &sl.done})
            if ((e = cudaEventCreateWithFlags(ev, cudaEventDisableTiming)) != cudaSuccess) return fail(e, "cudaEventCreateWithFlags");
    }
    *out = s;
    return 0;
}

// per-thread, per-device stream object of depth 1 behind the synchronous host entry points
struct ThreadStreams {
    F2PPairStream* by_device[64] = {};
    ~ThreadStreams() { for (auto* s : by_device) if (s) release(s); }
};
int thread_stream(F2PPairStream** out) {
    static thread_local ThreadStreams tls;
    int dev = -1;
    F2P_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) { set_error("device index %d out of range", dev); return F2P_ENOTSUP; }
    if (!tls.by_device[dev]) F2P_TRY(create(1, &tls.by_device[dev]));
    *out = tls.by_device[dev];
    return 0;
}

}  // namespace

extern "C" {

int f2p_pair_stream_create(int depth, F2PPairStream** stream_out) { return create(depth, stream_out); }

int f2p_pair_stream_destroy(F2PPairStream* s) {
    if (!s) return 0;
    const int rc = flush(s);
    release(s);
    return rc;
}

int f2p_pair_stream_submit_naive(F2PPairStream* s, const float* C_host, const float* evecs_x_host, int n1,
                                 const float* evecs_y_host, int n2, int k, unsigned flags, int64_t* out_host,
                                 long long* completed_out) {
    return submit(s, false, C_host, evecs_x_host, n1, evecs_y_host, n2, k, nullptr, 0, nullptr, 0, nullptr, flags, out_host,
                  nullptr, nullptr, completed_out);
}

int f2p_pair_stream_submit_spatial(F2PPairStream* s, const float* C_host, const float* evecs_x_host, int n1,
                                   const float* evecs_y_host, int n2, int k, const float* cad_host, int ncad,
                                   const float* pc_host, int npc, const float thr[4], unsigned flags, int64_t* out_host,
                                   int32_t* count_host, int32_t* status_host, long long* completed_out) {
    return submit(s, true, C_host, evecs_x_host, n1, evecs_y_host, n2, k, cad_host, ncad, pc_host, npc, thr, flags, out_host,
                  count_host, status_host, completed_out);
}

int f2p_pair_stream_flush(F2PPairStream* s, long long* completed_out) {
    if (!s) { set_error("null pair stream"); return F2P_EINVAL; }
    F2P_TRY(check_device(s));
    F2P_TRY(flush(s));
    if (completed_out) *completed_out = s->completed;
    return 0;
}

int f2p_naive_host(const float* C, const float* ex, int n1, const float* ey, int n2, int k, unsigned flags,
                   int64_t* out) {
    F2PPairStream* s = nullptr;
    F2P_TRY(thread_stream(&s));
    F2P_TRY(submit(s, false, C, ex, n1, ey, n2, k, nullptr, 0, nullptr, 0, nullptr, flags, out, nullptr, nullptr, nullptr));
    return flush(s);
}

int f2p_spatial_host(const float* C, const float* ex, int n1, const float* ey, int n2, int k, const float* cad,
                     int ncad, const float* pc, int npc, const float thr[4], unsigned flags, int64_t* out,
                     int32_t* count, int32_t* status) {
    if (count) *count = 0;
    if (status) *status = 0;
    F2PPairStream* s = nullptr;
    F2P_TRY(thread_stream(&s));
    F2P_TRY(submit(s, true, C, ex, n1, ey, n2, k, cad, ncad, pc, npc, thr, flags, out, count, status, nullptr));
    return flush(s);
}

}  // extern "C"

// api_host.cu — host-buffer convenience entry points: what a non-torch caller (C, C++, ctypes with numpy) binds.
// Copies host->device, runs the batched device path with one pair, copies the result back; synchronous.
#include <string.h>
#include "common.cuh"
#include "internal.h"

using namespace f2p;

namespace {
struct DevBuf {
    void* p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    int alloc(size_t bytes) { return (int)cudaMalloc(&p, bytes > 0 ? bytes : 1); }
};
}  // namespace

#define F2P_TRY(expr) do { int rc__ = (expr); if (rc__) { if (rc__ > 0) set_error("CUDA error %d (%s) in %s", rc__, cudaGetErrorString((cudaError_t)rc__), #expr); return rc__; } } while (0)

extern "C" {

int f2p_naive_host(const float* C, const float* ex, int n1, const float* ey, int n2, int k, unsigned flags,
                   int64_t* out) {
    if (!C || !ex || !ey || !out) { set_error("null pointer"); return F2P_EINVAL; }
    if (n2 == 0) return 0;
    F2PPairDesc d; memset(&d, 0, sizeof(d));
    d.n1 = n1; d.n2 = n2; d.ldx = k; d.ldy = k; d.ncad = n1; d.npc = n2;
    const size_t wsb = f2p_workspace_bytes(&d, 1, k, 0);
    if (wsb == 0) return F2P_EINVAL;
    DevBuf dC, dx, dy, dout, ws;
    F2P_TRY(dC.alloc(sizeof(float) * (size_t)k * k)); F2P_TRY(dx.alloc(sizeof(float) * (size_t)n1 * k));
    F2P_TRY(dy.alloc(sizeof(float) * (size_t)n2 * k)); F2P_TRY(dout.alloc(sizeof(int64_t) * 2 * (size_t)n2));
    F2P_TRY(ws.alloc(wsb));
    F2P_CUDA(cudaMemcpy(dC.p, C, sizeof(float) * (size_t)k * k, cudaMemcpyHostToDevice));
    F2P_CUDA(cudaMemcpy(dx.p, ex, sizeof(float) * (size_t)n1 * k, cudaMemcpyHostToDevice));
    F2P_CUDA(cudaMemcpy(dy.p, ey, sizeof(float) * (size_t)n2 * k, cudaMemcpyHostToDevice));
    int rc = f2p_naive_batched((float*)dx.p, (float*)dy.p, (float*)dC.p, &d, 1, k, flags, (int64_t*)dout.p, ws.p, wsb, nullptr);
    if (rc) return rc;
    F2P_CUDA(cudaMemcpy(out, dout.p, sizeof(int64_t) * 2 * (size_t)n2, cudaMemcpyDeviceToHost));
    return 0;
}

int f2p_spatial_host(const float* C, const float* ex, int n1, const float* ey, int n2, int k, const float* cad,
                     int ncad, const float* pc, int npc, const float thr[4], unsigned flags, int64_t* out,
                     int32_t* count, int32_t* status) {
    if (!C || !ex || !ey || !cad || !pc || !thr || !out || !count || !status) { set_error("null pointer"); return F2P_EINVAL; }
    *count = 0; *status = 0;
    if (n2 == 0) { *status = 1; return 0; }
    F2PPairDesc d; memset(&d, 0, sizeof(d));
    d.n1 = n1; d.n2 = n2; d.ldx = k; d.ldy = k; d.ncad = ncad; d.npc = npc;
    memcpy(d.thr, thr, sizeof(d.thr));
    const size_t wsb = f2p_workspace_bytes(&d, 1, k, 1);
    if (wsb == 0) return F2P_EINVAL;
    DevBuf dC, dx, dy, dcad, dpc, dout, dcnt, ws;
    F2P_TRY(dC.alloc(sizeof(float) * (size_t)k * k)); F2P_TRY(dx.alloc(sizeof(float) * (size_t)n1 * k));
    F2P_TRY(dy.alloc(sizeof(float) * (size_t)n2 * k)); F2P_TRY(dcad.alloc(sizeof(float) * 3 * (size_t)ncad));
    F2P_TRY(dpc.alloc(sizeof(float) * 3 * (size_t)npc)); F2P_TRY(dout.alloc(sizeof(int64_t) * 10 * (size_t)n2));
    F2P_TRY(dcnt.alloc(2 * sizeof(int32_t))); F2P_TRY(ws.alloc(wsb));
    F2P_CUDA(cudaMemcpy(dC.p, C, sizeof(float) * (size_t)k * k, cudaMemcpyHostToDevice));
    F2P_CUDA(cudaMemcpy(dx.p, ex, sizeof(float) * (size_t)n1 * k, cudaMemcpyHostToDevice));
    F2P_CUDA(cudaMemcpy(dy.p, ey, sizeof(float) * (size_t)n2 * k, cudaMemcpyHostToDevice));
    F2P_CUDA(cudaMemcpy(dcad.p, cad, sizeof(float) * 3 * (size_t)ncad, cudaMemcpyHostToDevice));
    F2P_CUDA(cudaMemcpy(dpc.p, pc, sizeof(float) * 3 * (size_t)npc, cudaMemcpyHostToDevice));
    int32_t* dc = (int32_t*)dcnt.p;
    int rc = f2p_spatial_batched((float*)dx.p, (float*)dy.p, (float*)dC.p, (float*)dcad.p, (float*)dpc.p, &d, 1, k, flags,
                                 (int64_t*)dout.p, dc, dc + 1, nullptr, ws.p, wsb, nullptr);
    if (rc) return rc;
    int32_t h[2];
    F2P_CUDA(cudaMemcpy(h, dc, sizeof(h), cudaMemcpyDeviceToHost));
    *count = h[0]; *status = h[1];
    F2P_CUDA(cudaMemcpy(out, dout.p, sizeof(int64_t) * 10 * (size_t)n2, cudaMemcpyDeviceToHost));
    return 0;
}

}  // extern "C"

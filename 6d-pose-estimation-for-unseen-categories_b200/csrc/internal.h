// internal.h — launcher prototypes shared between the translation units of libf2p_sm100.so.
#pragma once
#include "common.cuh"

namespace f2p {

// Row splits per column tile.  Two goals: (1) at least ~2 waves of CTAs (2 CTAs/SM) even for one small pair, with
// >= 2 row tiles per item; (2) for big problems ~40 waves so the last partial wave costs < 3%, while every item still
// streams >= 16 row tiles (amortises the resident column-operand load and the final cross-thread reduction).
static inline int choose_nsplit(long long base_items, int nrt, int sm) {
    if (base_items <= 0 || nrt <= 1) return 1;
    const long long slots = 2ll * sm;
    long long ns1 = (2 * slots + base_items - 1) / base_items;
    long long ns2 = (40 * slots + base_items - 1) / base_items;
    const long long cap1 = nrt / 2 > 0 ? nrt / 2 : 1, cap2 = nrt / 16 > 0 ? nrt / 16 : 1;
    if (ns1 > cap1) ns1 = cap1;
    if (ns2 > cap2) ns2 = cap2;
    long long ns = ns1 > ns2 ? ns1 : ns2;
    return (int)(ns < 1 ? 1 : ns);
}

// prep.cu
int launch_project(bool is_x, const float* evecs, const float* Cmat, const PairDev* pairs_dev, int npairs,
                   int uniform_tiles, int total_tiles, int k, float* outT, int64_t ld_total, float* nrm,
                   cudaStream_t stream);
// nn.cu
struct ArgminWs {
    unsigned long long* best; unsigned long long* best_exact; float* yTf; float* nyf;
    uint32_t* best2; int* flagged; int* nflag;
};
size_t nn_smem_bytes(int k, int mode);
size_t top5_ws_bytes(int64_t total_items, int64_t total_ytiles);
size_t argmin_ws_bytes(int64_t ld2_total, int k, int npairs);
void argmin_ws_carve(ArgminWs& w, unsigned char* base, int64_t ld2_total, int k, int npairs);
int launch_argmin(const float* xT, int64_t ld1_total, const float* nx, const float* yT, int64_t ld2_total,
                  const float* ny, int k, const PairDev* pairs_dev, int npairs, int uniform_items, int total_items,
                  int uniform_ytiles, int total_ytiles, double flops, unsigned flags, int64_t* out, const ArgminWs& w,
                  cudaStream_t stream);
int launch_top5(const float* xT, int64_t ld1_total, const float* nx, const float* yT, int64_t ld2_total, const float* ny,
                int k, const PairDev* pairs_dev, int npairs, int uniform_items, int total_items, int uniform_ytiles,
                int total_ytiles, unsigned flags, unsigned long long* part5, int64_t* cand, int cand_is_out_layout,
                cudaStream_t stream);
// filter.cu
size_t filter_pair_bytes(int M);
size_t filter_table_bytes(int npairs);
int run_filter_batch(const PairDev* pairs_host, int npairs, const float* cad, const float* pc, const int64_t* cand,
                     int cand_is_out_layout, int64_t* out, int32_t* counts, int32_t* status,
                     unsigned char* table_dev, unsigned char* ws_base, const size_t* filter_off, cudaStream_t st);
int run_filter_single(const float* cad, int ncad, const float* pc, int npc, const int64_t* p, int M, const float thr[4],
                      int64_t* out, int32_t* count, int32_t* status, uint8_t* mask, float* scores,
                      unsigned char* ws, size_t ws_bytes, cudaStream_t st);
// peak.cu
int measure_fp32_peak(double* packed, double* scalar, cudaStream_t st);

}  // namespace f2p

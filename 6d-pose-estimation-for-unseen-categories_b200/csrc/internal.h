// internal.h — launcher prototypes shared between the translation units of libf2p_sm100.so.
#pragma once
#include "common.cuh"

namespace f2p {

// prep.cu
int launch_project(bool is_x, const float* evecs, const float* Cmat, const PairDev* pairs_dev, int npairs,
                   int uniform_tiles, int total_tiles, int k, float* outT, int64_t ld_total, float* nrm,
                   cudaStream_t stream);
// nn.cu
size_t nn_smem_bytes(int k, int mode);
int launch_nn(int mode, const float* xT, int64_t ld1_total, const float* nx, const float* yT, int64_t ld2_total,
              const float* ny, int k, const PairDev* pairs_dev, int npairs, int uniform_items, int total_items,
              unsigned long long* best, unsigned long long* part5, cudaStream_t stream);
int launch_finalize_argmin(const unsigned long long* best, const PairDev* pairs_dev, int npairs, int uniform_tiles,
                           int total_ytiles, unsigned flags, int64_t* out, cudaStream_t stream);
int launch_finalize_top5(const unsigned long long* part5, const PairDev* pairs_dev, int npairs, int uniform_tiles,
                         int total_ytiles, unsigned flags, int64_t* cand, int cand_is_out_layout, cudaStream_t stream);
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

// internal.h — launcher prototypes shared between the translation units of libf2p_sm100.so.
#pragma once
#include <stdlib.h>
#include "common.cuh"

namespace f2p {

// Row splits per column tile.  Big problems: ~40 waves of 2 CTAs/SM so the last partial wave costs < 3 %, while every
// item still streams >= 16 row tiles (amortises the resident column-operand load and the per-CTA bootstrap).
// Small problems (a few waves at most, e.g. one 5k x 2k pair = 16 column tiles): the split count is chosen by a wave
// model — time ~ waves x (fixed per-CTA cost + row tiles per item), fixed ~ 1.5 tile times (descriptor fetch, barrier
// init, column-operand load, first-tile bootstrap) — so that the grid fills whole waves: 18 splits = 288 CTAs = one
// wave on 148 SMs instead of 20 splits = 320 CTAs = one wave plus a 24-CTA tail.
// Row splits of the FP32 kernels when the tensor-pipe screen decides the call and they only run for the column tiles it
// flags (usually none): every CTA of the launch still has to start and read its tile's flag, so the grid is kept at about
// two waves instead of forty (cfg4: 782 instead of 11 730 CTAs, 47 -> 4 us per call); a flagged tile then streams more
// row tiles per CTA.
static inline int fallback_nsplit(int ns, long long base_items, int sm) {
    if (base_items <= 0) return ns;
    long long cap = (4ll * sm + base_items - 1) / base_items;
    if (cap < 1) cap = 1;
    return ns < cap ? ns : (int)cap;
}
static inline int choose_nsplit(long long base_items, int nrt, int sm, int min_tiles = 16) {
    if (base_items <= 0 || nrt <= 1) return 1;
    {   // F2P_NSPLIT: fixed split count for experiments (tools/)
        static const int forced = [] { const char* e = getenv("F2P_NSPLIT"); return e ? atoi(e) : 0; }();
        if (forced > 0) return forced < nrt ? forced : nrt;
    }
    const long long slots = 2ll * sm;
    const long long cap1 = nrt / 2 > 0 ? nrt / 2 : 1, cap2 = nrt / min_tiles > 0 ? nrt / min_tiles : 1;
    long long ns1 = (2 * slots + base_items - 1) / base_items;
    long long ns2 = (40 * slots + base_items - 1) / base_items;
    if (ns1 > cap1) ns1 = cap1;
    if (ns2 > cap2) ns2 = cap2;
    long long ns_big = ns1 > ns2 ? ns1 : ns2;
    if (ns_big < 1) ns_big = 1;
    // few waves (small problems, or mid-size batches whose items may not be split further): whole-wave cost model;
    // the top-5 kernel (min_tiles > 16) pays ~4 tile times per item for its bootstrap cuts, the arg-min kernel ~1.5
    if (base_items * cap1 <= 4 * slots || base_items * ns_big <= 8 * slots) {
        const double fixed = min_tiles > 16 ? 4.0 : 1.5;
        long long best = 1;
        double best_cost = 1e300;
        for (long long ns = 1; ns <= cap1; ++ns) {
            const long long waves = (base_items * ns + slots - 1) / slots;
            const double cost = (double)waves * (fixed + (double)((nrt + ns - 1) / ns));
            if (cost < best_cost - 1e-9) { best_cost = cost; best = ns; }
        }
        return (int)best;
    }
    return (int)ns_big;
}

struct TCPlan;
// prep.cu
int launch_project(bool is_x, const float* evecs, const float* Cmat, const PairDev* pairs_dev, int npairs,
                   int uniform_tiles, int total_tiles, int k, float* outT, int64_t ld_total, float* nrm,
                   cudaStream_t stream);
// both operands in one launch; best_init / tilecnt (optional): arg-min merge keys and per-column-tile arrival counters,
// reset by the CTAs that prepare the scene operand (then launch_argmin(..., pre_init = true) needs no init launch)
int launch_prep(const float* ex, const float* Cmat, const float* ey, const PairDev* pairs_dev, int npairs, int uni_xtiles,
                int total_xtiles, int uni_ytiles, int total_ytiles, int k, float* xT, int64_t ld1_total, float* nx,
                float* yT, int64_t ld2_total, float* ny, unsigned long long* best_init, int* tilecnt, cudaStream_t stream);
// nn.cu
struct ArgminWs {
    unsigned long long* best; unsigned long long* best_exact; float* yTf; float* nyf;
    uint32_t* best2; int* flagged; int* nflag; int* tilecnt;
};
size_t nn_smem_bytes(int k, int mode);
size_t top5_ws_bytes(int64_t total_items, int64_t total_ytiles);
size_t argmin_ws_bytes(int64_t ld2_total, int k, int npairs);
void argmin_ws_carve(ArgminWs& w, unsigned char* base, int64_t ld2_total, int k, int npairs);
int launch_argmin(const float* xT, int64_t ld1_total, const float* nx, const float* yT, int64_t ld2_total,
                  const float* ny, int k, const PairDev* pairs_dev, int npairs, int uniform_items, int total_items,
                  int uniform_ytiles, int total_ytiles, double flops, unsigned flags, int64_t* out, const ArgminWs& w,
                  cudaStream_t stream, bool pre_init = false, const TCPlan* tc = nullptr);
int launch_argmin_keys(const float* xT, int64_t ld1, const float* nx, const float* yT, int64_t ld2, const float* ny, int k,
                       const PairDev* pair_dev, int items, int ytiles, int n2, uint32_t row0, unsigned long long* keys,
                       const ArgminWs& w, cudaStream_t stream, const TCPlan* tc);
int launch_keys_to_map(const unsigned long long* keys, int n2, unsigned flags, int64_t* out, cudaStream_t stream);
int launch_top5(const float* xT, int64_t ld1_total, const float* nx, const float* yT, int64_t ld2_total, const float* ny,
                int k, const PairDev* pairs_dev, int npairs, int uniform_items, int total_items, int uniform_ytiles,
                int total_ytiles, unsigned flags, unsigned long long* part5, int64_t* cand, int cand_is_out_layout,
                cudaStream_t stream, const TCPlan* tc = nullptr);
// tc.cu — tensor-pipe (tcgen05, TF32) screen + exact verification of kernel 2
constexpr int TC_ROWS = 256;   // CAD rows per accumulator tile of the screen
constexpr int TC_EH = 2;       // epilogue threads per scene point (4 * TC_EH epilogue warps)
constexpr int TC_CG = 2;       // CTAs per MMA (cta_group): 2 = CTA pairs, each item covers two adjacent column tiles
struct TCWs {
    float* XA; float* XR; float* YA; uint32_t* tau; unsigned long long* cand; int* ccnt;
    unsigned char* zero_base; size_t zero_bytes;      // stats | nxmax | tcflag | pair_bad, cleared per call
    unsigned long long* stats; uint32_t* nxmax; int* tcflag; int* pair_bad;
};
// what the batch plan hands to launch_argmin / launch_top5 when the tensor-pipe screen is used (null = FFMA2 only)
struct TCPlan {
    int uni_items, total_items;       // screen work items (column tile x row split)
    int uni_xtiles, total_xtiles;     // 128-row packing tiles of the XA operand
    int64_t ld1_tc_total;             // rows of XA (every pair padded to a multiple of TC_ROWS)
    TCWs ws;
};
int tc_kp(int k);
size_t tc_ws_bytes(int64_t ld1_tc_total, int64_t ld2_total, int k, int64_t tc_items, int64_t total_ytiles, int npairs, int top5);
void tc_ws_carve(TCWs& w, unsigned char* base, int64_t ld1_tc_total, int64_t ld2_total, int k, int64_t tc_items,
                 int64_t total_ytiles, int npairs, int top5);
int launch_tc(int top5, const float* xT, int64_t ld1_total, const float* nx, const float* yT, int64_t ld2_total,
              const float* ny, int k, const PairDev* pairs_dev, int npairs, int uniform_tc_items, int total_tc_items,
              int uniform_tc_xtiles, int total_tc_xtiles, int uniform_ytiles, int total_ytiles, int64_t ld1_tc_total,
              unsigned flags, unsigned long long* best, int64_t* out5, int cand_is_out_layout, const TCWs& w,
              cudaStream_t stream, const unsigned long long* seed_keys = nullptr);
// Row splits per column tile of the screen (one CTA per SM, each owning all of TMEM): whole waves on `sm` SMs, every
// item streaming >= 8 tiles of 256 rows when the problem allows (per-CTA fixed cost ~ 2.5 tile times: TMEM allocation,
// barrier init, the resident scene tile, pipeline fill and drain).
static inline int choose_tc_nsplit(long long base_items, int nrt, int sm) {   // base_items: column tiles (CTAs per row split)
    if (base_items <= 0 || nrt <= 1) return 1;
    long long cap = nrt / 8 > 0 ? nrt / 8 : 1;
    if (base_items * cap < sm) cap = nrt / 2 > 0 ? nrt / 2 : 1;      // tiny problems: fill the SMs first
    long long best = 1;
    double best_cost = 1e300;
    for (long long ns = 1; ns <= cap && ns <= 64; ++ns) {
        const long long waves = (base_items * ns + sm - 1) / sm;
        const double cost = (double)waves * (2.5 + (double)((nrt + ns - 1) / ns));
        if (cost < best_cost * (1.0 - 1e-9)) { best_cost = cost; best = ns; }
    }
    return (int)best;
}
// Tensor-pipe path selection (F2P_NN_PATH=tc forces it, =gated / exact / two_phase disable it): by default problems with
// at least TC_MIN_FLOPS of distance work whose k fills a 128-byte swizzle row.  Non-semantic: every path is exact.
constexpr double TC_MIN_FLOPS = 8.0e9;    // measured crossover (tools/sweep.py with F2P_NN_PATH forced both ways): batches win from 1.2e10, single pairs from 1.0e10, lose at 2.4e9
bool tc_enabled(double flops, int k, int top5);
bool tc_fits(int k, int top5);   // the resident scene tile + a >= 2-deep ring fit one SM's shared memory
int measure_tf32_peak(double* tflops, cudaStream_t st);
int tc_last_stats(unsigned long long out[8]);
int measure_tmem_ld(double* bytes_per_clk_sm, double* gbs, cudaStream_t st);

// filter.cu
size_t filter_pair_bytes(int M);
size_t filter_table_bytes(int npairs);
// Optional stage timing of the batched entry points (api.cu; f2p_stage_timing / f2p_stage_times): marks are CUDA events
// recorded on the caller's stream at the stage boundaries — 0 start, 1 projection done, 2 kernel 2 done, 3 filter init
// done, then (vote, combine) of rounds 0..2 = marks 4..9.  No-ops unless timing is enabled.
constexpr int STAGE_MARKS = 10;
bool stage_timing_on();
void stage_mark(int mark, cudaStream_t st);
void stage_filter_counts(int* const* cnt_dev, int npairs);   // where each pair's round counts live (cnt[0..2])

int run_filter_batch(const PairDev* pairs_host, int npairs, const float* cad, const float* pc, const int64_t* cand,
                     int cand_is_out_layout, int64_t* out, int32_t* counts, int32_t* status,
                     unsigned char* table_dev, unsigned char* ws_base, const size_t* filter_off, cudaStream_t st);
int run_filter_single(const float* cad, int ncad, const float* pc, int npc, const int64_t* p, int M, const float thr[4],
                      int64_t* out, int32_t* count, int32_t* status, uint8_t* mask, float* scores,
                      unsigned char* ws, size_t ws_bytes, cudaStream_t st);
void set_vote_band_override(double band);   // test hook (f2p_test_vote_band): < 0 restores the proven band
// peak.cu
int measure_fp32_peak(double* packed, double* scalar, cudaStream_t st);

// positives.cu: exclusive scan of per-row counts (single CTA) -> row offsets + total
int launch_rowscan(const int* rowcnt, int n, int64_t* rowoff, int64_t* total, cudaStream_t stream);

}  // namespace f2p

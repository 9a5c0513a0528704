// tc.cu — kernel 2 on the tensor pipe: tcgen05 (TF32, FP32 accumulate in TMEM) SCREEN of the all-pairs squared
// distances with a proven error band, followed by the exact FP32 re-evaluation of the few in-band (row, column)
// candidates.  The result is bit-identical to the FFMA2 kernels of nn.cu (and to the reference, naive.py:33-34 /
// spacial_filtering.py:32-35): the tensor cores only decide which elements can NOT be among a column's K nearest.
//
// 1. tc_pack_kernel rewrites the k-major FP32 operands (xT = -2*feat_x, yT = evecs_y, norms nx / ny) as K-major rows
//    rounded to TF32 (cvt.rna), augmented so that the accumulator is already a LOWER bound of the reference's value:
//        XA[i] = [ tf32(-2 x_i[0..k)) | a_hi a_mid a_lo | 1 1 1 | 0.. ]      a_i = RD(nx_i (1 - EPS))
//        YA[j] = [ tf32(y_j[0..k))    | 1 1 1 | b_hi b_mid b_lo | 0.. ]      b_j = RD(ny_j (1 - EPS)) - 2^-100
//    so D[j][i] = YA[j] . XA[i] = s_ij - E_ij with s_ij ~ -2 x.y + nx + ny and E_ij = EPS (nx_i + ny_j) + 2^-100.
// 2. nn_tc_screen_kernel: CTA = 128 scene points (UMMA M, resident in shared memory) x a range of 256-row CAD tiles
//    (UMMA N) streamed by TMA (SWIZZLE_128B) through an mbarrier ring; one thread issues tcgen05.mma.kind::tf32 into a
//    double-buffered 128 x 256 FP32 accumulator (all 512 TMEM columns); the epilogue warps read it back with tcgen05.ld
//    (lane = scene point, so the running threshold T_j is a per-thread scalar), reduce 32 values with FMNMX3 and take the
//    slow path only when the chunk minimum is below T_j.  T_j = K-th smallest UPPER bound seen (UB = LB + W); an element
//    is a candidate iff LB_ij < T_j.  Thresholds are shared between CTAs through a global ordered-uint array.
// 3. nn_tc_verify_kernel<K>: every recorded candidate with LB < final T_j is evaluated with the reference's arithmetic
//    (sequential-k fmaf, + nx, + ny, clamp, sqrt_rn) and merged by (distance, row) exactly like nn.cu does.
// Column tiles the screen cannot decide (candidate list overflow, operands outside the proven range) are flagged and
// redone by the FFMA2 kernels.
//
// Error band ("tensor-pipe screen"): with u = 2^-24, S = sum_c |2 x_c y_c| <= nx + ny (AM-GM on the FP32
// vectors; the computed norms are within (k+1)u of the real ones):
//   reference  c_ref  = fl(fl(acc + nx) + ny), acc = k sequential FMAs:  |c_ref - (P + nx + ny)| <= (2k + 4) u (nx + ny)
//   TF32 inputs (round to nearest, 11 significant bits each):            |P_tf32 - P| <= (2^-10 + 2^-22) S
//   tensor-core accumulation (assumption A, checked at run time by the verify kernel's violation counter): every
//   k-step (8 products + accumulator) is within 2^-19 of the sum of the magnitudes it adds -> (k/8 + 2) 2^-18 (nx + ny)
//   EPS(k) = 2^-10 + 2^-13 + (12 k + 128) u covers all three with > 15 % to spare for k <= 256.
// UB_ij = LB_ij + 2 E_ij must also clear the sqrt_rn rounding class of the K-th best (a lower-index row with the same
// rounded distance wins the tie): W = OMEGA (nxmax_tile + ny_j) + 2^-98, OMEGA = 2 EPS + 2^-19.
#include "common.cuh"
#include "internal.h"

namespace f2p {

constexpr int TC_N = 256;            // CAD rows per accumulator tile (UMMA N)
constexpr int TC_KB = 32;            // floats per 128-byte swizzle row (one k-block)
constexpr int TC_A_BLOCK = TILE * TC_KB;     // floats of one resident k-block of the scene tile (16 KB)
constexpr int TC_B_STAGE = TC_N * TC_KB;     // floats of one streamed stage (32 KB)
constexpr uint32_t ENC_INF = 0xff800000u;    // ordered encoding of +inf

__device__ __forceinline__ uint32_t enc_f(float f) {
    const uint32_t b = __float_as_uint(f);
    return b ^ ((b >> 31) ? 0xffffffffu : 0x80000000u);
}
__device__ __forceinline__ float dec_f(uint32_t e) {
    return __uint_as_float(e ^ ((e >> 31) ? 0x80000000u : 0xffffffffu));
}
__device__ __forceinline__ float min3f(float a, float b, float c) {
    float r; asm("min.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r;
}
__device__ __forceinline__ float tf32_rna(float a) {
    uint32_t r; asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(a)); return __uint_as_float(r);
}

// ---- tcgen05 wrappers --------------------------------------------------------------------------------------------
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_alloc512(uint32_t* slot) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_free512(uint32_t taddr) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(512u) : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc], TF32 inputs, FP32 accumulate; issued by ONE thread
__device__ __forceinline__ void umma_tf32(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// arrives on `bar` when all tcgen05.mma issued so far by this thread have completed (implies fence::before_thread_sync)
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&v)[32]) {
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,"
        "%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr) : "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}
// ---- CTA-pair (cta_group::2) variants: one MMA spans two SMs, each CTA holds its own 128 scene points (M half) and
// half of the streamed CAD rows (N half); the leader CTA (cluster rank 0) issues, both CTAs feed and drain -----------
__device__ __forceinline__ uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_alloc512_2sm(uint32_t* slot) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_free512_2sm(uint32_t taddr) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(512u) : "memory");
}
__device__ __forceinline__ void umma_tf32_2sm(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// arrives on the barrier at this shared-memory offset in BOTH CTAs of the pair
__device__ __forceinline__ void umma_commit_2sm(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"((uint16_t)3) : "memory");
}
// TMA load into this CTA's shared memory whose bytes complete on the LEADER CTA's barrier (peer bit of the address cleared)
__device__ __forceinline__ void tma_load_2d_2sm(void* smem_dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(map), "r"(smem_u32(bar) & 0xfeffffffu), "r"(c0), "r"(c1) : "memory");
}
// arrive on the barrier at the same offset in cluster rank `cta` (default .release.cta semantics: the explicit
// .release.cluster form compiles to MEMBAR.ALL.GPU + ERRBAR, which cost the epilogue warps a third of their time — ncu;
// the hand-over of a TMEM buffer is ordered by tcgen05.fence::before_thread_sync, not by a memory fence)
__device__ __forceinline__ void mbar_arrive_remote(uint64_t* bar, uint32_t cta) {
    asm volatile(
        "{\n\t.reg .b32 ra;\n\t"
        "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
        "mbarrier.arrive.shared::cluster.b64 _, [ra];\n\t}" ::"r"(smem_u32(bar)), "r"(cta) : "memory");
}

// true in exactly one lane of the (fully active) warp; keeps the surrounding control flow warp-uniform so that the
// descriptors of tcgen05.mma / TMA stay in uniform registers (a `lane == 0` branch makes the compiler wrap every such
// instruction in a uniformisation loop: the single issuing thread then needs ~1000 cycles per 4-MMA stage — ncu, v1)
__device__ __forceinline__ bool elect_one() {
    uint32_t pred = 0;
    asm volatile(
        "{\n\t.reg .b32 rx;\n\t.reg .pred px;\n\t"
        "elect.sync rx|px, %1;\n\t"
        "@px mov.s32 %0, 1;\n\t}" : "+r"(pred) : "r"(0xffffffffu));
    return pred != 0;
}

__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major, SWIZZLE_128B shared-memory matrix descriptor (cute::UMMA::SmemDescriptor): start address >> 4, leading byte
// offset 1 (unused for swizzled K-major), stride byte offset 1024 B (one 8-row swizzle atom) >> 4, version 1 (sm_100),
// layout type 2 (SWIZZLE_128B).  A k-step of 8 TF32 inside the 128-byte row advances the start address by 32 B.
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t saddr) {
    return (uint64_t)((saddr & 0x3ffffu) >> 4) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
// instruction descriptor (cute::UMMA::InstrDescriptor): D = F32 (bits 4-5 = 1), A = B = TF32 (bits 7-9, 10-12 = 2), both
// K-major (bits 15, 16 = 0), N >> 3 at bits 17-22, M >> 4 at bits 24-28
constexpr uint32_t TC_IDESC = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(TC_N >> 3) << 17) | ((uint32_t)(TILE >> 4) << 24);

// bounded mbarrier wait: a protocol bug becomes a trap (launch failure) instead of a hung GPU
__device__ __forceinline__ bool mbar_try(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait_tc(uint64_t* bar, uint32_t parity) {
    if (mbar_try(bar, parity)) return;
    long long t0 = 0;
    for (unsigned n = 1; !mbar_try(bar, parity); ++n) {
        if ((n & 255u) == 0) {       // try_wait suspends the thread in hardware; the clock is only a watchdog
            if (t0 == 0) t0 = clock64();
            else if (clock64() - t0 > 4000000000ll) { printf("f2p tc: mbarrier timeout (block %d thread %d)\n", blockIdx.x, threadIdx.x); __trap(); }
        }
    }
}

struct TCParams {
    const PairDev* pairs;
    int npairs;
    int uniform_items;      // >0: every pair has this many TC work items
    int k, kp;              // kp = round_up(k + 6, 8): augmented K
    int nstage;
    const float* ny;        // [ld2_total]
    const float* nxmax;     // [total 256-row tiles] largest valid nx of the tile
    uint32_t* tau;          // [ld2_total] ordered-uint thresholds (K-th smallest upper bound seen by any CTA)
    u64* cand;              // [items][EH][CAP][TILE] (enc(LB) << 32 | row)
    int* ccnt;              // [items][EH][TILE]
    int* tcflag;            // [total column tiles] 1 = the FFMA2 kernels must redo this tile
    const int* pair_bad;    // [npairs] operands outside the proven range
    float omega;
    unsigned long long* stats;   // [0] recorded candidates, [1] overflowed lists, [2] verified candidates, [3] LB violations
};

template <int K> struct TCCfg { static constexpr int CAP = (K == 1) ? 16 : 32; };

// ---------------------------------------------------------------------------------------------------------------
// Screen kernel.  Warp 0 lane 0: TMA producer; warp 1: TMEM allocation, lane 0 issues the MMAs; warps 2..2+EW:
// epilogue (TMEM lane quarter = warp % 4; with EW = 8 two warps share a quarter and split the accumulator's columns).
// CG = 2: the two CTAs of a cluster form one cta_group::2 MMA (M = 256 = two scene tiles, N = 256): each CTA keeps its
// own scene tile resident and stages only HALF of every CAD tile, so L2 -> SM traffic and shared-memory reads per MMA are
// halved (the single-CTA form is L2-bandwidth bound: ncu, profiles/r02_ncu_tc_screen_large_v1_summary.txt).  The leader
// (cluster rank 0) owns the `full` and `tempty` barriers and issues every MMA; tcgen05.commit multicasts to the `empty`
// / `tfull` barriers of both CTAs.
// ---------------------------------------------------------------------------------------------------------------
template <int K, int EW, int CG>
__global__ void __launch_bounds__(64 + 32 * EW, 1)
nn_tc_screen_kernel(const __grid_constant__ CUtensorMap mapXA, const __grid_constant__ CUtensorMap mapYA, const TCParams prm) {
    constexpr int CAP = TCCfg<K>::CAP;
    constexpr int EH = EW / 4;                 // epilogue threads per scene point (column halves of the accumulator)
    constexpr int CHUNKS = TC_N / 32 / EH;     // 32-column chunks per epilogue thread and tile
    constexpr int NB = TC_N / CG;              // CAD rows this CTA stages per tile
    constexpr int B_STAGE = NB * TC_KB;        // floats per stage
    constexpr uint32_t IDESC = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(TC_N >> 3) << 17) | ((uint32_t)((TILE * CG) >> 4) << 24);
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    const int kp = prm.kp;
    const int nkb = (kp + TC_KB - 1) / TC_KB;
    const int last_steps = (kp - (nkb - 1) * TC_KB) / 8;
    const int S = prm.nstage;
    float* sA = reinterpret_cast<float*>(smem_raw);                     // [nkb][128][32]  (SW128)
    float* sB = sA + (size_t)nkb * TC_A_BLOCK;                          // [S][NB][32]     (SW128)
    u64* s_list = reinterpret_cast<u64*>(sB + (size_t)S * B_STAGE);     // [EH][CAP][TILE]
    uint64_t* bars = reinterpret_cast<uint64_t*>(s_list + (size_t)EH * CAP * TILE);
    uint64_t* fullA = bars;             // 1
    uint64_t* full = bars + 1;          // S (<= 8)
    uint64_t* empty = bars + 9;         // S
    uint64_t* tfull = bars + 17;        // 2
    uint64_t* tempty = bars + 19;       // 2
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 21);
    int* s_flag = reinterpret_cast<int*>(s_tmem + 1);
    uint32_t* s_T = reinterpret_cast<uint32_t*>(s_tmem + 2);       // [TILE] smallest threshold any epilogue thread of the column holds

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t rank = (CG == 2) ? cluster_ctarank() : 0u;
    const bool leader = rank == 0;
    const int item = blockIdx.x / CG;
    const int pi = find_pair(item, prm.pairs, prm.npairs, prm.uniform_items, &PairDev::tc_item0);
    const PairDev* pdp = prm.pairs + pi;
    const int nsplit = pdp->tc_nsplit, nrt = pdp->tc_nrt, n2 = pdp->n2, nct = pdp->nct;
    const int local = item - pdp->tc_item0;
    const int ctp = local / nsplit, sp = local - ctp * nsplit;
    const int ct_raw = ctp * CG + (int)rank;
    const bool dummy = ct_raw >= nct;          // odd tile count: the pair's last cluster has one idle half
    const int ct = dummy ? nct - 1 : ct_raw;   // (it still stages its half of the CAD rows and keeps the barriers going)
    const int t_begin = (int)(((long long)sp * nrt) / nsplit);
    const int t_end = (int)(((long long)(sp + 1) * nrt) / nsplit);
    const int ntiles = t_end - t_begin;
    const int xrow0 = pdp->xa_row0, ycol = pdp->y_col0 + ct * TILE;
    const int ytile = pdp->ytile0 + ct;
    const size_t slot = (size_t)item * CG + rank;      // candidate lists of this CTA

    // a pair outside the proven operand range is left to the FFMA2 kernels (same decision in both CTAs of a cluster:
    // the flag was written by the packing kernels)
    if (tid == 0) {
        const int bad = prm.pair_bad[pi];
        if (bad && !dummy) prm.tcflag[ytile] = 1;
        *s_flag = bad;
    }
    __syncthreads();
    if (*s_flag) return;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&mapXA);
        tma_prefetch_desc(&mapYA);
        mbar_init(fullA, 1);
        for (int s = 0; s < S; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(&tfull[b], 1); mbar_init(&tempty[b], EW * CG); }
        fence_barrier_init();
    }
    if (warp == 1) { if (CG == 2) tmem_alloc512_2sm(s_tmem); else tmem_alloc512(s_tmem); }
    if (tid >= 64 && tid < 64 + TILE) s_T[tid - 64] = ENC_INF;
    tc_fence_before();
    if (CG == 2) cluster_sync_all(); else __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *s_tmem;

    if (warp == 0) {
        // ---- TMA producer (both CTAs; the leader arms the barriers with the bytes of the whole pair).  The whole warp
        // walks the loop (uniform control flow), one elected lane issues. ----
        if (elect_one()) {
            if (leader) mbar_arrive_expect_tx(fullA, (uint32_t)(CG * nkb * TC_A_BLOCK * sizeof(float)));
            for (int kb = 0; kb < nkb; ++kb) {
                if (CG == 2) tma_load_2d_2sm(sA + (size_t)kb * TC_A_BLOCK, &mapYA, kb * TC_KB, ycol, fullA);
                else tma_load_2d(sA + (size_t)kb * TC_A_BLOCK, &mapYA, kb * TC_KB, ycol, fullA);
            }
        }
        __syncwarp();
        int s = 0; uint32_t ph = 1;          // waiting on parity 1 of a fresh barrier returns at once
        for (int t = 0; t < ntiles; ++t) {
            const int row = xrow0 + (t_begin + t) * TC_N + (int)rank * NB;
            for (int kb = 0; kb < nkb; ++kb) {
                mbar_wait_tc(&empty[s], ph);
                if (elect_one()) {
                    if (leader) mbar_arrive_expect_tx(&full[s], (uint32_t)(CG * B_STAGE * sizeof(float)));
                    if (CG == 2) tma_load_2d_2sm(sB + (size_t)s * B_STAGE, &mapXA, kb * TC_KB, row, &full[s]);
                    else tma_load_2d(sB + (size_t)s * B_STAGE, &mapXA, kb * TC_KB, row, &full[s]);
                }
                __syncwarp();
                if (++s == S) { s = 0; ph ^= 1u; }
            }
        }
    } else if (warp == 1) {
        if (leader) {
            // ---- MMA issuer: warp-uniform loop, descriptors advanced by adding to their address field ----
            mbar_wait_tc(fullA, 0);
            const uint64_t a_desc0 = smem_desc_sw128(smem_u32(sA)), b_desc0 = smem_desc_sw128(smem_u32(sB));
            constexpr uint32_t A_BLK16 = (uint32_t)(TC_A_BLOCK * sizeof(float)) >> 4, B_STG16 = (uint32_t)(B_STAGE * sizeof(float)) >> 4;
            int s = 0; uint32_t ph = 0;
            for (int t = 0; t < ntiles; ++t) {
                const int buf = t & 1;
                mbar_wait_tc(&tempty[buf], (uint32_t)(((t >> 1) & 1) ^ 1));
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)(buf * TC_N);
                for (int kb = 0; kb < nkb; ++kb) {
                    mbar_wait_tc(&full[s], ph);
                    tc_fence_after();
                    const uint64_t ad = a_desc0 + (uint64_t)((uint32_t)kb * A_BLK16);
                    const uint64_t bd = b_desc0 + (uint64_t)((uint32_t)s * B_STG16);
                    const int nst = (kb == nkb - 1) ? last_steps : TC_KB / 8;
                    if (elect_one()) {
#pragma unroll
                        for (int st = 0; st < TC_KB / 8; ++st) {
                            if (st < nst) {
                                if (CG == 2) umma_tf32_2sm(d_tmem, ad + 2 * st, bd + 2 * st, IDESC, (uint32_t)((kb | st) != 0));
                                else umma_tf32(d_tmem, ad + 2 * st, bd + 2 * st, IDESC, (uint32_t)((kb | st) != 0));
                            }
                        }
                        if (CG == 2) umma_commit_2sm(&empty[s]); else umma_commit(&empty[s]);   // frees the stage (both CTAs)
                        if (kb == nkb - 1) { if (CG == 2) umma_commit_2sm(&tfull[buf]); else umma_commit(&tfull[buf]); }   // tile t complete
                    }
                    __syncwarp();
                    if (++s == S) { s = 0; ph ^= 1u; }
                }
            }
        }
    } else {
        // ---- epilogue ----
        const int ew = warp - 2;
        const int quarter = warp & 3;              // TMEM lanes this warp may read: [32 * (warp % 4), +32)
        const int half = (EH == 2) ? (ew >> 2) : 0;
        const int col = quarter * 32 + lane;       // scene point of this thread within the tile
        const bool live = !dummy && ct * TILE + col < n2;
        u64* my_list = s_list + (size_t)half * CAP * TILE + col;     // entry e at my_list[e * TILE]
        const float nyj = live ? prm.ny[ycol + col] : 0.f;
        const float wy = __fmaf_ru(prm.omega, nyj, 3.1554436e-30f);   // OMEGA * ny + 2^-98
        uint32_t* taup = prm.tau + ycol + col;
        const float NEG_INF = __int_as_float(0xff800000), INF = __int_as_float(0x7f800000);
        float Tg = live ? dec_f(*reinterpret_cast<volatile uint32_t*>(taup)) : NEG_INF;
        float ub[K];                               // K smallest upper bounds seen by this thread, ascending
#pragma unroll
        for (int r = 0; r < K; ++r) ub[r] = INF;
        float T = Tg;                              // = min(ub[K-1], Tg)
        int cnt = 0;
        bool ovf = false;
        int nrec = 0;

        auto compact = [&]() {
            int j = 0;
            for (int e = 0; e < cnt; ++e) {
                const u64 v = my_list[e * TILE];
                if (dec_f((uint32_t)(v >> 32)) < T) { my_list[j * TILE] = v; ++j; }
            }
            cnt = j;
        };
        auto record = [&](float v, uint32_t row, float Wt) {
            if (cnt == CAP) compact();
            if (cnt < CAP) { my_list[cnt * TILE] = ((u64)enc_f(v) << 32) | row; ++cnt; ++nrec; }
            else ovf = true;
            float u = __fadd_ru(v, Wt);
            // most records of a settled column are in band but above its K-th upper bound: they cannot move the threshold
            if (u < ub[K - 1]) {
#pragma unroll
                for (int r = 0; r < K; ++r) { const float t2 = ub[r]; const bool p = u < t2; ub[r] = p ? u : t2; u = p ? t2 : u; }
                T = fminf(ub[K - 1], Tg);
            }
        };

        const uint32_t lane_addr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(half * CHUNKS * 32);
        // value e (0..7) of eight registers without a local-memory round trip: 7 selects
        auto sel8 = [](const float* c, int e) -> float {
            const bool b0 = e & 1, b1 = e & 2, b2 = e & 4;
            const float a0 = b0 ? c[1] : c[0], a1 = b0 ? c[3] : c[2], a2 = b0 ? c[5] : c[4], a3 = b0 ? c[7] : c[6];
            const float d0 = b1 ? a1 : a0, d1 = b1 ? a3 : a2;
            return b2 ? d1 : d0;
        };
        for (int t = 0; t < ntiles; ++t) {
            const int buf = t & 1;
            const int tile = t_begin + t;
            const float Wt = __fmaf_ru(prm.omega, __ldg(prm.nxmax + (xrow0 / TC_N) + tile), wy);
            // thresholds found by the other CTAs of this column tile (one L2 read per tile)
            if (live) { Tg = fminf(Tg, dec_f(*reinterpret_cast<volatile uint32_t*>(taup))); T = fminf(T, Tg); }
            const float T_in = T;
            mbar_wait_tc(&tfull[buf], (uint32_t)((t >> 1) & 1));
            tc_fence_after();
            const uint32_t taddr = lane_addr + (uint32_t)(buf * TC_N);
            const uint32_t rowbase = (uint32_t)(tile * TC_N + half * CHUNKS * 32);
            float va[32], vb[32];
            if (K > 1 && t == 0) {
                // Cold start of a top-K scan: instead of tightening the threshold one record at a time (K ln(rows / K)
                // records per thread in the first tile), take the K-th smallest of this thread's 4 * CHUNKS group minima:
                // K different groups hold a value <= it, so it bounds the K-th smallest lower bound, and + W the K-th
                // smallest UPPER bound — a valid threshold found without a single record.  The tile is then read again.
                float gm[K];
#pragma unroll
                for (int r = 0; r < K; ++r) gm[r] = INF;
                for (int c = 0; c < CHUNKS; ++c) {
                    __syncwarp();
                    tmem_ld32(taddr + (uint32_t)(c * 32), va);
                    tmem_ld_wait();
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const float m0 = min3f(va[8 * q], va[8 * q + 1], va[8 * q + 2]);
                        const float m1 = min3f(va[8 * q + 3], va[8 * q + 4], va[8 * q + 5]);
                        float u = min3f(m0, m1, fminf(va[8 * q + 6], va[8 * q + 7]));
#pragma unroll
                        for (int r = 0; r < K; ++r) { const float t2 = gm[r]; const bool p = u < t2; gm[r] = p ? u : t2; u = p ? t2 : u; }
                    }
                }
                if (live && 4 * CHUNKS >= K) { Tg = fminf(Tg, __fadd_ru(gm[K - 1], Wt)); T = fminf(T, Tg); }
            }
            __syncwarp();
            tmem_ld32(taddr, va);
#pragma unroll
            for (int c = 0; c < CHUNKS; ++c) {
                float (&cur)[32] = (c & 1) ? vb : va;
                float (&nxt)[32] = (c & 1) ? va : vb;
                __syncwarp();           // the slow path below diverges; tcgen05.ld / wait::ld are warp-collective
                tmem_ld_wait();
                if (c + 1 < CHUNKS) tmem_ld32(taddr + (uint32_t)((c + 1) * 32), nxt);
                else {
                    // all of this thread's reads of the buffer have landed: hand it back to the MMA issuer
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) { if (CG == 2) mbar_arrive_remote(&tempty[buf], 0); else mbar_arrive(&tempty[buf]); }
                }
                // the other epilogue thread(s) of this column may have found a tighter threshold
                if (EH > 1 && live) { Tg = fminf(Tg, dec_f(s_T[col])); T = fminf(T, Tg); }
                const float T_c = T;
                float g[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const float m0 = min3f(cur[8 * q], cur[8 * q + 1], cur[8 * q + 2]);
                    const float m1 = min3f(cur[8 * q + 3], cur[8 * q + 4], cur[8 * q + 5]);
                    g[q] = min3f(m0, m1, fminf(cur[8 * q + 6], cur[8 * q + 7]));
                }
                const float cm = fminf(fminf(g[0], g[1]), fminf(g[2], g[3]));
                if (cm < T) {
                    // rare: walk the passing values of every group that holds one (mask + register select: no local memory)
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        if (g[q] < T) {
                            uint32_t m8 = 0;
#pragma unroll
                            for (int e = 0; e < 8; ++e) m8 |= (cur[8 * q + e] < T ? 1u : 0u) << e;
                            while (m8) {
                                const int e = __ffs(m8) - 1;
                                m8 &= m8 - 1;
                                const float v = sel8(&cur[8 * q], e);
                                if (v < T) record(v, rowbase + (uint32_t)(c * 32 + 8 * q + e), Wt);
                            }
                        }
                    }
                    if (EH > 1 && T < T_c && live) atomicMin(&s_T[col], enc_f(T));
                }
            }
            if (T < T_in && live) { atomicMin(taup, enc_f(T)); }
        }
        // survivors of this CTA: entries still below its final threshold
        if (!dummy) {
            if (live) {
                compact();
                u64* gl = prm.cand + ((slot * EH + half) * CAP) * TILE + col;
                for (int e = 0; e < cnt; ++e) gl[(size_t)e * TILE] = my_list[e * TILE];
                if (ovf) prm.tcflag[ytile] = 1;
            }
            prm.ccnt[(slot * EH + half) * TILE + col] = live ? cnt : 0;
        }
        if (prm.stats) {
            int tot = nrec;
#pragma unroll
            for (int m = 16; m >= 1; m >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, m);
            const unsigned any_ovf = __ballot_sync(0xffffffffu, ovf);
            if (lane == 0) { atomicAdd(prm.stats + 0, (unsigned long long)tot); if (any_ovf) atomicAdd(prm.stats + 1, (unsigned long long)__popc(any_ovf)); }
        }
    }

    // nobody leaves while the other CTA may still signal this one's barriers or read its shared memory
    tc_fence_before();
    if (CG == 2) cluster_sync_all(); else __syncthreads();
    if (warp == 1) { tc_fence_after(); if (CG == 2) tmem_free512_2sm(tmem_base); else tmem_free512(tmem_base); }
}

// ---------------------------------------------------------------------------------------------------------------
// Exact re-evaluation of the candidates (reference arithmetic, identical to nn.cu) — one thread per scene point.
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ u64 tc_make_key(float d, uint32_t i) {
    return ((u64)(__float_as_uint(d) & 0x7fffffffu) << 32) | (u64)i;
}
__device__ __forceinline__ void tc_insert5(u64 (&l)[5], u64 key) {
#pragma unroll
    for (int r = 0; r < 5; ++r) {
        const bool p = key < l[r];
        const u64 t = l[r];
        l[r] = p ? key : t;
        key = p ? t : key;
    }
}

struct TCVerifyParams {
    const PairDev* pairs; int npairs; int uniform_ytiles;
    int k, eh, cap;
    const float* XR;        // [ld1_tc_total][kr] exact FP32 rows of -2*feat_x (row-major copy made by the packing kernel;
                            // kr = round_up(k, 4), zero padded, so every row is read as float4)
    const float* nx;
    const float* yT; int64_t ld2_total; const float* ny;
    const uint32_t* tau; const u64* cand; const int* ccnt; const int* tcflag;
    u64* best;              // K = 1: merge keys of the arg-min
    int64_t* out5;          // K = 5: nn_query layout
    int cand_is_out_layout; unsigned flags;
    unsigned long long* stats;
};

// Block = one column tile (128 scene points).  The candidates of a tile are spread very unevenly over its columns (about
// one per column on average, a dozen for a few), and one exact evaluation is a dependent chain of k FMAs fed by a cold
// 4k-byte row: a thread per column made the whole launch wait for its unluckiest thread (ncu: 223 us at cfg4, 49
// long-scoreboard stalls per issue).  So the tile's (column, row) items are flattened into one shared list and dealt out
// to all 128 threads; the scene columns are staged in shared memory once, the CAD rows are L2-prefetched and read as
// float4.  Arithmetic per item is unchanged (sequential-k fmaf, + nx, + ny, clamp, sqrt_rn).
constexpr int VCAP = 2048;      // items per pass (a tile with more takes several passes)

template <int K>
__global__ void __launch_bounds__(TILE)
nn_tc_verify_kernel(const TCVerifyParams prm) {
    extern __shared__ __align__(16) float s_y[];       // [k][TILE] scene columns of this tile
    __shared__ uint32_t s_row[VCAP];
    __shared__ float s_lb[VCAP];
    __shared__ u64 s_key[VCAP];
    __shared__ uint16_t s_col[VCAP];
    __shared__ u64 s_best[TILE];
    __shared__ float s_ny[TILE];
    __shared__ int s_wsum[TILE / 32];
    const int pi = find_pair(blockIdx.x, prm.pairs, prm.npairs, prm.uniform_ytiles, &PairDev::ytile0);
    const PairDev* pd = prm.pairs + pi;
    const int ct = blockIdx.x - pd->ytile0;
    if (prm.tcflag[blockIdx.x]) return;         // redone by the FFMA2 kernels (uniform exit)
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int j = ct * TILE + tid, n2 = pd->n2;
    const int64_t yc = (int64_t)pd->y_col0 + j;   // padded columns exist in yT / ny / tau / ccnt (all zero / empty)
    const int k = prm.k;
    const uint32_t tau = prm.tau[yc];
    s_ny[tid] = prm.ny[yc];
    s_best[tid] = ~0ull;
    {   // stage the tile's scene columns: k coalesced asynchronous copies per thread, all in flight at once (a register-
        // staged loop kept eight in flight and paid k/8 memory round trips per CTA)
        const float* yp = prm.yT + yc;
        for (int c = 0; c < k; ++c)
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(s_y + c * TILE + tid)),
                         "l"(yp + (int64_t)c * prm.ld2_total) : "memory");
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    const int nsplit = pd->tc_nsplit;
    const int nlists = nsplit * prm.eh;
    auto list_of = [&](int sh, const u64*& cl) -> int {
        // lists of the CTA (item, cluster rank) that screened this column tile for row split sh / eh, accumulator half sh % eh
        const size_t item = (size_t)pd->tc_item0 + (size_t)(ct / TC_CG) * nsplit + (size_t)(sh / prm.eh);
        const size_t slot = (item * TC_CG + (size_t)(ct % TC_CG)) * prm.eh + (size_t)(sh % prm.eh);
        cl = prm.cand + slot * prm.cap * TILE + tid;
        return prm.ccnt[slot * TILE + tid];
    };
    // Walks this column's lists, eight entries per step with all eight loads issued before any is used (a loop with one
    // dependent L2 round trip per entry was most of this kernel's time on short scans: ncu, 8 % issue utilisation).
    // Slots past the end read as ~0, whose encoded bound is above every threshold.
    auto walk = [&](auto&& visit) {
        // two lists at a time: both counts, then eight entries of each in flight together; every walk visits the entries
        // in the same order (the second walk relies on it)
        for (int sh = 0; sh < nlists; sh += 2) {
            const u64 *c0, *c1 = nullptr;
            const int n0 = list_of(sh, c0);
            const int n1 = (sh + 1 < nlists) ? list_of(sh + 1, c1) : 0;
            for (int e0 = 0; e0 < n0 || e0 < n1; e0 += 8) {
                u64 v0[8], v1[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    v0[u] = (e0 + u < n0) ? __ldg(c0 + (size_t)(e0 + u) * TILE) : ~0ull;
                    v1[u] = (e0 + u < n1) ? __ldg(c1 + (size_t)(e0 + u) * TILE) : ~0ull;
                }
#pragma unroll
                for (int u = 0; u < 8; ++u)
                    if ((uint32_t)(v0[u] >> 32) < tau) visit(v0[u]);
#pragma unroll
                for (int u = 0; u < 8; ++u)
                    if ((uint32_t)(v1[u] >> 32) < tau) visit(v1[u]);
            }
        }
    };
    if constexpr (K == 1) {
        // Arg-min: ONE pass.  A passing entry takes the next slot of the CTA's item arrays through a shared counter (the
        // order of the items does not matter: keys merge with atomicMin); lists are read two at a time, both counts and
        // then both lists' first eight entries in flight together, so a column's walk is a handful of L2 round trips
        // instead of three per list and pass.  Entries beyond the item arrays' capacity are evaluated by their own thread.
        __shared__ int s_n;
        if (tid == 0) s_n = 0;
        __syncthreads();
        const int kr = (k + 3) & ~3, k4 = k & ~3;
        int mine = 0, nviol = 0;
        // one candidate in the reference's arithmetic (sequential-k fmaf, + nx, + ny, clamp, sqrt_rn)
        auto eval_one = [&](uint32_t i0, int col, float lb) -> u64 {
            const float* xp = prm.XR + ((int64_t)pd->xa_row0 + i0) * kr;
            const float* yq = s_y + col;
            float acc = 0.f;
#pragma unroll 4
            for (int c = 0; c < k4; c += 4) {
                const float4 xv = __ldg(reinterpret_cast<const float4*>(xp + c));
                acc = fmaf(xv.x, yq[c * TILE], acc);
                acc = fmaf(xv.y, yq[(c + 1) * TILE], acc);
                acc = fmaf(xv.z, yq[(c + 2) * TILE], acc);
                acc = fmaf(xv.w, yq[(c + 3) * TILE], acc);
            }
            if (k4 < k) {
                const float4 xv = __ldg(reinterpret_cast<const float4*>(xp + k4));
                const float xa[3] = {xv.x, xv.y, xv.z};
#pragma unroll
                for (int t = 0; t < 3; ++t)
                    if (k4 + t < k) acc = fmaf(xa[t], yq[(k4 + t) * TILE], acc);
            }
            const float v = __fadd_rn(__fadd_rn(acc, __ldg(prm.nx + pd->x_col0 + i0)), s_ny[col]);
            if (lb > v) ++nviol;                           // the lower bound must hold: assumption A's run-time check
            return tc_make_key(__fsqrt_rn(fmaxf(v, 0.f)), i0);
        };
        auto take = [&](u64 ce) {
            ++mine;
            const int slot = atomicAdd(&s_n, 1);
            if (slot < VCAP) {
                s_row[slot] = (uint32_t)(ce & 0xffffffffull);
                s_lb[slot] = dec_f((uint32_t)(ce >> 32));
                s_col[slot] = (uint16_t)tid;
            } else {
                asm volatile("cp.async.wait_all;" ::: "memory");    // this column of the scene tile was copied by this thread
                atomicMin(&s_best[tid], eval_one((uint32_t)(ce & 0xffffffffull), tid, dec_f((uint32_t)(ce >> 32))));
            }
        };
        for (int sh = 0; sh < nlists; sh += 2) {
            const u64 *c0, *c1 = nullptr;
            const int n0 = list_of(sh, c0);
            const int n1 = (sh + 1 < nlists) ? list_of(sh + 1, c1) : 0;
            for (int e0 = 0; e0 < n0 || e0 < n1; e0 += 8) {
                u64 v0[8], v1[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    v0[u] = (e0 + u < n0) ? __ldg(c0 + (size_t)(e0 + u) * TILE) : ~0ull;
                    v1[u] = (e0 + u < n1) ? __ldg(c1 + (size_t)(e0 + u) * TILE) : ~0ull;
                }
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    if ((uint32_t)(v0[u] >> 32) < tau) take(v0[u]);
                    if ((uint32_t)(v1[u] >> 32) < tau) take(v1[u]);
                }
            }
        }
        asm volatile("cp.async.wait_all;" ::: "memory");     // the scene tile (its copies ran behind the walk)
        __syncthreads();
        const int cnt = min(VCAP, s_n);
        for (int w = tid; w < cnt; w += TILE) {
            const int wn = w + TILE;
            if (wn < cnt) {                                  // the next item's row towards L1 while this one computes
                const float* q0 = prm.XR + ((int64_t)pd->xa_row0 + s_row[wn]) * kr;
                for (int c = 0; c < kr; c += 32) prefetch_l1(q0 + c);
            }
            const int col = s_col[w];
            atomicMin(&s_best[col], eval_one(s_row[w], col, s_lb[w]));
        }
        __syncthreads();
        if (j < n2 && s_best[tid] != ~0ull) atomicMin(prm.best + yc, s_best[tid]);
        if (prm.stats) {
            int nv = mine, vi = nviol;
#pragma unroll
            for (int m = 16; m >= 1; m >>= 1) { nv += __shfl_xor_sync(0xffffffffu, nv, m); vi += __shfl_xor_sync(0xffffffffu, vi, m); }
            if (lane == 0) {
                if (nv) atomicAdd(prm.stats + 2, (unsigned long long)nv);
                if (vi) atomicAdd(prm.stats + 3, (unsigned long long)vi);
            }
        }
        return;
    }
    // pass 0: how many entries of this column are below its final threshold
    int mine = 0;
    walk([&](u64) { ++mine; });
    // exclusive scan over the 128 threads
    int incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += v; }
    if (lane == 31) s_wsum[warp] = incl;
    asm volatile("cp.async.wait_all;" ::: "memory");         // the scene tile (its copies ran behind the counting pass)
    __syncthreads();
    int offset = incl - mine, total = 0;
#pragma unroll
    for (int w = 0; w < TILE / 32; ++w) { const int v = s_wsum[w]; if (w < warp) offset += v; total += v; }

    u64 best[K];
#pragma unroll
    for (int r = 0; r < K; ++r) best[r] = ~0ull;
    int nviol = 0;
    for (int base = 0; base < total; base += VCAP) {
        // this thread's items that fall into [base, base + VCAP)
        if (offset < base + VCAP && offset + mine > base) {
            int idx = offset;
            walk([&](u64 ce) {
                if (idx >= base && idx < base + VCAP) {
                    s_row[idx - base] = (uint32_t)(ce & 0xffffffffull);
                    s_lb[idx - base] = dec_f((uint32_t)(ce >> 32));
                    s_col[idx - base] = (uint16_t)tid;
                }
                ++idx;
            });
        }
        __syncthreads();
        const int cnt = min(VCAP, total - base);
        const int kr = (k + 3) & ~3, k4 = k & ~3;
        // two items per thread and step: two independent k-long FMA chains in flight (each one is the reference's own
        // sequential chain; nothing is reassociated)
        for (int w = tid; w < cnt; w += 2 * TILE) {
            const int w2 = w + TILE;
            const bool two = w2 < cnt;
            const uint32_t i0 = s_row[w], i1 = two ? s_row[w2] : i0;
            const int col0 = s_col[w], col1 = two ? s_col[w2] : col0;
            const float* xp0 = prm.XR + ((int64_t)pd->xa_row0 + i0) * kr;
            const float* xp1 = prm.XR + ((int64_t)pd->xa_row0 + i1) * kr;
            {   // rows of the NEXT step's two items towards L2 / L1 while this step computes
                const int wn = w + 2 * TILE, wn2 = wn + TILE;
                if (wn < cnt) {
                    const float* q0 = prm.XR + ((int64_t)pd->xa_row0 + s_row[wn]) * kr;
                    for (int c = 0; c < kr; c += 32) prefetch_l1(q0 + c);
                    if (wn2 < cnt) {
                        const float* q1 = prm.XR + ((int64_t)pd->xa_row0 + s_row[wn2]) * kr;
                        for (int c = 0; c < kr; c += 32) prefetch_l1(q1 + c);
                    }
                }
            }
            const float* yq0 = s_y + col0;
            const float* yq1 = s_y + col1;
            float acc0 = 0.f, acc1 = 0.f;
#pragma unroll 4
            for (int c = 0; c < k4; c += 4) {
                const float4 xv = __ldg(reinterpret_cast<const float4*>(xp0 + c));
                const float4 xw = __ldg(reinterpret_cast<const float4*>(xp1 + c));
                acc0 = fmaf(xv.x, yq0[c * TILE], acc0);             acc1 = fmaf(xw.x, yq1[c * TILE], acc1);
                acc0 = fmaf(xv.y, yq0[(c + 1) * TILE], acc0);       acc1 = fmaf(xw.y, yq1[(c + 1) * TILE], acc1);
                acc0 = fmaf(xv.z, yq0[(c + 2) * TILE], acc0);       acc1 = fmaf(xw.z, yq1[(c + 2) * TILE], acc1);
                acc0 = fmaf(xv.w, yq0[(c + 3) * TILE], acc0);       acc1 = fmaf(xw.w, yq1[(c + 3) * TILE], acc1);
            }
            if (k4 < k) {
                const float4 xv = __ldg(reinterpret_cast<const float4*>(xp0 + k4));
                const float4 xw = __ldg(reinterpret_cast<const float4*>(xp1 + k4));
                const float xa[3] = {xv.x, xv.y, xv.z}, xb[3] = {xw.x, xw.y, xw.z};
#pragma unroll
                for (int t = 0; t < 3; ++t)
                    if (k4 + t < k) { acc0 = fmaf(xa[t], yq0[(k4 + t) * TILE], acc0); acc1 = fmaf(xb[t], yq1[(k4 + t) * TILE], acc1); }
            }
            const float v0 = __fadd_rn(__fadd_rn(acc0, __ldg(prm.nx + pd->x_col0 + i0)), s_ny[col0]);
            const float v1 = __fadd_rn(__fadd_rn(acc1, __ldg(prm.nx + pd->x_col0 + i1)), s_ny[col1]);
            if (s_lb[w] > v0) ++nviol;         // the lower bound must hold: assumption A's run-time check
            const u64 key0 = tc_make_key(__fsqrt_rn(fmaxf(v0, 0.f)), i0);
            if (K == 1) atomicMin(&s_best[col0], key0);
            else s_key[w] = key0;
            if (two) {
                if (s_lb[w2] > v1) ++nviol;
                const u64 key1 = tc_make_key(__fsqrt_rn(fmaxf(v1, 0.f)), i1);
                if (K == 1) atomicMin(&s_best[col1], key1);
                else s_key[w2] = key1;
            }
        }
        __syncthreads();
        if (K != 1) {
            const int lo = max(offset, base), hi = min(offset + mine, base + VCAP);
            for (int w = lo; w < hi; ++w) {
                const u64 key = s_key[w - base];
                if (key < best[K - 1]) tc_insert5(reinterpret_cast<u64(&)[5]>(best), key);
            }
            __syncthreads();
        }
    }
    if (j < n2) {
        if (K == 1) {
            if (s_best[tid] != ~0ull) atomicMin(prm.best + yc, s_best[tid]);
        } else {
            const int64_t M = 5ll * n2;
            int64_t* o = prm.out5 + (prm.cand_is_out_layout ? pd->out_off : 2ll * pd->cand0);
            const int64_t r1 = (prm.flags & F2P_INT16_COMPAT) ? (int64_t)(int16_t)(uint16_t)(j & 0xffff) : (int64_t)j;
#pragma unroll
            for (int r = 0; r < K; ++r) {
                o[5ll * j + r] = (int64_t)(uint32_t)(best[r] & 0xffffffffull);
                if (!(prm.flags & F2P_ROW0_ONLY)) o[M + 5ll * j + r] = r1;
            }
        }
    }
    if (prm.stats) {
        int nv = mine, vi = nviol;
#pragma unroll
        for (int m = 16; m >= 1; m >>= 1) { nv += __shfl_xor_sync(0xffffffffu, nv, m); vi += __shfl_xor_sync(0xffffffffu, vi, m); }
        if (lane == 0) {
            if (nv) atomicAdd(prm.stats + 2, (unsigned long long)nv);
            if (vi) atomicAdd(prm.stats + 3, (unsigned long long)vi);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Operand packing: k-major FP32 -> K-major TF32 rows with the augmentation columns.  CTA = 128 rows.
//   IS_X: rows of a pair's TC extent (multiples of 256; rows >= n1 get a = 1e30 so they are never candidates);
//         folds the largest valid nx of every 256-row tile into nxmax (atomicMax on the bits of a non-negative float)
//   !IS_X: scene points; also resets tau of its 128 columns.
// ---------------------------------------------------------------------------------------------------------------
constexpr int PACK_LD = TILE + 1;

template <bool IS_X>
__global__ void __launch_bounds__(256)
tc_pack_kernel(const float* __restrict__ srcT, int64_t ld_total, const float* __restrict__ nrm,
               const PairDev* __restrict__ pairs, int npairs, int uniform_tiles, int k, int kp, float one_minus_eps,
               float* __restrict__ dst, float* __restrict__ dst_exact, uint32_t* __restrict__ nxmax, uint32_t* __restrict__ tau,
               int* __restrict__ pair_bad, const u64* __restrict__ seed_keys) {
    __shared__ float s[TC_KB * PACK_LD];
    __shared__ float s_aug[3][TILE];
    __shared__ float s_live[TILE];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    int pi, tile;            // tile = 128-row tile within the pair's packed extent
    if (IS_X) {
        pi = find_pair(blockIdx.x, pairs, npairs, uniform_tiles, &PairDev::tc_xtile0);
        tile = blockIdx.x - pairs[pi].tc_xtile0;
    } else {
        pi = find_pair(blockIdx.x, pairs, npairs, uniform_tiles, &PairDev::ytile0);
        tile = blockIdx.x - pairs[pi].ytile0;
    }
    const PairDev* pd = pairs + pi;
    const int n = IS_X ? pd->n1 : pd->n2;
    const int src_tiles = IS_X ? pd->nrt : pd->nct;           // 128-row tiles the k-major source holds
    const int64_t src_col0 = (int64_t)(IS_X ? pd->x_col0 : pd->y_col0) + (int64_t)tile * TILE;
    const int64_t dst_row0 = (int64_t)(IS_X ? pd->xa_row0 : pd->y_col0) + (int64_t)tile * TILE;
    const bool have_src = tile < src_tiles;

    if (tid < TILE) {
        const int r = tile * TILE + tid;
        const bool live = r < n;
        const float nv = (live && have_src) ? nrm[src_col0 + tid] : 0.f;
        float a;
        if (live) {
            a = __fmul_rd(nv, one_minus_eps);
            if (!IS_X) a = __fadd_rd(a, -7.8886091e-31f);     // - 2^-100
            if (!(nv <= 1.2676506e30f)) pair_bad[pi] = 1;     // > 2^100, inf or NaN
        } else {
            a = IS_X ? 1.0e30f : 0.f;
        }
        const float hi = tf32_rna(a);
        const float r1 = __fadd_rn(a, -hi);                    // exact
        const float mid = tf32_rna(r1);
        const float lo = __fadd_rn(r1, -mid);                  // exact, <= 3 significant bits
        s_aug[0][tid] = hi; s_aug[1][tid] = mid; s_aug[2][tid] = lo;
        s_live[tid] = live ? 1.f : 0.f;
        if (IS_X) {
            float m = live ? nv : 0.f;
#pragma unroll
            for (int o = 16; o >= 1; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
            if (lane == 0 && m > 0.f) atomicMax(nxmax + (dst_row0 / TC_N), __float_as_uint(m));
        } else {
            // Threshold at the start of the scan.  A caller that already holds an exact (distance, row) key per scene point
            // from other CAD rows of the same pair (blocked arg-min) passes it: a row of this call matters only if its
            // rounded distance is <= the held one, i.e. c < (d + 1 ulp)^2 <= T0, and LB <= c, so the strict gate LB < T0
            // keeps every such row (the same bound the FP32 kernel uses for its threshold).
            uint32_t t0 = ENC_INF;
            if (seed_keys && live) {                  // keys exist for the n2 scene points only, not for the padding
                const u64 key = seed_keys[dst_row0 + tid];
                const float d = __uint_as_float((uint32_t)(key >> 32));
                if (key != ~0ull && d >= 0.f && d < 3.0e38f) {
                    const float dn = __uint_as_float(__float_as_uint(d) + 1u);
                    t0 = enc_f(__fmul_ru(dn, dn));
                }
            }
            tau[dst_row0 + tid] = t0;
        }
    }
    const int nkb = (kp + TC_KB - 1) / TC_KB;
    for (int kb = 0; kb < nkb; ++kb) {
        __syncthreads();
        // load 32 k-rows x 128 points (coalesced along the points)
        for (int cc = warp; cc < TC_KB; cc += 8) {
            const int c = kb * TC_KB + cc;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int r = q * 32 + lane;
                s[cc * PACK_LD + r] = (c < k && have_src) ? srcT[(int64_t)c * ld_total + src_col0 + r] : 0.f;
            }
        }
        __syncthreads();
        const int c = kb * TC_KB + lane;
        if (c < kp) {
            for (int r = warp; r < TILE; r += 8) {
                float v;
                const bool live = s_live[r] != 0.f;
                if (c < k) v = tf32_rna(s[lane * PACK_LD + r]);
                else if (c < k + 3) v = IS_X ? s_aug[c - k][r] : (live ? 1.f : 0.f);
                else if (c < k + 6) v = IS_X ? 1.f : s_aug[c - k - 3][r];
                else v = 0.f;
                if (!IS_X && !live) v = 0.f;
                dst[(dst_row0 + r) * kp + c] = v;
                if (IS_X && c < ((k + 3) & ~3)) dst_exact[(dst_row0 + r) * ((k + 3) & ~3) + c] = (c < k) ? s[lane * PACK_LD + r] : 0.f;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// TF32 tensor-pipe peak of this device with the screen's instruction shape (roofline denominator): every CTA issues
// `iters` x 4 dependent-free M128 N256 K8 MMAs on shared-memory operands it filled itself; no loads, no epilogue.
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128, 1) tc_peak_kernel(int iters, float* sink) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    float* sA = reinterpret_cast<float*>(smem_raw);
    float* sB = sA + TC_A_BLOCK;
    uint64_t* bar = reinterpret_cast<uint64_t*>(sB + TC_B_STAGE);
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bar + 1);
    const int tid = threadIdx.x, warp = tid >> 5;
    uint32_t x = 0x9e3779b9u * (uint32_t)(blockIdx.x * 128 + tid + 1);
    for (int i = tid; i < TC_A_BLOCK + TC_B_STAGE; i += 128) {
        x = x * 1664525u + 1013904223u;
        sA[i] = tf32_rna((float)(int)(x >> 8) * (1.0f / 8388608.0f) - 1.0f);     // [-1, 1)
    }
    if (tid == 0) { mbar_init(bar, 1); fence_barrier_init(); }
    if (warp == 1) tmem_alloc512(s_tmem);
    fence_proxy_async();          // generic-proxy smem writes -> visible to the tensor core's async proxy
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *s_tmem;
    if (warp == 0) {
        const uint64_t ad = smem_desc_sw128(smem_u32(sA)), bd = smem_desc_sw128(smem_u32(sB));
        for (int it = 0; it < iters; ++it) {
            const uint32_t d = tmem_base + (uint32_t)((it & 1) * TC_N);
            if (elect_one()) {
#pragma unroll
                for (int st = 0; st < 4; ++st) umma_tf32(d, ad + 2 * st, bd + 2 * st, TC_IDESC, (uint32_t)(it > 1 || st > 0));
            }
            __syncwarp();
        }
        if (elect_one()) umma_commit(bar);
        __syncwarp();
        mbar_wait_tc(bar, 0);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == 0) {
        float v[32];
        tmem_ld32(tmem_base, v);
        tmem_ld_wait();
        if (sink && v[0] == 123.456f) sink[tid] = v[1];
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) { tc_fence_after(); tmem_free512(tmem_base); }
}

// TMEM -> register bandwidth (tcgen05.ld 32x32b.x32) with 8 warps per SM and no MMA running: the rate that bounds the
// screen's epilogue (every accumulator element is read once).
__global__ void __launch_bounds__(256, 1) tc_ldtm_kernel(int iters, float* sink) {
    __shared__ uint32_t s_tmem;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) tmem_alloc512(&s_tmem);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t base = s_tmem + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)((warp >> 2) * 256);
    float acc = 0.f;
    for (int it = 0; it < iters; ++it) {
        float v[32];
        tmem_ld32(base + (uint32_t)((it & 7) * 32), v);
        tmem_ld_wait();
        acc += v[0] + v[31];
    }
    if (sink && acc == 123.456f) sink[threadIdx.x] = acc;
    tc_fence_before();
    __syncthreads();
    if (warp == 0) { tc_fence_after(); tmem_free512(s_tmem); }
}

int measure_tmem_ld(double* bytes_per_clk_sm, double* gbs, cudaStream_t st) {
    int dev = 0, sm = 0, khz = 0;
    F2P_CUDA(cudaGetDevice(&dev));
    F2P_CUDA(cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, dev));
    F2P_CUDA(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev));
    F2P_CUDA(cudaFuncSetAttribute(tc_ldtm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 120 * 1024));
    cudaEvent_t e0, e1;
    F2P_CUDA(cudaEventCreate(&e0)); F2P_CUDA(cudaEventCreate(&e1));
    const int iters = 20000;
    double best = 0.0;
    for (int rep = 0; rep < 3; ++rep) {
        F2P_CUDA(cudaEventRecord(e0, st));
        tc_ldtm_kernel<<<sm, 256, 120 * 1024, st>>>(iters, nullptr);
        F2P_CUDA(cudaEventRecord(e1, st));
        F2P_CUDA(cudaEventSynchronize(e1));
        float ms = 0.f;
        F2P_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        const double bps = 8.0 * 4096.0 * iters * sm / (ms * 1e-3);      // 8 warps x 4 KB per load
        if (bps > best) best = bps;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (gbs) *gbs = best / 1e9;
    if (bytes_per_clk_sm) *bytes_per_clk_sm = best / sm / (khz * 1e3);
    return 0;
}

int measure_tf32_peak(double* tflops, cudaStream_t st) {
    int dev = 0, sm = 0;
    F2P_CUDA(cudaGetDevice(&dev));
    F2P_CUDA(cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, dev));
    const size_t smem = 120 * 1024;     // one CTA per SM
    F2P_CUDA(cudaFuncSetAttribute(tc_peak_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaEvent_t e0, e1;
    F2P_CUDA(cudaEventCreate(&e0)); F2P_CUDA(cudaEventCreate(&e1));
    const int iters = 20000;
    double best = 0.0;
    for (int rep = 0; rep < 4; ++rep) {
        F2P_CUDA(cudaEventRecord(e0, st));
        tc_peak_kernel<<<sm, 128, smem, st>>>(iters, nullptr);
        F2P_CUDA(cudaEventRecord(e1, st));
        F2P_CUDA(cudaEventSynchronize(e1));
        float ms = 0.f;
        F2P_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        const double tf = 2.0 * TILE * TC_N * 8.0 * 4.0 * iters * sm / (ms * 1e-3) / 1e12;
        if (rep > 0 && tf > best) best = tf;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (tflops) *tflops = best;
    return 0;
}

// statistics block of the most recent screen (measurement aid, f2p_tc_last_stats)
static unsigned long long* g_last_stats = nullptr;
int tc_last_stats(unsigned long long out[8]) {
    for (int i = 0; i < 8; ++i) out[i] = 0;
    if (!g_last_stats) { set_error("no tensor-pipe call recorded"); return F2P_EINVAL; }
    F2P_CUDA(cudaMemcpy(out, g_last_stats, 8 * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    return 0;
}

// ---- host side ---------------------------------------------------------------------------------------------------
static double tc_eps(int k) { return 1.0 / 1024.0 + 1.0 / 8192.0 + (12.0 * k + 128.0) / 16777216.0; }

int tc_kp(int k) { return (k + 6 + 7) / 8 * 8; }

static size_t tc_smem_bytes(int kp, int eh, int cap, int* nstage_out) {
    const int nkb = (kp + TC_KB - 1) / TC_KB;
    const size_t stage = (size_t)(TC_B_STAGE / TC_CG) * sizeof(float);
    const size_t fixed = (size_t)nkb * TC_A_BLOCK * sizeof(float) + (size_t)eh * cap * TILE * sizeof(u64) + 24 * sizeof(uint64_t) + TILE * sizeof(uint32_t);
    const size_t budget = 227 * 1024 - 1024;            // dynamic shared memory is 1024-aligned by the declaration
    int S = fixed >= budget ? 0 : (int)((budget - fixed) / stage);
    if (S > 8) S = 8;
    *nstage_out = S;
    size_t b = fixed + (size_t)S * stage;
    if (b < 120 * 1024) b = 120 * 1024;                 // one CTA per SM: it owns all 512 TMEM columns
    return b;
}

bool tc_fits(int k, int top5) {
    int S = 0;
    tc_smem_bytes(tc_kp(k), TC_EH, top5 ? TCCfg<5>::CAP : TCCfg<1>::CAP, &S);
    return S >= 2;
}

size_t tc_ws_bytes(int64_t ld1_tc_total, int64_t ld2_total, int k, int64_t tc_items, int64_t total_ytiles, int npairs, int top5) {
    const int kp = tc_kp(k);
    const int cap = top5 ? TCCfg<5>::CAP : TCCfg<1>::CAP;
    size_t b = 0;
    auto add = [&](size_t n) { b += (n + 255) & ~(size_t)255; };
    add(sizeof(float) * (size_t)ld1_tc_total * kp);                 // XA
    add(sizeof(float) * (size_t)ld1_tc_total * ((k + 3) & ~3));     // XR
    add(sizeof(float) * (size_t)ld2_total * kp);                    // YA
    add(sizeof(uint32_t) * (size_t)ld2_total);                      // tau
    add(sizeof(u64) * (size_t)tc_items * TC_CG * TC_EH * cap * TILE);       // cand
    add(sizeof(int) * (size_t)tc_items * TC_CG * TC_EH * TILE);             // ccnt
    add(sizeof(uint32_t) * (size_t)(ld1_tc_total / TC_N) + sizeof(int) * (size_t)(total_ytiles + (npairs > 0 ? npairs : 1)) + 64);   // zeroed block
    return b;
}

void tc_ws_carve(TCWs& w, unsigned char* base, int64_t ld1_tc_total, int64_t ld2_total, int k, int64_t tc_items,
                 int64_t total_ytiles, int npairs, int top5) {
    const int kp = tc_kp(k);
    const int cap = top5 ? TCCfg<5>::CAP : TCCfg<1>::CAP;
    unsigned char* p = base;
    auto take = [&](size_t n) { unsigned char* q = p; p += (n + 255) & ~(size_t)255; return q; };
    w.XA = reinterpret_cast<float*>(take(sizeof(float) * (size_t)ld1_tc_total * kp));
    w.XR = reinterpret_cast<float*>(take(sizeof(float) * (size_t)ld1_tc_total * ((k + 3) & ~3)));
    w.YA = reinterpret_cast<float*>(take(sizeof(float) * (size_t)ld2_total * kp));
    w.tau = reinterpret_cast<uint32_t*>(take(sizeof(uint32_t) * (size_t)ld2_total));
    w.cand = reinterpret_cast<u64*>(take(sizeof(u64) * (size_t)tc_items * TC_CG * TC_EH * cap * TILE));
    w.ccnt = reinterpret_cast<int*>(take(sizeof(int) * (size_t)tc_items * TC_CG * TC_EH * TILE));
    // one zero-initialised block: stats[8] | nxmax[tiles] | tcflag[ytiles] | pair_bad[npairs]
    w.zero_bytes = 64 + sizeof(uint32_t) * (size_t)(ld1_tc_total / TC_N) + sizeof(int) * (size_t)(total_ytiles + (npairs > 0 ? npairs : 1));
    unsigned char* z = take(w.zero_bytes);
    w.zero_base = z;
    w.stats = reinterpret_cast<unsigned long long*>(z);
    w.nxmax = reinterpret_cast<uint32_t*>(z + 64);
    w.tcflag = reinterpret_cast<int*>(w.nxmax + (size_t)(ld1_tc_total / TC_N));
    w.pair_bad = w.tcflag + total_ytiles;
}

template <int K>
static int launch_screen(const TCWs& w, const CUtensorMap& mapXA, const CUtensorMap& mapYA, TCParams prm, int total_items,
                         cudaStream_t stream) {
    int S = 0;
    const size_t smem = tc_smem_bytes(prm.kp, TC_EH, TCCfg<K>::CAP, &S);
    if (S < 2) { set_error("tensor-pipe screen: k too large for the shared-memory ring"); return F2P_ENOTSUP; }
    prm.nstage = S;
    auto kern = nn_tc_screen_kernel<K, 4 * TC_EH, TC_CG>;
    F2P_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)total_items * TC_CG); cfg.blockDim = dim3(64 + 32 * 4 * TC_EH);
    cfg.dynamicSmemBytes = smem; cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = TC_CG; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    F2P_CUDA(cudaLaunchKernelEx(&cfg, kern, mapXA, mapYA, prm));
    return 0;
}

// Packs the operands, screens on the tensor pipe and verifies the candidates.  K = 1: merges exact keys into `best`
// (pre-initialised to ~0); K = 5: writes nn_query's candidates.  Afterwards w.tcflag marks the column tiles the caller
// must still run through the FFMA2 kernels.
int launch_tc(int top5, const float* xT, int64_t ld1_total, const float* nx, const float* yT, int64_t ld2_total,
              const float* ny, int k, const PairDev* pairs_dev, int npairs, int uniform_tc_items, int total_tc_items,
              int uniform_tc_xtiles, int total_tc_xtiles, int uniform_ytiles, int total_ytiles, int64_t ld1_tc_total,
              unsigned flags, u64* best, int64_t* out5, int cand_is_out_layout, const TCWs& w, cudaStream_t stream,
              const u64* seed_keys) {
    if (total_tc_items <= 0 || total_ytiles <= 0) return 0;
    const int kp = tc_kp(k);
    const double eps = tc_eps(k);
    float ome = (float)(1.0 - eps);
    if ((double)ome > 1.0 - eps) ome = nextafterf(ome, 0.f);
    float omega = (float)(2.0 * eps + 1.0 / 524288.0);
    if ((double)omega < 2.0 * eps + 1.0 / 524288.0) omega = nextafterf(omega, 1.f);

    F2P_CUDA(cudaMemsetAsync(w.zero_base, 0, w.zero_bytes, stream));
    g_last_stats = w.stats;
    tc_pack_kernel<true><<<total_tc_xtiles, 256, 0, stream>>>(xT, ld1_total, nx, pairs_dev, npairs, uniform_tc_xtiles, k, kp,
                                                             ome, w.XA, w.XR, w.nxmax, nullptr, w.pair_bad, nullptr);
    tc_pack_kernel<false><<<total_ytiles, 256, 0, stream>>>(yT, ld2_total, ny, pairs_dev, npairs, uniform_ytiles, k, kp,
                                                           ome, w.YA, nullptr, nullptr, w.tau, w.pair_bad, seed_keys);
    F2P_CUDA(cudaGetLastError());

    CUtensorMap mapXA, mapYA;
    int rc = make_tensor_map_2d(&mapXA, w.XA, (uint64_t)kp, (uint64_t)ld1_tc_total, (uint64_t)kp, TC_KB, TC_N / TC_CG, 1);
    if (rc) return rc;
    rc = make_tensor_map_2d(&mapYA, w.YA, (uint64_t)kp, (uint64_t)ld2_total, (uint64_t)kp, TC_KB, TILE, 1);
    if (rc) return rc;

    TCParams prm;
    prm.pairs = pairs_dev; prm.npairs = npairs; prm.uniform_items = uniform_tc_items; prm.k = k; prm.kp = kp; prm.nstage = 0;
    prm.ny = ny; prm.nxmax = reinterpret_cast<const float*>(w.nxmax); prm.tau = w.tau; prm.cand = w.cand; prm.ccnt = w.ccnt;
    prm.tcflag = w.tcflag; prm.pair_bad = w.pair_bad; prm.omega = omega; prm.stats = w.stats;
    rc = top5 ? launch_screen<5>(w, mapXA, mapYA, prm, total_tc_items, stream)
              : launch_screen<1>(w, mapXA, mapYA, prm, total_tc_items, stream);
    if (rc) return rc;

    TCVerifyParams vp;
    vp.pairs = pairs_dev; vp.npairs = npairs; vp.uniform_ytiles = uniform_ytiles; vp.k = k; vp.eh = TC_EH;
    vp.cap = top5 ? TCCfg<5>::CAP : TCCfg<1>::CAP;
    vp.XR = w.XR; vp.nx = nx; vp.yT = yT; vp.ld2_total = ld2_total; vp.ny = ny;
    vp.tau = w.tau; vp.cand = w.cand; vp.ccnt = w.ccnt; vp.tcflag = w.tcflag; vp.best = best; vp.out5 = out5;
    vp.cand_is_out_layout = cand_is_out_layout; vp.flags = flags; vp.stats = w.stats;
    const size_t vsmem = (size_t)k * TILE * sizeof(float);
    if (top5) {
        F2P_CUDA(cudaFuncSetAttribute(nn_tc_verify_kernel<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)vsmem));
        nn_tc_verify_kernel<5><<<total_ytiles, TILE, vsmem, stream>>>(vp);
    } else {
        F2P_CUDA(cudaFuncSetAttribute(nn_tc_verify_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)vsmem));
        nn_tc_verify_kernel<1><<<total_ytiles, TILE, vsmem, stream>>>(vp);
    }
    F2P_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace f2p

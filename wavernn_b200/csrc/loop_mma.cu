// Per-sample loop, tensor-core kernel for large fold batches (config 5: hundreds of folds per GPU).
//
// `generate`'s hot loop (models/fatchord_version.py:201-241) for B folds at once is, per step, a chain of skinny
// GEMMs  out[B x rows] = act[B x K] . W[rows x K]^T:  rnn1 (:210), rnn2 (:213-214), fc1 (:217-218), fc2 (:220-221),
// fc3 (:223).  The FMA kernels (loop_res.cu ...) spend the same time per fold whatever B is; here all five layers run
// on the 5th-generation tensor cores:
//
//   * folds are the M operand: a fold tile is 128 folds = the 128 TMEM lanes, weight rows are N (96 = [r|z|n] x 32 units
//     for a GRU job, 64 for fc1 / fc2, 32 for fc3); `tcgen05.mma` cta_group::1 kind::f16 issued by one thread,
//     accumulators in TMEM, read back with `tcgen05.ld` by the thread that owns the fold: gates, ReLU and the MoL draw
//     (utils/distribution.py:87-123) are thread-local, no shared-memory staging and no shuffles;
//   * precision: every operand is a bf16 hi + bf16 lo pair (x = hi + lo to ~2^-17), three MMAs per k-step
//     (hi.hi + hi.lo + lo.hi) accumulated in fp32: products carry ~1e-5 relative error, far inside the 1e-3 bound
//     on teacher-forced logits that plain bf16 breaks after a few thousand recurrent steps;
//   * operands live in global memory (L2-resident) already in the K-major, no-swizzle shared-memory image the MMA
//     descriptors address, "[k-chunk of 8][row][8 elements]", so a pipeline stage is two plain `cp.async.bulk`
//     copies (one weight block, one activation block) completing on an mbarrier: no tensor maps, no swizzle;
//   * fold tiles are independent recurrences (:194-196), so nothing is grid-wide: a tile is served by 16 GRU CTAs
//     (32 units each) and 8 fc CTAs (64 rows each, one of them also runs fc3 + the sampler), which hand operands to one
//     another through L2 and five per-tile counters (rnn1, rnn2, fc1, fc2, sample published);
//   * only what depends on the step before is on the sample-to-sample chain.  The conditioning columns (mel, aux) of
//     every layer are known for all steps: a pre-pass lays them out as MMA operands for the whole call and their
//     products are issued one step AHEAD into the accumulators; the recurrent products (W_hh . h) are formed one step
//     ahead too; the previous sample, the one scalar that is new, enters rnn1 / rnn2 / fc1 as a rank-1 fp32 term in
//     the epilogues.  On the chain per step: rnn1 gates -> 8 stages of K = 64 (rnn2) -> gates -> 8 (fc1) -> 8 (fc2) -> 8 (fc3)
//     -> draw;
//   * algebra: the I layer and both residual adds are folded into composed matrices formed on the host in double
//     (W_ih1 W_I etc., as in loop_res.cu).
//
// Warp roles in a CTA (320 threads): warp 0 = TMA producer (8 lanes), warp 1 = MMA issuer (owns TMEM), warps 2..9 =
// epilogues (TMEM -> registers -> global operands).
#include <cuda_bf16.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "internal.h"
#include "loop_common.cuh"
#include "exchange.cuh"

namespace wrnn {

namespace mma {

constexpr int H = 512, FCD = 512, AUX = 32, FEAT = 80;
constexpr int UT = 32;                     // GRU units per row tile: weight rows [r(32) | z(32) | n(32)] = N = 96 of an MMA
constexpr int GRU_ROWS = 3 * UT, FC_ROWS = 64;
constexpr int FT = 128;                    // folds per tile = M of the MMA = TMEM lanes: lane i of the accumulators is fold i
constexpr int MT_GRU = H / UT, MT_FC = FCD / FC_ROWS;     // 16, 8 row tiles
constexpr int MAXTILES = 16;               // fold tiles per launch (counter table; a B200 holds 9)
constexpr int KC = 8;                      // k-chunks (of 8 elements) per pipeline stage: K = 64 = four MMA k-steps.  The TMA engine
                                           // spends ~360 clocks on a bulk copy of 4..16 KB whatever its size: few, large copies
constexpr int MAXSTAGE = 10;               // ring depth is chosen at launch: as many stages as shared memory holds
constexpr int W_GRU = 2 * KC * GRU_ROWS * 16, W_FC = 2 * KC * FC_ROWS * 16;     // weight block of a stage, hi + lo: 24576 / 16384 B
constexpr int ACT_STAGE = 2 * KC * FT * 16;                                      // activation block of a stage, hi + lo: 32768 B
constexpr int CGROUPS = 4;                 // conditioning operand: 32 chunks = 256 elements per fold (209 used)
constexpr int HGROUPS = H / (8 * KC);      // 8 stage groups in a 512-wide operand
constexpr int EPI_PARTS = 2;               // epilogue warps per TMEM lane quarter: each takes 1 / EPI_PARTS of an accumulator's columns.
                                           // 4 (16 warps, 96 registers) was measured: no faster.  The GRU gate math is bound by the
                                           // SFU: 128 folds x 32 units x 6 MUFU at 16 per clock = 1.5 k clk per epilogue however many warps
constexpr int EPI_WARPS = 4 * EPI_PARTS, EPI_THREADS = 32 * EPI_WARPS, THREADS = 64 + EPI_THREADS;
constexpr int UPT = UT / EPI_PARTS, RPT = FC_ROWS / EPI_PARTS;      // GRU units / fc rows per epilogue thread: 16 / 32
constexpr int NCP = 32;                    // fc3 rows padded (MoL: 30) = N of the fc3 job
// this CTA's bias rows and previous-sample columns in shared memory
constexpr int SB_G = 0;                                  // rnn1 [input 96 | recurrent 96], rnn2 [input 96 | recurrent 96]
constexpr int SB_XG = SB_G + 4 * GRU_ROWS;               // previous-sample column of rnn1 [96], rnn2 [96]
constexpr int SBIAS_GRU = SB_XG + 2 * GRU_ROWS;          // a CTA without an fc job keeps only these (and gets a fourth ring stage)
constexpr int SB_F = SBIAS_GRU;                          // fc1 [64], fc2 [64]
constexpr int SB_XF = SB_F + 2 * FC_ROWS;                // previous-sample column of fc1 [64]
constexpr int SB_B3 = SB_XF + FC_ROWS;                   // fc3 bias [32]
constexpr int SBIAS = SB_B3 + NCP;
// fc3 inside the fc CTAs: this CTA's 64 fc2 outputs of the 128 folds as an MMA operand (K = 64, hi + lo) and the matching 64
// columns of fc3 (32 rows, hi + lo), both in shared memory for the whole call, between the stage ring and the bias rows
constexpr int P3_A = 2 * KC * FT * 16, P3_B = 2 * KC * NCP * 16;       // 32768 + 8192 B
constexpr int TMEM_COLS = 512;
constexpr int FIXED_SMEM = P3_A + P3_B + SBIAS * 4;
// conditioning operand element order (which features each layer concatenates):
//   [a4 (32) | a2 (32) | m (80) | a1 (32) | a3 (32) | zero (48)]
// in groups of 64: rnn1 reads groups 1..2, rnn2 0..2, fc1 1..3, fc2 group 0 (weights of the features a layer does not
// read are zero); every 8-element chunk is 8 consecutive features of one source.  The previous sample has no slot: it
// enters in the epilogues.
constexpr int C_A4 = 0, C_A2 = 32, C_M = 64, C_A1 = 144, C_A3 = 176, C_END = 208;

// Job types.  A job = one MMA chain of (row tile, fold tile) into one TMEM accumulator.
//   ..A ("ahead"): the conditioning columns of a layer for step t + 1, issued during step t (start the accumulator);
//   ..C ("chain"): the columns that depend on the step's own activations (accumulate on top);
//   S1 / S2: recurrent products W_hh . h' of the NEXT step (loop_res.cu's "shadow products", on tensor cores).
enum { J_R1A = 0, J_R2A, J_R2C, J_S1, J_S2, J_F1A, J_F1C, J_F2A, J_F2C, NJOB };
enum { ARR_C = 0, ARR_H1NEW, ARR_H2NEW, ARR_S12, ARR_Y1 };
enum { PH_RNN1 = 0, PH_RNN2, PH_FC1, PH_FC2, PH_SAMPLE, PH_H2, NPHASE };     // per-tile counters: who has published what
// (PH_RNN2 = h1' + h2' is out, which is all fc1 waits for; PH_H2 = rnn2's own new state is out too, for the recurrent products)
enum { TF_R1 = 0, TF_R2, TF_F1, TF_F2, TF_P3, NTF };                   // accumulator-ready barriers, one per epilogue
// TMEM accumulator (first column) of each job: 128 lanes = folds, columns = weight rows.  rnn1's gates read X1 + H1,
// rnn2's X2 + H2 (96 columns each: [r | z | n] x 32 units); fc1, fc2 64 columns each; fc3 reuses fc1's first 32.
__host__ __device__ inline int job_acc(int jt) {
    switch (jt) {
        case J_R1A: return 0;
        case J_S1: return 96;
        case J_R2A: case J_R2C: return 192;
        case J_S2: return 288;
        case J_F2A: case J_F2C: return 448;
        default: return 384;                // fc1
    }
}
__host__ __device__ inline int job_stages(int jt) {
    switch (jt) {
        case J_R1A: return 2;
        case J_R2A: case J_F1A: return 3;
        case J_F2A: return 1;
        default: return HGROUPS;
    }
}
__host__ __device__ inline int job_rows(int jt) { return jt >= J_F1A ? FC_ROWS : GRU_ROWS; }
__host__ __device__ inline int job_tiles(int jt) { return jt >= J_F1A ? MT_FC : MT_GRU; }
__host__ __device__ inline bool job_starts_acc(int jt) { return jt != J_R2C && jt != J_F1C && jt != J_F2C; }
__host__ __device__ inline int job_commit(int jt) {      // which accumulator-ready barrier a finished job arrives on, or -1
    return jt == J_R1A ? TF_R1 : jt == J_R2C ? TF_R2 : jt == J_F1C ? TF_F1 : jt == J_F2C ? TF_F2 : -1;
}
struct StageInfo { int arr, g; };
__host__ __device__ inline StageInfo stage_info(int jt, int s) {
    switch (jt) {
        case J_R1A: return StageInfo{ARR_C, 1 + s};
        case J_R2A: return StageInfo{ARR_C, s};
        case J_R2C: case J_S1: return StageInfo{ARR_H1NEW, s};
        case J_S2: return StageInfo{ARR_H2NEW, s};
        case J_F1A: return StageInfo{ARR_C, 1 + s};
        case J_F1C: return StageInfo{ARR_S12, s};
        case J_F2A: return StageInfo{ARR_C, 0};
        default: return StageInfo{ARR_Y1, s};          // J_F2C
    }
}

struct Params {
    LoopArgs a;
    int NTILES, G, NC;
    int dbg_state;                 // WRNN_MMA_STEPS set: also store the fp32 state / fc2 output (wrnn_debug_mma_state)
    int64_t steps_run;             // == a.S except under WRNN_MMA_STEPS (bring-up: stop after n steps and inspect the state)
    int smem_total;                // dynamic shared memory of a CTA: stage ring | fc3 operands (fc CTAs) | bias rows
    const uint8_t *Aimg;           // packed weights: job type j at offA[j], [row tile][stage][hi|lo][KC][rows][8] bf16
    int64_t offA[NJOB];
    int64_t offP3;                 // fc3 columns as an operand: [8 fc row tiles][hi|lo][KC][32 rows][8] bf16, in Aimg
    const float *bias;             // rnn1 [16][2][96] | rnn2 [16][2][96] | fc1 [512] | fc2 [512] | x columns rnn1 [16][96] | rnn2 [16][96] | fc1 [512] | fc3 bias [32]
    uint8_t *H1[2], *H2[2], *S12, *Y1;          // activation operands, [fold tile][group][hi|lo][KC][128][8] bf16
    float *part3;                  // fc3 partial sums of every fc CTA: [fold tile][8 row tiles][8 groups of 4 logits][128 folds] float4
    const uint8_t *Cop;            // conditioning operand of every step: [S][fold tile][7 groups][hi|lo][KC][128][8] bf16
    uint2 *xp;                     // previous sample of every fold, tagged: {float bits, step it is the input of}, [NTILES * 128]
    float *h1f, *h2f, *y2f;        // debug only: [NTILES * 128][512] fp32 state / fc2 output
    GridSync gs;
    long long *prof;               // debug (WRNN_MMA_PROF=step): clock64 timeline of that step from CTA 0 and CTA G-1, else null
    int64_t prof_step;
    int exp;                       // debug (WRNN_MMA_EXP): timing experiments that break the results; 0 in production
};

#ifndef WRNN_STEP_MARK_MASK
#define WRNN_STEP_MARK_MASK 0x1ff
#endif
#define STEP_MARK(k)                                                                                  \
    do {                                                                                              \
        if (((WRNN_STEP_MARK_MASK) >> (k)) & 1)                                                       \
            if (P.prof && t >= P.prof_step && t < P.prof_step + 64) P.prof[192 + (k) * 64 + (t - P.prof_step)] = clock64(); \
    } while (0)
#define PROF_MARK(slot)                                                                              \
    do {                                                                                             \
        if (P.prof && t == P.prof_step && (cta == 0 || cta == G - 1)) P.prof[(cta == 0 ? 0 : 64) + (slot)] = clock64(); \
    } while (0)

// ---------------------------------------------------------------------------- PTX
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(void *b, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(void *b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(void *b) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(void *b, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(smem_u32(b)), "r"(parity) : "memory");
    return ok != 0;
}
// non-blocking test (`try_wait` may suspend the thread): used where the wait is short and on the sample-to-sample chain
__device__ __forceinline__ bool mbar_test_wait(void *b, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(smem_u32(b)), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, void *b) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(b)) : "memory");
}
// generic-proxy writes (st.global by the epilogue threads) are read by the TMA unit (async proxy) of other SMs: EVERY writing
// thread orders its own stores against the async proxy before the CTA publishes the phase, and the producer thread does the
// same after it has acquired the grid counter (the pattern of a TMA store after generic smem writes, applied to global)
// (the .global form: the plain `fence.proxy.async` also covers shared memory and cost 2.2 us of a 36 us step, five publishes)
__device__ __forceinline__ void fence_proxy_async_global() { asm volatile("fence.proxy.async.global;" ::: "memory"); }
__device__ __forceinline__ void tmem_alloc(uint32_t *dst_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accum) : "memory");
}
__device__ __forceinline__ void umma_commit(void *bar) {     // arrives on `bar` when every MMA issued so far has completed
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {   // 32 lanes x 32 columns: thread i <- lane i
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
          "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
          "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
          "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_ld8_nowait(uint32_t taddr, uint32_t (&r)[8]) {    // thread i <- lane i, 8 consecutive columns
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// Shared-memory matrix descriptor, K-major, no swizzle: element (row r, k) of the operand sits at
//   start + (r % 8) * 16 + (r / 8) * SBO + (k / 8) * LBO + (k % 8) * 2   bytes.
// With the image [k-chunk][rows][8]: SBO = 128 (eight 16-byte rows), LBO = rows * 16.
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((saddr >> 4) & 0x3fff) | ((uint64_t)((lbo >> 4) & 0x3fff) << 16) | ((uint64_t)((sbo >> 4) & 0x3fff) << 32) |
           ((uint64_t)1 << 46);
}
// instruction descriptor: fp32 accumulate, A = B = bf16, both K-major, N >> 3 at bit 17, M >> 4 at bit 24
__device__ __forceinline__ uint32_t make_idesc(int M, int N) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// ---------------------------------------------------------------------------- small helpers
struct Watch {                  // a lost arrival becomes an error flag, not a hung GPU box
    int *abort_flag;
    bool dead;
    int site;                   // which wait this thread is in (reported in the abort code): role * 16 + detail
};
__device__ __forceinline__ bool give_up(Watch &w, long long t0, unsigned &spins) {
    if ((++spins & 1023u) == 0u) {
        if (clock64() - t0 > WATCHDOG_CLOCKS) { atomicCAS(w.abort_flag, 0, 0x10000 | (w.site << 8) | (int)(blockIdx.x & 0xff)); w.dead = true; }
        else if (ld_volatile_s32(w.abort_flag)) w.dead = true;
    }
    return w.dead;
}
__device__ __forceinline__ void mbar_wait(void *b, uint32_t parity, Watch &w) {
    if (w.dead || mbar_try_wait(b, parity)) return;
    const long long t0 = clock64();
    unsigned spins = 0;
    while (!mbar_try_wait(b, parity))
        if (give_up(w, t0, spins)) return;
}
// Hand-over between the CTAs of one fold tile: ONE MONOTONIC COUNTER PER (phase, fold tile), each on its own 128-byte
// line.  The CTAs that produce a phase's operands (16 GRU CTAs, 8 fc CTAs, 1 sampler) add one each per step once their
// stores are ordered; a consumer waits for arrivals * (t + 1).  Fold tiles never wait for one another.
__device__ __forceinline__ unsigned long long *tile_counter(const GridSync &gs, int phase, int nt) {
    return gs.counter + 16 * (1 + phase * MAXTILES + nt);
}
__device__ __forceinline__ int phase_arrivals(int phase) { return phase <= PH_RNN2 || phase == PH_H2 ? MT_GRU : phase <= PH_FC2 ? MT_FC : 1; }
__device__ __forceinline__ void tile_wait(const GridSync &gs, int phase, int nt, int64_t steps_done, Watch &w) {
    if (w.dead || steps_done <= 0) return;
    const unsigned long long want = (unsigned long long)phase_arrivals(phase) * (unsigned long long)steps_done;
    const long long t0 = clock64();
    unsigned spins = 0;
    const unsigned long long *c = tile_counter(gs, phase, nt);
    while (ld_acquire_u64(c) < want)
        if (give_up(w, t0, spins)) return;
}
// Reconverge the warp before the named barrier (see publish() in the kernel: a warp that stays split in two fragments
// across barriers and collectives runs them in turns on the hardware's forward-progress timer, ~20 k clocks a turn).
__device__ __forceinline__ void epi_bar() {
    __syncwarp();
    bar_sync(1, EPI_THREADS);
}
// The previous sample travels as ONE 8-byte word {value, step}: the reader polls the word itself (one L2 round trip once it
// is there) instead of a counter followed by the data.  Steps only grow, so the slot needs neither reset nor a second buffer:
// the sampler of step t writes {x, t + 1} after every reader of {.., t} is through (they all precede its own inputs).
__device__ __forceinline__ uint2 ld_volatile_u2(const uint2 *p) {
    uint2 v;
    asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float wait_prev_sample(const uint2 *slot, int64_t t, Watch &w) {
    uint2 v = ld_volatile_u2(slot);
    if (v.y != (uint32_t)t && !w.dead) {
        const long long t0 = clock64();
        unsigned spins = 0;
        while ((v = ld_volatile_u2(slot)).y != (uint32_t)t)
            if (give_up(w, t0, spins)) break;
    }
    return __uint_as_float(v.x);
}

// bf16 hi / lo split of four floats, packed: .x/.y = hi pairs, .z/.w = lo pairs
__device__ __forceinline__ void split4(const float (&v)[4], uint2 &hi, uint2 &lo) {
    __nv_bfloat16 h[4], l[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        h[i] = __float2bfloat16_rn(v[i]);
        l[i] = __float2bfloat16_rn(v[i] - __bfloat162float(h[i]));
    }
    hi.x = (uint32_t)__bfloat16_as_ushort(h[0]) | ((uint32_t)__bfloat16_as_ushort(h[1]) << 16);
    hi.y = (uint32_t)__bfloat16_as_ushort(h[2]) | ((uint32_t)__bfloat16_as_ushort(h[3]) << 16);
    lo.x = (uint32_t)__bfloat16_as_ushort(l[0]) | ((uint32_t)__bfloat16_as_ushort(l[1]) << 16);
    lo.y = (uint32_t)__bfloat16_as_ushort(l[2]) | ((uint32_t)__bfloat16_as_ushort(l[3]) << 16);
}
// eight floats -> one 16-byte chunk of the hi image and one of the lo image
__device__ __forceinline__ void split8(const float (&v)[8], uint4 &hi, uint4 &lo) {
    const float a[4] = {v[0], v[1], v[2], v[3]}, b[4] = {v[4], v[5], v[6], v[7]};
    uint2 h0, l0, h1, l1;
    split4(a, h0, l0);
    split4(b, h1, l1);
    hi = make_uint4(h0.x, h0.y, h1.x, h1.y);
    lo = make_uint4(l0.x, l0.y, l1.x, l1.y);
}
// byte offset of the 16-byte slot (fold fi of tile nt, feature chunk `chunk`) in a B operand with NG stage groups
__device__ __forceinline__ size_t b_slot(int NT, int NG, int nt, int chunk, int fi, int hl) {
    return ((size_t)(nt * NG + chunk / KC) * 2 + hl) * (size_t)(KC * NT * 16) + (size_t)((chunk % KC) * NT + fi) * 16;
}

// ---------------------------------------------------------------------------- MoL draw, one thread per fold
// Same arithmetic as mol_sample_warp (loop_common.cuh; utils/distribution.py:87-123): Gumbel-max over the mixture
// logits (first index wins a tie), then mean + exp(max(log_scale, -32.236...)) * (log u - log(1 - u)), clamped to [-1, 1].
constexpr int NMIX = 10;        // MOL: n_classes = 30 = 3 x 10 (fatchord_version.py:103-104)
// The noise terms do not depend on the logits: `gum[i]` = log(-log(u1[i])) and `lgt` = log u2 - log(1 - u2) are formed
// while the fc3 products are still under way (mol_noise_thread), the same values the warp sampler subtracts / multiplies.
__device__ __forceinline__ float mol_sample_thread(const float (&lg)[NCP], const float (&gum)[NMIX], float lgt, int *pick) {
    float best = -INFINITY, mean = 0.f, ls = 0.f;
    int idx = 0;
#pragma unroll
    for (int i = 0; i < NMIX; i++) {
        const float tv = lg[i] - gum[i];
        if (tv > best) { best = tv; idx = i; mean = lg[NMIX + i]; ls = lg[2 * NMIX + i]; }
    }
    ls = fmaxf(ls, -32.23619130191664f);
    float x = mean + expf(ls) * lgt;
    x = fminf(fmaxf(x, -1.f), 1.f);
    *pick = idx;
    return x;
}

// ---------------------------------------------------------------------------- conditioning operand, all steps (pre-pass)
// The conditioning features of every (step, fold) as the MMA operand image, straight from the frame-rate tensors: the
// mel part is formed from the padded frames by the polyphase table (UpsampleNetwork's repeat + FIR stages and the crop,
// fatchord_version.py:86-88; same taps, same order of operations as fold_cond_kernel, so the values are the ones the
// FMA kernels read), the x275 nearest-repeat of aux (Stretch2d, :83-85) and the fold offsets (:293-340) are indices.
// No fp32 fold batch is written for this kernel.  Thread = (step, fold tile, chunk of 8 features, fold): the 128 folds of
// a tile are consecutive threads, so every store instruction writes 2 KB of contiguous operand.
struct CondSrc {
    const float *melT, *res_out, *poly;         // [frames][80], [frames][128], [hop][taps]
    const FoldMeta *meta;                       // per fold of this launch
    int hop, indent, taps, kmin;
};
__global__ void __launch_bounds__(256) cond_operand_kernel(const CondSrc c, int64_t B, int64_t S, int ntiles, uint8_t *__restrict__ Cop) {
    const int fi = threadIdx.x & (FT - 1);
    const int64_t blk = blockIdx.x;
    const int chunk = (int)(blk % 16) * 2 + (threadIdx.x >> 7);         // 0..31
    const int64_t rest = blk / 16;
    const int nt = (int)(rest % ntiles);
    const int64_t t = rest / ntiles, b = (int64_t)nt * FT + fi;
    float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    const int e = chunk * 8;
    if (b < B && e < C_END) {
        const FoldMeta m = c.meta[b];
        const int64_t l = m.start + t;
        if (l < m.L) {
            const int li = (int)l;
            if (e >= C_M && e < C_A1) {
                const int f = e - C_M, ii = li + c.indent, fr = ii / c.hop, ph = ii - fr * c.hop;
                float4 v0 = make_float4(0.f, 0.f, 0.f, 0.f), v1 = v0;
                for (int k = 0; k < c.taps; k++) {
                    const int j = fr + c.kmin + k;
                    if (j < 0 || j >= m.Tp) continue;
                    const float wv = __ldg(c.poly + ph * c.taps + k);
                    const float4 *src = reinterpret_cast<const float4 *>(c.melT + (size_t)(m.pm_off + j) * FEAT + f);
                    const float4 q0 = __ldg(src), q1 = __ldg(src + 1);
                    v0.x = fmaf(wv, q0.x, v0.x); v0.y = fmaf(wv, q0.y, v0.y); v0.z = fmaf(wv, q0.z, v0.z); v0.w = fmaf(wv, q0.w, v0.w);
                    v1.x = fmaf(wv, q1.x, v1.x); v1.y = fmaf(wv, q1.y, v1.y); v1.z = fmaf(wv, q1.z, v1.z); v1.w = fmaf(wv, q1.w, v1.w);
                }
                v[0] = v0.x; v[1] = v0.y; v[2] = v0.z; v[3] = v0.w; v[4] = v1.x; v[5] = v1.y; v[6] = v1.z; v[7] = v1.w;
            } else {
                // aux rows are [a1 | a2 | a3 | a4] (fatchord_version.py:203-206); the operand order is [a4 | a2 | . | a1 | a3]
                const int a = e < C_A2 ? 3 * AUX + (e - C_A4) : e < C_M ? AUX + (e - C_A2) : e < C_A3 ? (e - C_A1) : 2 * AUX + (e - C_A3);
                const float4 *src = reinterpret_cast<const float4 *>(c.res_out + (size_t)(m.res_off + li / c.hop) * (4 * AUX) + a);
                const float4 q0 = __ldg(src), q1 = __ldg(src + 1);
                v[0] = q0.x; v[1] = q0.y; v[2] = q0.z; v[3] = q0.w; v[4] = q1.x; v[5] = q1.y; v[6] = q1.z; v[7] = q1.w;
            }
        }
    }
    uint4 hi, lo;
    split8(v, hi, lo);
    uint8_t *base = Cop + ((size_t)t * ntiles) * CGROUPS * ACT_STAGE;
    __stcs(reinterpret_cast<uint4 *>(base + b_slot(FT, CGROUPS, nt, chunk, fi, 0)), hi);      // streamed: read once, a step at a time
    __stcs(reinterpret_cast<uint4 *>(base + b_slot(FT, CGROUPS, nt, chunk, fi, 1)), lo);
}

// ---------------------------------------------------------------------------- pipeline: ring, producer, issuer
// stage block a job reads at step t.  Activation arrays alternate with the step; the conditioning operand is indexed by it.
__device__ __forceinline__ const uint8_t *stage_src(const Params &P, StageInfo si, int nt, int64_t t) {
    switch (si.arr) {
        case ARR_C: {
            const int64_t tt = t < P.a.S ? t : P.a.S - 1;
            return P.Cop + (((size_t)tt * P.NTILES + nt) * CGROUPS + si.g) * ACT_STAGE;
        }
        case ARR_H1NEW: return P.H1[(t + 1) & 1] + (size_t)(nt * HGROUPS + si.g) * ACT_STAGE;
        case ARR_H2NEW: return P.H2[(t + 1) & 1] + (size_t)(nt * HGROUPS + si.g) * ACT_STAGE;
        case ARR_S12: return P.S12 + (size_t)(nt * HGROUPS + si.g) * ACT_STAGE;
        default: return P.Y1 + (size_t)(nt * HGROUPS + si.g) * ACT_STAGE;
    }
}

struct Ring {                   // the stage ring of this CTA and a position in it (producer and MMA issuer keep one each)
    uint32_t nstage, w_slot, stage_bytes;       // depth, bytes reserved for the weight block of a stage, bytes per stage
    uint32_t slot, par;
    __device__ __forceinline__ void advance() {
        if (++slot == nstage) { slot = 0; par ^= 1u; }
    }
};
// A CTA that only runs fc jobs (8 KB weight blocks) packs its stages tighter than one with a GRU job (12 KB); the ring
// is as deep as shared memory allows: 4 stages of 56 KB for a CTA with only a GRU job, 3 of 48 KB beside the fc3 operands.
__device__ __forceinline__ int ring_bytes_of(const Params &P, bool has_fc) {
    return (P.smem_total - (has_fc ? P3_A + P3_B + SBIAS * 4 : SBIAS_GRU * 4)) & ~127;
}
__device__ __forceinline__ Ring make_ring(const Params &P, bool has_gru, bool has_fc) {
    Ring r;
    r.w_slot = has_gru ? W_GRU : W_FC;
    r.stage_bytes = r.w_slot + ACT_STAGE;
    r.nstage = min((uint32_t)MAXSTAGE, (uint32_t)ring_bytes_of(P, has_fc) / r.stage_bytes);
    r.slot = 0;
    r.par = 0;
    return r;
}

// One job of the TMA producer lanes: (wait for a tile counter,) stream the stages of (job type, row tile, fold tile).
// A bulk-copy INSTRUCTION occupies its warp for ~320 clocks whatever its size and however many lanes are active, and a
// thread completes about one copy per 700 clocks (scripts/mb_umma.cu; measured here: lanes taking turns with one copy
// per instruction moved a stage per 640 clocks whatever its size; eight lanes issuing four stages at once reach the
// L2 -> SM limit of ~64 B/clk but deliver the first stage late).  So a stage is ONE instruction of a lane pair, the even
// lane bringing the weight block, the odd lane the activation block, and the four pairs of the NPROD = 8 lanes take the
// stages in turn (min(4, ring depth) of them): in order, a stage per ~320 clocks of issue, ~900 of transfer.  The even lane arms the barrier with the
// byte count of the stage; a copy may complete its bytes before the arrive (the phase cannot end before it).
// The weight blocks do not depend on anybody: for a gated job up to a ring-full of stages is armed and gets its weight
// block BEFORE the wait on the tile counter, so that only activation bytes (half a block per lane) move afterwards.
// `t`: the step whose operands the job reads (conditioning row t; activation buffers written during step t).
// `gate`: phase whose counter must show `gate_steps` published steps of fold tile nt before the loads start, or -1.
constexpr int NPROD = 8;
struct Pipe {
    Ring r;
    uint32_t count;
    uint8_t *stages;
    uint64_t *bar_full, *bar_empty;
};
__device__ __forceinline__ void produce_job(const Params &P, int jt, int mt, int nt, int64_t t, int gate, int64_t gate_steps, Pipe &pp,
                                            int lane, Watch &w, long long *pacc = nullptr, long long *pgate = nullptr) {
    const int nst = ((P.exp & 1) && (jt == J_S1 || jt == J_S2)) ? 1 : job_stages(jt);
    const uint32_t wbytes = 2u * KC * (uint32_t)job_rows(jt) * 16u;
    const uint8_t *Wbase = P.Aimg + P.offA[jt] + (size_t)mt * nst * wbytes;
    Ring &r = pp.r;
    const bool wlane = (lane & 1) == 0;
    long long waited = 0;           // debug timeline only
    // The lane pairs run independently, each waiting on the "slot empty" barriers of its own stages by phase parity: a pair
    // must never be a whole ring ahead of the pair that owns the stage before it in the same slot, or its parity test
    // would be answered by the phase before.  Its own previous stage bounds its lead to (pairs) stages: pairs <= ring depth.
    const uint32_t npair = min(4u, r.nstage);
    const int early = (gate >= 0 && !(P.exp & 2)) ? min(nst, (int)r.nstage) : 0;
    {
        Ring re = r;
        uint32_t cnt = pp.count;
        for (int s = 0; s < early; s++, cnt++) {
            if (cnt % npair == (uint32_t)(lane >> 1)) {
                w.site = 0x10 + jt;
                mbar_wait(&pp.bar_empty[re.slot], re.par ^ 1u, w);
                if (wlane) {
                    mbar_expect_tx(&pp.bar_full[re.slot], wbytes + ACT_STAGE);
                    bulk_g2s(pp.stages + (size_t)re.slot * re.stage_bytes, Wbase + (size_t)s * wbytes, wbytes, &pp.bar_full[re.slot]);
                }
            }
            re.advance();
        }
    }
    if (gate >= 0) {            // the operands were written with generic stores by other SMs: acquire, then cross to the async proxy
        w.site = 0x20 + jt;
        tile_wait(P.gs, gate, nt, gate_steps, w);
        fence_proxy_async_global();
        if (pgate && lane == 0) *pgate = clock64();
    }
    for (int s = 0; s < nst; s++, pp.count++) {
        if (pp.count % npair == (uint32_t)(lane >> 1)) {
            uint8_t *dst = pp.stages + (size_t)r.slot * r.stage_bytes;
            const uint8_t *act = stage_src(P, stage_info(jt, s), nt, t);
            if (s >= early) {       // the whole stage: one instruction, two copies
                const long long c0 = pacc ? clock64() : 0;
                w.site = 0x30 + jt;
                mbar_wait(&pp.bar_empty[r.slot], r.par ^ 1u, w);
                if (pacc) waited += clock64() - c0;
                if (wlane) mbar_expect_tx(&pp.bar_full[r.slot], wbytes + ACT_STAGE);
                bulk_g2s(wlane ? dst : dst + r.w_slot, wlane ? Wbase + (size_t)s * wbytes : act, wlane ? wbytes : (uint32_t)ACT_STAGE,
                         &pp.bar_full[r.slot]);
            } else {                // armed, weights under way: the activation block, half per lane
                const uint32_t half = wlane ? 0u : (uint32_t)ACT_STAGE / 2;
                bulk_g2s(dst + r.w_slot + half, act + half, ACT_STAGE / 2, &pp.bar_full[r.slot]);
            }
        }
        r.advance();
    }
    if (pacc && lane == 0) *pacc = waited;
}

// One job of the MMA issuer thread: three MMAs (hi.hi, hi.lo, lo.hi) per k-step into the job's accumulator.
struct Issue {
    Ring r;
    uint32_t stage0, tb;
    uint64_t *bar_full, *bar_empty, *bar_tf;
    int exp;
};
__device__ __forceinline__ void issue_job(int jt, Issue &is, Watch &w, long long *macc = nullptr) {
    const int nst = ((is.exp & 1) && (jt == J_S1 || jt == J_S2)) ? 1 : job_stages(jt);
    const uint32_t rows = job_rows(jt);
    // A = activations (M = 128 folds), B = weights (N = rows); both K-major, no swizzle, [k-chunk][row][8].  Descriptors
    // differ only in the 14-bit start-address field (16-byte units): constant part once per job.
    const uint64_t dW0 = make_desc(0, rows * 16u, 128), dX0 = make_desc(0, FT * 16u, 128);
    const uint32_t stepW = 2u * rows, stepX = 2u * FT;          // two k-chunks, in 16-byte units
    const uint32_t loW = KC * rows, loX = KC * FT;               // hi -> lo image
    const uint32_t d = is.tb + (uint32_t)job_acc(jt);
    const uint32_t idesc = make_idesc(FT, (int)rows);
    const bool fresh = job_starts_acc(jt);
    Ring &r = is.r;
    long long a0 = 0, a1 = 0, a2 = 0;       // debug timeline only
    for (int s = 0; s < nst; s++) {
        const long long c0 = macc ? clock64() : 0;
        w.site = 0x40 + jt;
        mbar_wait(&is.bar_full[r.slot], r.par, w);
        tc_fence_after();
        const long long c1 = macc ? clock64() : 0;
        const uint32_t aW = (is.stage0 + r.slot * r.stage_bytes) >> 4, aX = aW + (r.w_slot >> 4);
#pragma unroll
        for (int ks = 0; ks < KC / 2; ks++) {
            const uint64_t wh = dW0 + (aW + ks * stepW), wl = dW0 + (aW + loW + ks * stepW);
            const uint64_t xh = dX0 + (aX + ks * stepX), xl = dX0 + (aX + loX + ks * stepX);
            umma_bf16(d, xh, wh, idesc, (fresh && s == 0 && ks == 0) ? 0u : 1u);
            umma_bf16(d, xh, wl, idesc, 1u);
            umma_bf16(d, xl, wh, idesc, 1u);
        }
        const long long c2 = macc ? clock64() : 0;
        umma_commit(&is.bar_empty[r.slot]);          // the stage is free once these MMAs have read it
        if (macc) { const long long c3 = clock64(); a0 += c1 - c0; a1 += c2 - c1; a2 += c3 - c2; }
        r.advance();
    }
    if (macc) { macc[0] = a0; macc[1] = a1; macc[2] = a2; }
    const int tf = job_commit(jt);
    if (tf >= 0) umma_commit(&is.bar_tf[tf]);        // this accumulator (and every MMA issued before) is complete
}

// fc3 contribution of an fc CTA: 128 folds x 32 logits = [128 x 64] (this CTA's fc2 outputs, written to shared memory by
// its epilogue) . [32 x 64]^T (its columns of fc3, resident), 4 k-steps x 3 MMAs into fc1's idle accumulator columns.
__device__ __forceinline__ void issue_p3(Issue &is, uint32_t p3a, uint32_t p3b, uint64_t *bar_y2, uint32_t par, Watch &w) {
    w.site = 0x4e;
    if (!w.dead && !mbar_test_wait(bar_y2, par)) {       // short wait on the chain: spin, do not sleep
        const long long t0 = clock64();
        unsigned spins = 0;
        while (!mbar_test_wait(bar_y2, par))
            if (give_up(w, t0, spins)) break;
    }
    tc_fence_after();
    const uint64_t dX0 = make_desc(0, FT * 16u, 128), dW0 = make_desc(0, NCP * 16u, 128);
    const uint32_t aX = p3a >> 4, aW = p3b >> 4, d = is.tb + (uint32_t)job_acc(J_F1C), idesc = make_idesc(FT, NCP);
#pragma unroll
    for (int ks = 0; ks < KC / 2; ks++) {
        const uint64_t xh = dX0 + (aX + ks * 2u * FT), xl = dX0 + (aX + KC * FT + ks * 2u * FT);
        const uint64_t wh = dW0 + (aW + ks * 2u * NCP), wl = dW0 + (aW + KC * NCP + ks * 2u * NCP);
        umma_bf16(d, xh, wh, idesc, ks == 0 ? 0u : 1u);
        umma_bf16(d, xh, wl, idesc, 1u);
        umma_bf16(d, xl, wh, idesc, 1u);
    }
    umma_commit(&is.bar_tf[TF_P3]);
}

// ---------------------------------------------------------------------------- epilogues (256 threads, warps 2..9)
// Accumulators are [128 folds (TMEM lanes)] x [weight rows (columns)], so a thread owns ONE fold (lane 32 q + lane of the
// warp's lane quarter q = warp % 4) and reads the r, z, n columns of a unit straight from its own lane: no shared-memory
// staging, no barrier inside an epilogue; the EPI_PARTS warps of a quarter split the columns (`part`).  Operand stores are one
// 16-byte chunk (8 units) per image per thread, consecutive lanes = consecutive folds = consecutive addresses.
//
// GRU gates (fatchord_version.py:210-216): X (input side: conditioning columns, issued ahead, + for rnn2 the h1' columns)
// and H (recurrent side, formed one step ahead) hold columns [r | z | n] x 32 units; the previous sample `xp` adds its
// column of the composed input matrix in fp32.  The thread keeps the state of its (fold, 16 units) in registers for the
// whole run (`hst`).  rnn2 also emits s12 = h1' + h2' (fc1 reads x + h1 + h2; x is folded into fc1's weights).
template <int LAYER>
__device__ __forceinline__ void epi_gru(const Params &P, int mt, int nt, int64_t t, uint32_t tb, const float *sb, float xp,
                                        float (&hst)[UPT], float (&h1new)[UPT], int q, int part, int lane, long long *pm = nullptr) {
    const int fi = q * 32 + lane, u0 = part * UPT, gu = mt * UT + u0;
    const size_t fold = (size_t)nt * FT + fi;
    // sb: input-side biases of r, z, n (the recurrent biases of r and z are folded into them on the host: the two sides
    // are only ever added), recurrent bias of n, previous-sample column
    const float *bi = sb + SB_G + LAYER * 2 * GRU_ROWS, *bhn = bi + GRU_ROWS + 2 * UT, *wx = sb + SB_XG + LAYER * GRU_ROWS;
    const uint32_t base = tb + ((uint32_t)(q * 32) << 16), colX = base + (uint32_t)job_acc(LAYER == 0 ? J_R1A : J_R2A) + u0,
                   colH = base + (uint32_t)job_acc(LAYER == 0 ? J_S1 : J_S2) + u0;
#pragma unroll
    for (int g = 0; g < UPT / 8; g++) {
        uint32_t xr[8], xz[8], xn[8], hr[8], hz[8], hn_[8];
        tmem_ld8_nowait(colX + 8 * g, xr);
        tmem_ld8_nowait(colX + UT + 8 * g, xz);
        tmem_ld8_nowait(colX + 2 * UT + 8 * g, xn);
        tmem_ld8_nowait(colH + 8 * g, hr);
        tmem_ld8_nowait(colH + UT + 8 * g, hz);
        tmem_ld8_nowait(colH + 2 * UT + 8 * g, hn_);
        tmem_ld_wait();
        if (pm && g == 0) pm[0] = clock64();
        float hn[8];
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const int u = u0 + 8 * g + i;
            hn[i] = gru_update_fast(fmaf(wx[u], xp, __uint_as_float(xr[i]) + bi[u]), fmaf(wx[UT + u], xp, __uint_as_float(xz[i]) + bi[UT + u]),
                                    fmaf(wx[2 * UT + u], xp, __uint_as_float(xn[i]) + bi[2 * UT + u]), __uint_as_float(hr[i]),
                                    __uint_as_float(hz[i]), __uint_as_float(hn_[i]) + bhn[u], hst[8 * g + i]);
            hst[8 * g + i] = hn[i];
        }
        if (P.dbg_state) {                  // fp32 copy of the state for wrnn_debug_mma_state only
            float4 *dst = reinterpret_cast<float4 *>((LAYER == 0 ? P.h1f : P.h2f) + fold * H + gu + 8 * g);
            dst[0] = make_float4(hn[0], hn[1], hn[2], hn[3]);
            dst[1] = make_float4(hn[4], hn[5], hn[6], hn[7]);
        }
        if (pm && g == 0) pm[1] = clock64();
        uint4 hi, lo;
        const int chunk = (gu >> 3) + g;
        if (LAYER == 1) {                   // what fc1 waits for goes out first; the new state itself follows (store_state)
            float sm[8];
#pragma unroll
            for (int i = 0; i < 8; i++) sm[i] = h1new[8 * g + i] + hn[i];
            split8(sm, hi, lo);
            *reinterpret_cast<uint4 *>(P.S12 + b_slot(FT, HGROUPS, nt, chunk, fi, 0)) = hi;
            *reinterpret_cast<uint4 *>(P.S12 + b_slot(FT, HGROUPS, nt, chunk, fi, 1)) = lo;
        } else {
            split8(hn, hi, lo);
            uint8_t *Hn = P.H1[(t + 1) & 1];
            *reinterpret_cast<uint4 *>(Hn + b_slot(FT, HGROUPS, nt, chunk, fi, 0)) = hi;
            *reinterpret_cast<uint4 *>(Hn + b_slot(FT, HGROUPS, nt, chunk, fi, 1)) = lo;
#pragma unroll
            for (int i = 0; i < 8; i++) h1new[8 * g + i] = hn[i];
        }
    }
}
// rnn2's new state as the operand of its recurrent products: off the chain, after h1' + h2' has been published
__device__ __forceinline__ void store_state2(const Params &P, int mt, int nt, int64_t t, const float (&hst)[UPT], int q, int part, int lane) {
    const int fi = q * 32 + lane, gu = mt * UT + part * UPT;
    uint8_t *Hn = P.H2[(t + 1) & 1];
#pragma unroll
    for (int g = 0; g < UPT / 8; g++) {
        float hn[8];
#pragma unroll
        for (int i = 0; i < 8; i++) hn[i] = hst[8 * g + i];
        uint4 hi, lo;
        split8(hn, hi, lo);
        *reinterpret_cast<uint4 *>(Hn + b_slot(FT, HGROUPS, nt, (gu >> 3) + g, fi, 0)) = hi;
        *reinterpret_cast<uint4 *>(Hn + b_slot(FT, HGROUPS, nt, (gu >> 3) + g, fi, 1)) = lo;
    }
}

// fc1 + ReLU (:217-218): 64 rows = 64 columns, RPT per thread, out as the bf16 hi/lo operand of fc2
__device__ __forceinline__ void epi_fc1(const Params &P, int mt, int nt, uint32_t tb, const float *sb, float xp, int q, int part, int lane) {
    const int fi = q * 32 + lane, r0 = part * RPT, gr = mt * FC_ROWS + r0;
    const float *bb = sb + SB_F + r0, *wx = sb + SB_XF + r0;
    const uint32_t col = tb + ((uint32_t)(q * 32) << 16) + (uint32_t)job_acc(J_F1C) + r0;
#pragma unroll
    for (int g = 0; g < RPT / 8; g += 2) {
        uint32_t a[8], b[8];
        tmem_ld8_nowait(col + 8 * g, a);
        tmem_ld8_nowait(col + 8 * g + 8, b);
        tmem_ld_wait();
        float va[8], vb[8];
#pragma unroll
        for (int i = 0; i < 8; i++) {
            va[i] = fmaxf(fmaf(wx[8 * g + i], xp, __uint_as_float(a[i]) + bb[8 * g + i]), 0.f);
            vb[i] = fmaxf(fmaf(wx[8 * g + 8 + i], xp, __uint_as_float(b[i]) + bb[8 * g + 8 + i]), 0.f);
        }
        uint4 hi, lo;
        split8(va, hi, lo);
        *reinterpret_cast<uint4 *>(P.Y1 + b_slot(FT, HGROUPS, nt, (gr >> 3) + g, fi, 0)) = hi;
        *reinterpret_cast<uint4 *>(P.Y1 + b_slot(FT, HGROUPS, nt, (gr >> 3) + g, fi, 1)) = lo;
        split8(vb, hi, lo);
        *reinterpret_cast<uint4 *>(P.Y1 + b_slot(FT, HGROUPS, nt, (gr >> 3) + g + 1, fi, 0)) = hi;
        *reinterpret_cast<uint4 *>(P.Y1 + b_slot(FT, HGROUPS, nt, (gr >> 3) + g + 1, fi, 1)) = lo;
    }
}

// fc2 + ReLU (:220-221), first half of the fused fc2 / fc3 epilogue.  The 64 outputs of a fold never leave the SM: they
// go to shared memory as the A operand (K = 64) of a 12-MMA job against this CTA's 64 columns of fc3 (:223), whose
// accumulator (128 folds x 32 logits) is this CTA's contribution to the fold's logits (second half: epi_fc3_part).  The
// sampler adds the eight contributions of a tile instead of running a sixth MMA phase on an operand that would have to be
// published and gathered first: one hand-over and 8 pipeline stages less per sample.
__device__ __forceinline__ void epi_fc2(const Params &P, int mt, int nt, uint32_t tb, const float *sb, uint8_t *p3a, int q, int part, int lane) {
    const int fi = q * 32 + lane, r0 = part * RPT, gr = mt * FC_ROWS + r0;
    const size_t fold = (size_t)nt * FT + fi;
    const float *bb = sb + SB_F + FC_ROWS + r0;
    const uint32_t col = tb + ((uint32_t)(q * 32) << 16) + (uint32_t)job_acc(J_F2C) + r0;
#pragma unroll
    for (int g = 0; g < RPT / 8; g += 2) {
        uint32_t a[8], b[8];
        tmem_ld8_nowait(col + 8 * g, a);
        tmem_ld8_nowait(col + 8 * g + 8, b);
        tmem_ld_wait();
        float va[8], vb[8];
#pragma unroll
        for (int i = 0; i < 8; i++) {
            va[i] = fmaxf(__uint_as_float(a[i]) + bb[8 * g + i], 0.f);
            vb[i] = fmaxf(__uint_as_float(b[i]) + bb[8 * g + 8 + i], 0.f);
        }
        if (P.dbg_state) {
            float4 *dst = reinterpret_cast<float4 *>(P.y2f + fold * FCD + gr + 8 * g);
            dst[0] = make_float4(va[0], va[1], va[2], va[3]);
            dst[1] = make_float4(va[4], va[5], va[6], va[7]);
            dst[2] = make_float4(vb[0], vb[1], vb[2], vb[3]);
            dst[3] = make_float4(vb[4], vb[5], vb[6], vb[7]);
        }
        // [hi|lo][k-chunk][fold][8]: consecutive lanes, consecutive 16-byte words
        const int chunk = (r0 >> 3) + g;
        uint4 hi, lo;
        split8(va, hi, lo);
        *reinterpret_cast<uint4 *>(p3a + (size_t)(chunk * FT + fi) * 16) = hi;
        *reinterpret_cast<uint4 *>(p3a + (size_t)((KC + chunk) * FT + fi) * 16) = lo;
        split8(vb, hi, lo);
        *reinterpret_cast<uint4 *>(p3a + (size_t)((chunk + 1) * FT + fi) * 16) = hi;
        *reinterpret_cast<uint4 *>(p3a + (size_t)((KC + chunk + 1) * FT + fi) * 16) = lo;
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // generic shared-memory stores, read by the tensor core
}
// second half: the contribution (128 folds x 32 logits) from TMEM to global, [tile][row tile][group of 4 logits][fold]:
// consecutive folds are consecutive 16-byte words on the writing and on the reading side
__device__ __forceinline__ void epi_fc3_part(const Params &P, int mt, int nt, uint32_t tb, int q, int part, int lane) {
    if (part != 0) return;
    const int fi = q * 32 + lane;
    uint32_t acc[32];
    tmem_ld32(tb + ((uint32_t)(q * 32) << 16) + (uint32_t)job_acc(J_F1C), acc);      // fc1's columns, free until the next step's F1A
    float4 *dst = reinterpret_cast<float4 *>(P.part3) + ((size_t)nt * MT_FC + mt) * (NCP / 4) * FT + fi;
#pragma unroll
    for (int i = 0; i < NCP / 4; i++)
        dst[(size_t)i * FT] = make_float4(__uint_as_float(acc[4 * i]), __uint_as_float(acc[4 * i + 1]), __uint_as_float(acc[4 * i + 2]),
                                          __uint_as_float(acc[4 * i + 3]));
}

// Noise of (step t, fold of this thread), injected or Philox, reduced to the terms the draw uses.  Off the chain: called
// before the wait for fc3's accumulator.
__device__ __forceinline__ void mol_noise_thread(const Params &P, int nt, int64_t t, int fi, float (&gum)[NMIX], float &lgt) {
    const int64_t b = (int64_t)nt * FT + fi;
    float u1[NMIX], u2;
    if (b >= P.a.B) {
#pragma unroll
        for (int i = 0; i < NMIX; i++) gum[i] = 0.f;
        lgt = 0.f;
        return;
    }
    if (P.a.u1) {
        const float *src = P.a.u1 + ((size_t)t * P.a.Btot + P.a.b0 + b) * NMIX;
#pragma unroll
        for (int i = 0; i < NMIX; i++) u1[i] = __ldg(src + i);
        u2 = __ldg(P.a.u2 + (size_t)t * P.a.Btot + P.a.b0 + b);
    } else {
        const uint2 key = make_uint2((uint32_t)P.a.seed, (uint32_t)(P.a.seed >> 32));
#pragma unroll
        for (int j = 0; j < 3; j++) {
            const uint4 r4 = philox4x32(make_uint4((uint32_t)t, (uint32_t)(P.a.b0 + b), (uint32_t)j, 1u), key);
            const uint32_t rr[4] = {r4.x, r4.y, r4.z, r4.w};
#pragma unroll
            for (int k = 0; k < 4; k++)
                if (4 * j + k < NMIX) u1[4 * j + k] = u_clip(rr[k]);
        }
        u2 = u_clip(philox4x32(make_uint4((uint32_t)t, (uint32_t)(P.a.b0 + b), 0u, 2u), key).x);
    }
#pragma unroll
    for (int i = 0; i < NMIX; i++) gum[i] = logf(-logf(u1[i]));
    lgt = logf(u2) - logf(1.f - u2);
}

// fc3 (:223) summed up + MoL draw (utils/distribution.py:87-123), once the tile has published fc2.  The fc CTAs of row
// tiles 0 .. 3 each sample a quarter (32 folds) of the tile, so that a CTA gathers 32 KB, not 128.  Lane = (contribution
// m = lane / 4, fold = 4 * warp + lane % 4): a thread loads the 8 x 16 bytes of one (fold, contribution), three
// xor-shuffles add the eight contributions (fixed order: deterministic), the lanes with m == 0 add the bias, draw, and
// write the sample, the mixture pick and the previous sample of step t + 1.
__device__ __forceinline__ int sampler_fold(int quarter, int ewarp, int lane) { return quarter * 32 + (ewarp & 7) * 4 + (lane & 3); }
__device__ __forceinline__ void epi_sample(const Params &P, int nt, int quarter, int64_t t, const float *sb, const float (&gum)[NMIX], float lgt,
                                           int ewarp, int lane, long long *pm = nullptr) {
    if (ewarp >= 8) return;                                  // warp-uniform (EPI_PARTS > 2: the first eight epilogue warps sample)
    const int m = lane >> 2, fi = sampler_fold(quarter, ewarp, lane);
    const int64_t b = (int64_t)nt * FT + fi, S = P.a.S;
    const float4 *src = reinterpret_cast<const float4 *>(P.part3) + ((size_t)nt * MT_FC + m) * (NCP / 4) * FT + fi;
    if (pm) pm[0] = clock64();
    float4 pv[NCP / 4];
#pragma unroll
    for (int i = 0; i < NCP / 4; i++) pv[i] = __ldcg(src + (size_t)i * FT);
    float lg[NCP];
#pragma unroll
    for (int i = 0; i < NCP / 4; i++) { lg[4 * i] = pv[i].x; lg[4 * i + 1] = pv[i].y; lg[4 * i + 2] = pv[i].z; lg[4 * i + 3] = pv[i].w; }
#pragma unroll
    for (int o = 4; o <= 16; o <<= 1)
#pragma unroll
        for (int i = 0; i < 3 * NMIX; i++) lg[i] += __shfl_xor_sync(0xffffffffu, lg[i], o);
    // Only the lanes with m == 0 draw.  No early return: lanes that left a divergent region early would run ahead to the
    // next CTA barrier, and a `bar.sync` reached by part of a warp left the rest of it unscheduled until the hardware's
    // forward-progress timer fired (~20 k clocks, seen as steps of 52 k / 73 k / 93 k clocks): reconverge explicitly.
    if (pm) pm[1] = clock64();
    if (m == 0) {
        uint2 *xnext = P.xp + b;
        if (b >= P.a.B) {
            *xnext = make_uint2(0u, (uint32_t)(t + 1));            // padded fold of the last tile
        } else {
#pragma unroll
            for (int i = 0; i < 3 * NMIX; i++) lg[i] += sb[SB_B3 + i];
            const bool live = t < P.a.fold_len[b];
            if (P.a.logits_out && live) {
                float *dst = P.a.logits_out + ((size_t)b * S + t) * P.NC;
#pragma unroll
                for (int i = 0; i < 3 * NMIX; i++) dst[i] = lg[i];
            }
            int pick = 0;
            float xv = mol_sample_thread(lg, gum, lgt, &pick);
            if (live) {
                P.a.samples[(size_t)b * S + t] = xv;
                P.a.picks[(size_t)b * S + t] = pick;
            } else {
                xv = 0.f;
            }
            if (P.a.teacher) xv = t + 1 < S ? P.a.teacher[(size_t)b * S + t + 1] : 0.f;
            *xnext = make_uint2(__float_as_uint(xv), (uint32_t)(t + 1));
        }
    }
    __syncwarp();
}

__global__ void __launch_bounds__(THREADS, 1) mma_loop_kernel(const Params P) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t bar_full[MAXSTAGE], bar_empty[MAXSTAGE], bar_tf[NTF], bar_done, bar_y2;
    __shared__ uint32_t tmem_base_s;
    __shared__ int dead_s;
    uint8_t *stages = smem;
    // (roles first: the layout depends on them)

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, cta = blockIdx.x;
    const int G = P.G;
    Watch w{P.gs.abort_flag, false, 0};
    // static roles: GRU (row tile, fold tile) jobs fill CTAs from 0 up, fc jobs from G - 1 down; at most one of each per CTA;
    // the fc CTA of row tile 0 also runs the tile's fc3 job and its sampler
    const bool has_gru = cta < MT_GRU * P.NTILES, has_fc = G - 1 - cta < MT_FC * P.NTILES;
    const int mt_g = cta % MT_GRU, nt_g = cta / MT_GRU, mt_f = (G - 1 - cta) % MT_FC, nt_f = (G - 1 - cta) / MT_FC;
    const bool has_f3 = has_fc && mt_f < 4;
    uint8_t *p3a = smem + ring_bytes_of(P, has_fc), *p3b = p3a + P3_A;          // fc3-inside-fc operands (fc CTAs only)
    float *sbias = reinterpret_cast<float *>(has_fc ? p3b + P3_B : p3a);       // this CTA's bias rows / previous-sample columns (SB_*)          // row tiles 0 .. 3 of the fc CTAs each sample a quarter of the tile's folds
    if (!has_gru && !has_fc) return;

    if (tid == 0) {
        for (int i = 0; i < MAXSTAGE; i++) { mbar_init(&bar_full[i], 1); mbar_init(&bar_empty[i], 1); }
        for (int i = 0; i < NTF; i++) mbar_init(&bar_tf[i], 1);
        mbar_init(&bar_done, 1);
        mbar_init(&bar_y2, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) tmem_alloc(&tmem_base_s, TMEM_COLS);
    for (int i = tid; i < (has_fc ? SBIAS : SBIAS_GRU); i += THREADS) {
        const float *gb = P.bias;                                               // rnn1 [16][2][96] | rnn2 [16][2][96]
        const float *fb = gb + (size_t)2 * MT_GRU * 2 * GRU_ROWS;               // fc1 [512] | fc2 [512]
        const float *xg = fb + 2 * FCD;                                         // x columns rnn1 [16][96] | rnn2 [16][96]
        const float *xf = xg + (size_t)2 * MT_GRU * GRU_ROWS;                   // x column fc1 [512]
        const float *b3 = xf + FCD;                                             // fc3 bias [32]
        float v = 0.f;
        if (i < SB_XG) {
            if (has_gru) v = gb[(size_t)((i / (2 * GRU_ROWS)) * MT_GRU + mt_g) * 2 * GRU_ROWS + i % (2 * GRU_ROWS)];
        } else if (i < SBIAS_GRU) {
            const int j = i - SB_XG;
            if (has_gru) v = xg[(size_t)((j / GRU_ROWS) * MT_GRU + mt_g) * GRU_ROWS + j % GRU_ROWS];
        } else if (i < SB_XF) {
            const int j = i - SB_F;
            v = fb[(size_t)(j / FC_ROWS) * FCD + mt_f * FC_ROWS + j % FC_ROWS];
        } else if (i < SB_B3) {
            v = xf[mt_f * FC_ROWS + (i - SB_XF)];
        } else {
            v = b3[i - SB_B3];
        }
        sbias[i] = v;
    }
    if (has_fc) {       // this CTA's 64 columns of fc3 as an MMA operand; read by the tensor core (async proxy) from step 0 on
        const uint4 *src = reinterpret_cast<const uint4 *>(P.Aimg + P.offP3 + (size_t)mt_f * P3_B);
        for (int i = tid; i < P3_B / 16; i += THREADS) reinterpret_cast<uint4 *>(p3b)[i] = src[i];
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tb = tmem_base_s;
    const bool profiled_cta = P.prof && (cta == 0 || cta == G - 1);
    long long *prof_base = P.prof + (cta == 0 ? 0 : 64);

    if (warp == 0) {
        // ======================================================================= TMA producer
        if (lane < NPROD) {
            Pipe pp{make_ring(P, has_gru, has_fc), 0u, stages, bar_full, bar_empty};
            // step "-1": recurrent products of step 0 on the zero state (h = 0, :194-195) and step 0's conditioning columns
            if (has_gru) {
                produce_job(P, J_S1, mt_g, nt_g, -1, -1, 0, pp, lane, w);
                produce_job(P, J_R1A, mt_g, nt_g, 0, -1, 0, pp, lane, w);
                produce_job(P, J_S2, mt_g, nt_g, -1, -1, 0, pp, lane, w);
                produce_job(P, J_R2A, mt_g, nt_g, 0, -1, 0, pp, lane, w);
            }
            if (has_fc) {
                produce_job(P, J_F1A, mt_f, nt_f, 0, -1, 0, pp, lane, w);
                produce_job(P, J_F2A, mt_f, nt_f, 0, -1, 0, pp, lane, w);
            }
            for (int64_t t = 0; t < P.steps_run && !w.dead; t++) {
                long long *pa = (profiled_cta && t == P.prof_step) ? prof_base + 32 : nullptr;
                const bool more = t + 1 < P.steps_run;
                if (has_gru) {
                    PROF_MARK(0);
                    produce_job(P, J_R2C, mt_g, nt_g, t, PH_RNN1, t + 1, pp, lane, w, pa ? pa + 0 : nullptr, pa ? prof_base + 8 : nullptr);
                    PROF_MARK(1);
                    if (more) produce_job(P, J_S1, mt_g, nt_g, t, -1, 0, pp, lane, w, pa ? pa + 1 : nullptr);
                    if (more) produce_job(P, J_R1A, mt_g, nt_g, t + 1, -1, 0, pp, lane, w);
                    PROF_MARK(2);
                }
                if (has_fc) {
                    PROF_MARK(3);
                    produce_job(P, J_F1C, mt_f, nt_f, t, PH_RNN2, t + 1, pp, lane, w, pa ? pa + 2 : nullptr, pa ? prof_base + 9 : nullptr);
                    PROF_MARK(4);
                    produce_job(P, J_F2C, mt_f, nt_f, t, PH_FC1, t + 1, pp, lane, w, pa ? pa + 3 : nullptr, pa ? prof_base + 19 : nullptr);
                    PROF_MARK(5);
                    if (more) {
                        // fc1's accumulator columns hold the fc3 contribution until this CTA's own epilogue has stored it, fc2's
                        // until it has been read: both implied by the tile's fc2 counter
                        produce_job(P, J_F1A, mt_f, nt_f, t + 1, PH_FC2, t + 1, pp, lane, w);
                        produce_job(P, J_F2A, mt_f, nt_f, t + 1, -1, 0, pp, lane, w);
                    }
                    PROF_MARK(6);
                }
                if (has_gru && more) {         // nothing reads the recurrent products of the step after the last
                    produce_job(P, J_S2, mt_g, nt_g, t, PH_H2, t + 1, pp, lane, w, pa ? pa + 5 : nullptr);
                    produce_job(P, J_R2A, mt_g, nt_g, t + 1, -1, 0, pp, lane, w);
                    PROF_MARK(7);
                }
            }
        }
    } else if (warp == 1) {
        // ======================================================================= MMA issuer: the producer's job order, no gates
        if (lane == 0) {
            Issue is{make_ring(P, has_gru, has_fc), smem_u32(stages), tb, bar_full, bar_empty, bar_tf, P.exp};
            if (has_gru) {
                issue_job(J_S1, is, w);
                issue_job(J_R1A, is, w);
                issue_job(J_S2, is, w);
                issue_job(J_R2A, is, w);
            }
            if (has_fc) {
                issue_job(J_F1A, is, w);
                issue_job(J_F2A, is, w);
            }
            for (int64_t t = 0; t < P.steps_run && !w.dead; t++) {
                long long *ma = (profiled_cta && t == P.prof_step) ? prof_base + 40 : nullptr;
                const bool more = t + 1 < P.steps_run;
                if (has_gru) {
                    issue_job(J_R2C, is, w, ma ? ma + 0 : nullptr);
                    PROF_MARK(10);
                    if (more) issue_job(J_S1, is, w, ma ? ma + 3 : nullptr);
                    if (more) issue_job(J_R1A, is, w);
                    PROF_MARK(11);
                }
                if (has_fc) {
                    issue_job(J_F1C, is, w, ma ? ma + 6 : nullptr);
                    PROF_MARK(12);
                    issue_job(J_F2C, is, w, ma ? ma + 9 : nullptr);
                    PROF_MARK(13);
                    issue_p3(is, smem_u32(p3a), smem_u32(p3b), &bar_y2, (uint32_t)(t & 1), w);
                    PROF_MARK(14);
                    if (more) {
                        issue_job(J_F1A, is, w);
                        issue_job(J_F2A, is, w);
                    }
                }
                if (has_gru && more) {
                    issue_job(J_S2, is, w, ma ? ma + 15 : nullptr);
                    issue_job(J_R2A, is, w);
                    PROF_MARK(15);
                }
            }
            // every MMA issued (hence every copy it consumed) is through before the CTA gives up TMEM and shared memory
            umma_commit(&bar_done);
            w.site = 0x4f;
            mbar_wait(&bar_done, 0u, w);
        }
        __syncwarp();
    } else {
        // ======================================================================= epilogue warps
        const int etid = tid - 64, ewarp = warp - 2;
        const int q = warp & 3, part = ewarp >> 2;          // TMEM lane quarter this warp may read; which part of the columns
        const int fi = q * 32 + lane;
        uint32_t tfpar = 0;                                  // every accumulator-ready barrier completes once per step
        // recurrent state of this thread's (fold, 16 units), both layers, and rnn1's new state for s12: registers for the whole run
        float hst1[UPT], hst2[UPT], h1new[UPT];
#pragma unroll
        for (int i = 0; i < UPT; i++) hst1[i] = hst2[i] = h1new[i] = 0.f;        // h1 = h2 = 0 (:194-195)
        auto publish = [&](int phase, int nt) {              // every writing thread fences, one thread releases the counter
            fence_proxy_async_global();
            epi_bar();
            if (etid == 0) red_release_add_u64(tile_counter(P.gs, phase, nt), 1ull);
            // Close the one-lane region here, and keep one-lane regions short everywhere in the epilogue (the sampler's poll
            // of the tile counter is done by a whole warp).  With a long one-lane region after this point (lane 0 polling
            // while lanes 1..31 went on to the next barrier), the first epilogue warp stayed split in two fragments through
            // the sampler's barrier, loads and shuffles, and the fragments took turns on the hardware's forward-progress
            // timer: steps of 52 k / 73 k / 93 k clocks instead of 52 k (36 us instead of 28.5 us per step), in some builds
            // and not in others: a compiler barrier or an unrelated branch in between was enough to hide it, a 20 ns sleep
            // of lane 0 made it 40 us.
            __syncwarp();
        };
        auto wait_acc = [&](int tf) {
            w.site = 0x50 + tf;
            mbar_wait(&bar_tf[tf], tfpar, w);
            __syncwarp();       // lanes leave the spin loop one by one; tcgen05.ld is .sync.aligned: converge first
            tc_fence_after();
        };
        for (int64_t t = 0; t < P.steps_run; t++) {
            if (w.dead) break;                               // uniform: set below from a flag read after a barrier
            if (has_gru) {
                wait_acc(TF_R1);
                if (etid == 0) PROF_MARK(20);
                w.site = 0x60;
                const float xp = wait_prev_sample(P.xp + (size_t)nt_g * FT + fi, t, w);      // the sample of step t - 1 is out
                if (etid == 0) PROF_MARK(21);
                if (etid == 0 && cta == 0) STEP_MARK(5);
                __syncwarp();
                epi_gru<0>(P, mt_g, nt_g, t, tb, sbias, xp, hst1, h1new, q, part, lane, (profiled_cta && t == P.prof_step && etid == 0) ? prof_base + 58 : nullptr);
                tc_fence_before();
                if (etid == 0) PROF_MARK(22);
                publish(PH_RNN1, nt_g);
                if (etid == 0) PROF_MARK(23);
                if (etid == 0 && cta == 0) STEP_MARK(6);
                wait_acc(TF_R2);
                if (etid == 0) PROF_MARK(24);
                if (etid == 0 && cta == 0) STEP_MARK(7);
                epi_gru<1>(P, mt_g, nt_g, t, tb, sbias, xp, hst2, h1new, q, part, lane);
                tc_fence_before();
                if (etid == 0) PROF_MARK(25);
                publish(PH_RNN2, nt_g);
                if (etid == 0) PROF_MARK(26);
                if (etid == 0 && cta == 0) STEP_MARK(8);
                store_state2(P, mt_g, nt_g, t, hst2, q, part, lane);
                publish(PH_H2, nt_g);
            }
            if (has_fc) {
                wait_acc(TF_F1);
                if (etid == 0) PROF_MARK(27);
                if (etid == 0 && cta == G - 1) STEP_MARK(0);
                // published before rnn1 of this step began; the loads that filled fc1's stages were ordered behind rnn2's counter
                const float xp = __uint_as_float(ld_volatile_u2(P.xp + (size_t)nt_f * FT + fi).x);
                epi_fc1(P, mt_f, nt_f, tb, sbias, xp, q, part, lane);
                tc_fence_before();
                publish(PH_FC1, nt_f);
                if (etid == 0) PROF_MARK(28);
                float gum[NMIX], lgt = 0.f;
                if (has_f3 && ewarp < 8 && (lane >> 2) == 0) mol_noise_thread(P, nt_f, t, sampler_fold(mt_f, ewarp, lane), gum, lgt);      // off the chain
                __syncwarp();
                wait_acc(TF_F2);
                if (etid == 0) PROF_MARK(29);
                if (etid == 0 && cta == G - 1) STEP_MARK(1);
                epi_fc2(P, mt_f, nt_f, tb, sbias, p3a, q, part, lane);
                tc_fence_before();
                epi_bar();
                if (etid == 0) mbar_arrive(&bar_y2);                  // the MMA thread forms this CTA's fc3 contribution
                wait_acc(TF_P3);
                epi_fc3_part(P, mt_f, nt_f, tb, q, part, lane);
                tc_fence_before();
                publish(PH_FC2, nt_f);
                if (etid == 0) PROF_MARK(30);
                if (etid == 0 && cta == G - 1) STEP_MARK(2);
                __syncwarp();                                            // see publish(): no one-lane region may reach the sampler
                if (has_f3) {
                    if (ewarp == 0) {                                   // the whole warp polls (one request): no one-lane region
                        w.site = 0x61;
                        tile_wait(P.gs, PH_FC2, nt_f, t + 1, w);       // the eight contributions to fc3 are out
                        if (lane == 0) PROF_MARK(16);
                        if (lane == 0 && cta == G - 1) STEP_MARK(3);
                        __syncwarp();
                    }
                    epi_bar();
                    epi_sample(P, nt_f, mt_f, t, sbias, gum, lgt, ewarp, lane, (profiled_cta && t == P.prof_step && etid == 0) ? prof_base + 62 : nullptr);
                    if (etid == 0) PROF_MARK(17);
                    if (etid == 0 && profiled_cta && t == P.prof_step + 100) prof_base[61] = clock64();      // 100 steps later: the period
                    if (etid == 0 && P.prof && cta == G - 1 && t >= P.prof_step - 8 && t < P.prof_step + 56) P.prof[128 + (t - P.prof_step + 8)] = clock64();
                    if (etid == 0 && cta == G - 1) STEP_MARK(4);
                }
            }
            tfpar ^= 1u;
            // uniform exit on a fired watchdog
            if (etid == 0) dead_s = (w.dead || ld_volatile_s32(P.gs.abort_flag)) ? 1 : 0;
            epi_bar();
            w.dead = dead_s != 0;
            epi_bar();
        }
    }
    // after an abort the waits fall through: give copies and MMAs still in flight time to land before shared memory and
    // TMEM go away, so that the host can read the abort code instead of a launch failure
    if (w.dead || ld_volatile_s32(P.gs.abort_flag)) {
        const long long t0 = clock64();
        while (clock64() - t0 < 4000000ll) {}
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tb, TMEM_COLS);
}

}  // namespace mma

// ============================================================================ host: pack + launch
static inline uint16_t f2bf(float x) {       // round to nearest even
    uint32_t u;
    memcpy(&u, &x, 4);
    if ((u & 0x7f800000u) == 0x7f800000u) return (uint16_t)(u >> 16);
    u += 0x7fffu + ((u >> 16) & 1u);
    return (uint16_t)(u >> 16);
}
static inline float bf2f(uint16_t b) {
    uint32_t u = (uint32_t)b << 16;
    float x;
    memcpy(&x, &u, 4);
    return x;
}

bool mma_supported(const wrnn_handle *h) {
    const Dims &d = h->d;
    return h->mma.ok && d.mode == WRNN_MODE_MOL && d.H == mma::H && d.FC == mma::FCD && d.A == mma::AUX && d.F == mma::FEAT &&
           d.NC == 3 * mma::NMIX;
}
int mma_min_folds() { return 41; }      // measured (scripts/mma_crossover.py): 27.4 us per step up to 120 folds; two resident groups of 20 take 23.8, three 35.7

int mma_plan(wrnn_handle *h) {
    using namespace mma;
    const Dims &d = h->d;
    MmaPlan &pl = h->mma;
    pl.ok = false;
    if (!(d.mode == WRNN_MODE_MOL && d.H == H && d.FC == FCD && d.A == AUX && d.F == FEAT && d.NC == 3 * NMIX)) return 0;
    if (h->num_sms < 16) return 0;
    auto get = [&](const char *k) -> const HostTensor * {
        auto it = h->w.find(k);
        return it == h->w.end() ? nullptr : &it->second;
    };
    const HostTensor *WI = get("I.weight"), *BI = get("I.bias");
    const HostTensor *W1i = get("rnn1.weight_ih_l0"), *W1h = get("rnn1.weight_hh_l0"), *B1i = get("rnn1.bias_ih_l0"),
                     *B1h = get("rnn1.bias_hh_l0");
    const HostTensor *W2i = get("rnn2.weight_ih_l0"), *W2h = get("rnn2.weight_hh_l0"), *B2i = get("rnn2.bias_ih_l0"),
                     *B2h = get("rnn2.bias_hh_l0");
    const HostTensor *F1 = get("fc1.weight"), *F1b = get("fc1.bias"), *F2 = get("fc2.weight"), *F2b = get("fc2.bias"),
                     *F3 = get("fc3.weight"), *F3b = get("fc3.bias");
    if (!WI || !BI || !W1i || !W1h || !B1i || !B1h || !W2i || !W2h || !B2i || !B2h || !F1 || !F1b || !F2 || !F2b || !F3 || !F3b)
        return 0;      // fma_plan reports missing keys
    const int KI = 1 + FEAT + AUX;     // 113: [x_prev, m, a1]
    // composed matrices in double (fatchord_version.py:208-218): W . W_I and W . b_I for the three consumers of x
    auto compose = [&](const HostTensor *W, int rows, int ldw, std::vector<double> &Wc, std::vector<double> &bc) {
        Wc.assign((size_t)rows * KI, 0.0);
        bc.assign(rows, 0.0);
        for (int r = 0; r < rows; r++) {
            const float *wr = W->v.data() + (size_t)r * ldw;
            double *o = Wc.data() + (size_t)r * KI;
            double bsum = 0.0;
            for (int k = 0; k < H; k++) {
                const double wk = wr[k];
                const float *ir = WI->v.data() + (size_t)k * KI;
                for (int c = 0; c < KI; c++) o[c] += wk * ir[c];
                bsum += wk * BI->v[k];
            }
            bc[r] = bsum;
        }
    };
    std::vector<double> Wc1, bc1, Wc2, bc2, Wc3, bc3;
    compose(W1i, 3 * H, H, Wc1, bc1);
    compose(W2i, 3 * H, H + AUX, Wc2, bc2);
    compose(F1, FCD, H + AUX, Wc3, bc3);
    // value of weight element (phase p, global row, conditioning element e of the operand order) / (512-wide operand, k)
    auto cond_w = [&](int p, int grow, int e) -> double {
        const bool in_x = e >= C_M && e < C_A3;        // m, a1: columns 1 .. 112 of the composed matrix (column 0 = x_prev: epilogues)
        const int xc = 1 + (e - C_M);
        if (p == 0) return in_x ? Wc1[(size_t)grow * KI + xc] : 0.0;
        if (p == 1) {
            if (e >= C_A2 && e < C_M) return W2i->v[(size_t)grow * (H + AUX) + H + (e - C_A2)];
            return in_x ? Wc2[(size_t)grow * KI + xc] : 0.0;
        }
        if (p == 2) {
            if (e >= C_A3 && e < C_END) return F1->v[(size_t)grow * (H + AUX) + H + (e - C_A3)];
            return in_x ? Wc3[(size_t)grow * KI + xc] : 0.0;
        }
        return (e >= C_A4 && e < C_A2) ? F2->v[(size_t)grow * (FCD + AUX) + FCD + (e - C_A4)] : 0.0;
    };
    // weight on a 512-wide operand column k for job type jt
    auto main_w = [&](int jt, int grow, int k) -> double {
        switch (jt) {
            case J_R2C: return W2i->v[(size_t)grow * (H + AUX) + k];
            case J_S1: return W1h->v[(size_t)grow * H + k];
            case J_S2: return W2h->v[(size_t)grow * H + k];
            case J_F1C: return F1->v[(size_t)grow * (H + AUX) + k];
            default: return F2->v[(size_t)grow * (FCD + AUX) + k];
        }
    };
    auto cond_layer = [](int jt) { return jt == J_R1A ? 0 : jt == J_R2A ? 1 : jt == J_F1A ? 2 : 3; };     // J_F2A: 3
    size_t total = 0;
    for (int jt = 0; jt < NJOB; jt++) {
        pl.offA[jt] = (int64_t)total;
        total += (size_t)job_tiles(jt) * job_stages(jt) * 2 * KC * job_rows(jt) * 16;
    }
    pl.offP3 = (int64_t)total;
    total += (size_t)MT_FC * P3_B;
    std::vector<uint16_t> img(total / 2, 0);
    for (int mt = 0; mt < MT_FC; mt++) {      // fc3 columns of fc2 row tile mt: [hi|lo][k-chunk][32 rows][8], rows 30, 31 zero
        uint16_t *blk = img.data() + (size_t)pl.offP3 / 2 + (size_t)mt * (P3_B / 2);
        for (int r = 0; r < d.NC; r++)
            for (int k = 0; k < FC_ROWS; k++) {
                const float wf = F3->v[(size_t)r * FCD + mt * FC_ROWS + k];
                const uint16_t hi = f2bf(wf), lo = f2bf(wf - bf2f(hi));
                const size_t o = ((size_t)(k / 8) * NCP + r) * 8 + k % 8;
                blk[o] = hi;
                blk[(size_t)KC * NCP * 8 + o] = lo;
            }
    }
    for (int jt = 0; jt < NJOB; jt++) {
        const int rows = job_rows(jt), mts = job_tiles(jt), nst = job_stages(jt);
        for (int mt = 0; mt < mts; mt++)
            for (int s = 0; s < nst; s++) {
                const StageInfo si = stage_info(jt, s);
                uint16_t *blk = img.data() + (size_t)pl.offA[jt] / 2 + ((size_t)mt * nst + s) * (2 * KC * rows * 8);
                for (int r = 0; r < rows; r++) {
                    const int grow = rows == GRU_ROWS ? (r / UT) * H + mt * UT + (r % UT) : mt * FC_ROWS + r;
                    for (int kl = 0; kl < KC * 8; kl++) {
                        const int feat = si.g * (KC * 8) + kl;
                        const double wv = si.arr == ARR_C ? cond_w(cond_layer(jt), grow, feat) : main_w(jt, grow, feat);
                        const float wf = (float)wv;
                        const uint16_t hi = f2bf(wf), lo = f2bf((float)(wv - (double)bf2f(hi)));
                        const size_t o = ((size_t)(kl / 8) * rows + r) * 8 + kl % 8;
                        blk[o] = hi;
                        blk[(size_t)KC * rows * 8 + o] = lo;
                    }
                }
            }
    }
    // biases, then the previous-sample column of the three composed input matrices (column 0 of W . W_I), then fc3's bias
    const size_t off_fc = (size_t)2 * MT_GRU * 2 * GRU_ROWS, off_xg = off_fc + 2 * FCD, off_xf = off_xg + (size_t)2 * MT_GRU * GRU_ROWS,
                 off_b3 = off_xf + FCD;
    std::vector<float> bias(off_b3 + NCP, 0.f);
    for (int p = 0; p < 2; p++)
        for (int mt = 0; mt < MT_GRU; mt++)
            for (int r = 0; r < GRU_ROWS; r++) {
                const int grow = (r / UT) * H + mt * UT + (r % UT);
                float *b = bias.data() + (size_t)(p * MT_GRU + mt) * 2 * GRU_ROWS;
                const double bhh = (p == 0 ? B1h : B2h)->v[grow];
                const bool n_gate = r >= 2 * UT;      // r and z: both sides are only ever added, one bias; n keeps its recurrent bias apart
                b[r] = (float)((p == 0 ? bc1[grow] : bc2[grow]) + (double)(p == 0 ? B1i : B2i)->v[grow] + (n_gate ? 0.0 : bhh));
                b[GRU_ROWS + r] = n_gate ? (float)bhh : 0.f;
                bias[off_xg + (size_t)(p * MT_GRU + mt) * GRU_ROWS + r] = (float)(p == 0 ? Wc1 : Wc2)[(size_t)grow * KI];
            }
    for (int r = 0; r < FCD; r++) {
        bias[off_fc + r] = (float)(bc3[r] + (double)F1b->v[r]);
        bias[off_fc + FCD + r] = F2b->v[r];
        bias[off_xf + r] = (float)Wc3[(size_t)r * KI];
    }
    for (int r = 0; r < d.NC; r++) bias[off_b3 + r] = F3b->v[r];
    if (pl.Aimg) cudaFree(pl.Aimg);
    if (pl.bias) cudaFree(pl.bias);
    pl.Aimg = nullptr; pl.bias = nullptr;
    WRNN_CUDA(h, cudaMalloc((void **)&pl.Aimg, img.size() * 2));
    WRNN_CUDA(h, cudaMemcpy(pl.Aimg, img.data(), img.size() * 2, cudaMemcpyHostToDevice));
    WRNN_CUDA(h, cudaMalloc((void **)&pl.bias, bias.size() * 4));
    WRNN_CUDA(h, cudaMemcpy(pl.bias, bias.data(), bias.size() * 4, cudaMemcpyHostToDevice));
    pl.img_bytes = total;
    // everything the device offers minus the static part (barriers): the stage ring takes what the bias rows leave
    pl.smem_bytes = (h->smem_optin - 512) / 128 * 128;       // static shared memory: 27 mbarriers and two words
    if (pl.smem_bytes < (size_t)FIXED_SMEM + 3 * (W_GRU + ACT_STAGE)) return 0;
    WRNN_CUDA(h, cudaFuncSetAttribute(mma::mma_loop_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem_bytes));
    pl.ok = true;
    return 0;
}

// The fold tile is the M of the MMA: 128 folds = the 128 TMEM lanes, fixed.  16 GRU jobs + 4 fc jobs per fold tile; up to
// num_sms / 16 tiles (9 on a B200: 1152 folds) so that every CTA has at most one GRU job; fc jobs fill CTAs from the top.
int64_t mma_max_folds(const wrnn_handle *h) { return (int64_t)128 * (h->num_sms / mma::MT_GRU); }

// initial previous sample: 0, or the teacher's first sample (fatchord_version.py:197, forward(): x[:, 0])
__global__ void xp_init_kernel(uint2 *xp, const float *teacher, int64_t B, int64_t S, int64_t NF) {
    const int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (b < NF) xp[b] = make_uint2(__float_as_uint((b < B && teacher) ? teacher[(size_t)b * S] : 0.f), 0u);
}

int mma_run(wrnn_handle *h, const LoopArgs &a, cudaStream_t st) {
    using namespace mma;
    const MmaPlan &pl = h->mma;
    Params P;
    P.a = a;
    P.G = h->num_sms;
    P.NTILES = (int)((a.B + FT - 1) / FT);
    if (MT_GRU * P.NTILES > P.G || P.NTILES > MAXTILES)
        return fail(h, WRNN_ERR_ARG, "tensor-core loop kernel: %lld folds in tiles of %d need %d CTAs, the device has %d",
                    (long long)a.B, FT, MT_GRU * P.NTILES, P.G);
    P.smem_total = (int)pl.smem_bytes;
    P.NC = h->d.NC;
    P.steps_run = a.S;
    P.dbg_state = 0;
    if (const char *e = getenv("WRNN_MMA_STEPS")) {
        P.steps_run = std::min<int64_t>(a.S, std::max(1, atoi(e)));
        P.dbg_state = 1;
    }
    P.Aimg = pl.Aimg;
    for (int j = 0; j < NJOB; j++) P.offA[j] = pl.offA[j];
    P.offP3 = pl.offP3;
    P.bias = pl.bias;
    const size_t NF = (size_t)P.NTILES * FT;
    const size_t bH = NF * H * 4, bF = NF * H * sizeof(float);       // operands: hi + lo bf16 = 4 B per element
    const size_t bX = (NF * sizeof(uint2) + 1023) / 1024 * 1024;
    const size_t bC = (size_t)a.S * P.NTILES * CGROUPS * ACT_STAGE;  // conditioning operand of every step
    const size_t bP = (size_t)P.NTILES * MT_FC * FT * NCP * sizeof(float);
    const size_t state = 6 * bH + bX + 3 * bF + bP;
    WRNN_TRY(ws_reserve(h, h->ws_act, state + bC + 4096));
    WRNN_CUDA(h, cudaMemsetAsync(h->ws_act.p, 0, state, st));       // h1 = h2 = 0 (:194-195), padded folds stay zero
    uint8_t *p8 = (uint8_t *)h->ws_act.p;
    P.H1[0] = p8; p8 += bH; P.H1[1] = p8; p8 += bH;
    P.H2[0] = p8; p8 += bH; P.H2[1] = p8; p8 += bH;
    P.S12 = p8; p8 += bH; P.Y1 = p8; p8 += bH;
    P.xp = (uint2 *)p8; p8 += bX;
    P.h1f = (float *)p8; p8 += bF; P.h2f = (float *)p8; p8 += bF; P.y2f = (float *)p8; p8 += bF;
    P.part3 = (float *)p8; p8 += bP;
    uint8_t *Cop = p8;
    P.Cop = Cop;
    h->mma_dbg[0] = P.h1f; h->mma_dbg[1] = P.h2f; h->mma_dbg[2] = P.y2f;
    {       // pre-pass: the conditioning operand of all steps, and the first previous sample
        if (!a.melT || !a.res_out || !a.fold_meta) return fail(h, WRNN_ERR_STATE, "tensor-core loop kernel: conditioning tensors not staged");
        const CondSrc cs{a.melT, a.res_out, h->cp.poly, (const FoldMeta *)a.fold_meta, h->d.total_scale, (int)h->d.indent, h->cp.poly_taps,
                         h->cp.poly_kmin};
        const int64_t blocks = a.S * (int64_t)P.NTILES * 16;
        if (blocks >= 0x7fffffff) return fail(h, WRNN_ERR_ARG, "tensor-core loop kernel: too many steps x fold tiles for one conditioning launch");
        cond_operand_kernel<<<(unsigned)blocks, 256, 0, st>>>(cs, a.B, a.S, P.NTILES, Cop);
        xp_init_kernel<<<(unsigned)((NF + 255) / 256), 256, 0, st>>>(P.xp, a.teacher, a.B, a.S, (int64_t)NF);
        WRNN_CUDA(h, cudaGetLastError());
        h->launches += 2;
    }
    const size_t sync_bytes = (size_t)(1 + NPHASE * MAXTILES) * 128;  // one 128-byte line per (phase, fold tile) counter, from +128
    WRNN_TRY(ws_reserve(h, h->ws_mma_sync, sync_bytes));
    WRNN_CUDA(h, cudaMemsetAsync(h->ws_mma_sync.p, 0, sync_bytes, st));
    P.gs.counter = (unsigned long long *)h->ws_mma_sync.p;
    P.gs.abort_flag = (int *)((char *)h->ws_sync.p + 64);             // the call's sticky abort flag (api_run.cu)
    WRNN_CUDA(h, cudaFuncSetAttribute(mma_loop_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem_bytes));
    P.prof = nullptr;
    P.prof_step = -1;
    P.exp = getenv("WRNN_MMA_EXP") ? atoi(getenv("WRNN_MMA_EXP")) : 0;
    const char *pe = getenv("WRNN_MMA_PROF");
    if (pe) {
        WRNN_TRY(ws_reserve(h, h->ws_prof, 832 * sizeof(long long)));
        WRNN_CUDA(h, cudaMemsetAsync(h->ws_prof.p, 0, 832 * sizeof(long long), st));
        P.prof = (long long *)h->ws_prof.p;
        P.prof_step = std::min<int64_t>(atoi(pe), P.steps_run - 1);
    }
    void *args[] = {(void *)&P};
    WRNN_CUDA(h, cudaLaunchCooperativeKernel((const void *)mma_loop_kernel, dim3(P.G), dim3(THREADS), args, pl.smem_bytes, st));
    h->launches++;
    if (pe) {     // debug: print the timeline of the profiled step (clocks relative to the CTA's first event of the step)
        std::vector<long long> pr(832);
        WRNN_CUDA(h, cudaStreamSynchronize(st));
        WRNN_CUDA(h, cudaMemcpy(pr.data(), P.prof, 832 * sizeof(long long), cudaMemcpyDeviceToHost));
        if (getenv("WRNN_MMA_PROF_STEPS")) {
            fprintf(stderr, "[mma steps] per step: interval to the same event of the step before | CTA G-1 intervals: fc1 ready->fc2 ready->published->seen->drawn | CTA 0: seen->rnn1 pub->rnn2 ready->rnn2 pub\n");
            for (int i = 1; i < 64; i++) {
                auto v = [&](int k) { return pr[192 + k * 64 + i]; };
                fprintf(stderr, "[mma steps] %2d  period %6lld | %6lld %6lld %6lld %6lld | %6lld %6lld %6lld | fc1 ready - drawn(prev) %6lld\n", i, v(4) - pr[192 + 4 * 64 + i - 1],
                        v(1) - v(0), v(2) - v(1), v(3) - v(2), v(4) - v(3), v(6) - v(5), v(7) - v(6), v(8) - v(7), v(0) - pr[192 + 4 * 64 + i - 1]);
            }
        }
        {
            fprintf(stderr, "[mma prof] clocks between consecutive draws of CTA G-1, from 7 steps before the profiled step (its own duration is the 8th number):");
            for (int i = 1; i < 64 && pr[128 + i]; i++) fprintf(stderr, " %lld", pr[128 + i] - pr[128 + i - 1]);
            fprintf(stderr, "\n");
        }
        static const char *names[32] = {
            "producer: step begin", "producer: rnn2 issued", "producer: recurrent-1 + next rnn1 cond issued", "producer: fc begin",
            "producer: fc1 issued", "producer: fc2 issued", "producer: fc3 issued", "producer: recurrent-2 + next rnn2 cond issued", "producer: rnn1 of the tile seen (rnn2 loads start)",
            "producer: rnn2 of the tile seen (fc1 loads start)",
            "mma: rnn2 issued", "mma: recurrent-1 + next rnn1 cond issued", "mma: fc1 issued", "mma: fc2 issued", "mma: fc3 issued",
            "mma: recurrent-2 + next rnn2 cond issued",
            "sampler: fc3 accumulator ready", "sampler: drawn", "sampler: published", "producer: fc1 of the tile seen (fc2 loads start)",
            "rnn1: accumulators ready", "rnn1: previous sample seen", "rnn1: gates done (warp 2)", "rnn1: published",
            "rnn2: accumulators ready", "rnn2: gates done (warp 2)", "rnn2: published",
            "fc1: accumulator ready", "fc1: published", "fc2: accumulator ready", "fc2: published", "producer: fc2 of the tile seen (fc3 loads start)"};
        for (int c = 0; c < 2; c++) {
            const long long *q = pr.data() + 64 * c;
            long long t0 = 0;
            for (int i = 0; i < 32; i++)
                if (q[i] && (!t0 || q[i] < t0)) t0 = q[i];
            fprintf(stderr, "[mma prof] CTA %s, step %lld, folds %lld, fold tile %d\n", c == 0 ? "0" : "G-1", (long long)P.prof_step,
                    (long long)a.B, FT);
            std::vector<std::pair<long long, int>> ev;
            for (int i = 0; i < 32; i++)
                if (q[i]) ev.push_back({q[i] - t0, i});
            std::sort(ev.begin(), ev.end());
            for (auto &e : ev) fprintf(stderr, "[mma prof]   %8lld  %s\n", e.first, names[e.second]);
            if (q[61] && q[17]) fprintf(stderr, "[mma prof]   100 steps later the sampler is at +%lld clk: %lld clk per step\n", q[61] - q[17], (q[61] - q[17]) / 100);
            if (q[62]) fprintf(stderr, "[mma prof]   inside the sampler (warp 2): past the barrier %lld, contributions added %lld (same clock)\n", q[62] - t0, q[63] - t0);
            if (q[58]) fprintf(stderr, "[mma prof]   inside rnn1 gates (warp 2): accumulators in registers %lld, gate math done %lld (same clock)\n", q[58] - t0, q[59] - t0);
            static const char *jn[6] = {"rnn2", "recurrent-1", "fc1", "fc2", "fc3", "recurrent-2"};
            static const int jtype[6] = {J_R2C, J_S1, J_F1C, J_F2C, J_F2C, J_S2};
            for (int j = 0; j < 6; j++)
                if (q[40 + 3 * j + 1])
                    fprintf(stderr, "[mma prof]   %-12s %2d stages: MMA thread full-wait %6lld, issue %6lld, commit %5lld; producer lane 0 empty-wait %6lld\n",
                            jn[j], job_stages(jtype[j]), q[40 + 3 * j], q[40 + 3 * j + 1], q[40 + 3 * j + 2], q[32 + j]);
        }
    }
    return 0;
}

}  // namespace wrnn

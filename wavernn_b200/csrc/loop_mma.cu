// Per-sample loop, tensor-core kernel for large fold batches (config 5: hundreds of folds per GPU).
//
// `generate`'s hot loop (models/fatchord_version.py:201-241) for B folds at once is, per step, a chain of skinny
// GEMMs  out[rows x B] = W[rows x K] . act[K x B]:  rnn1 (:210), rnn2 (:213-214), fc1 (:217-218), fc2 (:220-221),
// fc3 (:223).  The FMA kernels (loop_res.cu ...) spend the same time per fold whatever B is; here the four big
// layers run on the 5th-generation tensor cores:
//
//   * swap-AB: weights are the M operand (128-row tiles), folds are N (NT = 16..128 per tile), `tcgen05.mma`
//     cta_group::1 kind::f16 issued by one thread, accumulators in TMEM, read back with `tcgen05.ld` into the fused
//     GRU-gate / ReLU epilogue;
//   * precision: every operand is a bf16 hi + bf16 lo pair (x = hi + lo to ~2^-17), three MMAs per k-step
//     (hi.hi + hi.lo + lo.hi) accumulated in fp32: products carry ~1e-5 relative error, far inside the 1e-3 bound
//     on teacher-forced logits that plain bf16 breaks after a few thousand recurrent steps;
//   * operands live in global memory (L2-resident) already in the K-major, no-swizzle shared-memory image the MMA
//     descriptors address, "[k-chunk of 8][row][8 elements]", so a pipeline stage is two plain `cp.async.bulk`
//     copies (one weight block, one activation block) completing on an mbarrier: no tensor maps, no swizzle;
//   * one persistent cooperative launch for all S steps; per step four MMA phases (rnn1, rnn2, fc1, fc2) and one
//     fp32 FFMA phase (fc3 + MoL sampling + next step's conditioning operand), separated by split-phase grid
//     barriers (epilogue warps arrive, the TMA producer waits); every phase is spread over all SMs as
//     (row tile, fold tile) jobs;
//   * algebra: the I layer and both residual adds are folded into composed matrices formed on the host in double
//     (W_ih1 W_I etc., as in loop_res.cu), so the conditioning features and the previous sample enter every layer as
//     extra K columns of the same MMA chain.
//
// Warp roles in a CTA (320 threads): warp 0 = TMA producer, warp 1 = MMA issuer (owns TMEM), warps 2..9 = epilogue
// (TMEM -> registers -> shared staging -> gates -> global operands) and the FFMA phase.
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
constexpr int GRU_ROWS = 3 * UT, FC_ROWS = 128;
constexpr int FT = 128;                    // folds per tile = M of the MMA = TMEM lanes: lane i of the accumulators is fold i
constexpr int MT_GRU = H / UT, MT_FC = FCD / FC_ROWS;     // 16, 4 row tiles
constexpr int KC = 4;                      // k-chunks (of 8 elements) per pipeline stage: K = 32 = two MMA k-steps
constexpr int MAXSTAGE = 8;                // ring depth is chosen at launch: as many stages as shared memory holds, up to 8
constexpr int W_GRU = 2 * KC * GRU_ROWS * 16, W_FC = 2 * KC * FC_ROWS * 16;     // weight block of a stage, hi + lo: 12288 / 16384 B
constexpr int ACT_STAGE = 2 * KC * FT * 16;                                      // activation block of a stage, hi + lo: 16384 B
constexpr int CGROUPS = 7;                 // conditioning operand: 28 chunks = 224 elements per fold
constexpr int HGROUPS = H / 32;            // 16 stage groups in a 512-wide operand
constexpr int THREADS = 320, EPI_THREADS = 256, EPI_WARPS = 8;
constexpr int SBIAS = 2 * 2 * GRU_ROWS + 2 * FC_ROWS;     // this CTA's biases in shared memory: rnn1 / rnn2 (input, recurrent), fc1 / fc2
constexpr int NCP = 32;                    // fc3 rows padded (MoL: 30)
constexpr int TMEM_COLS = 512;
constexpr int FIXED_SMEM = (SBIAS + NCP * FCD + EPI_WARPS * 256) * 4;            // biases + fc3 rows + per-warp scratch
// conditioning operand element order (which features each layer concatenates):
//   [a4 (32) | a2 (32) | x_prev (1) | m (80) | a1 (32) | a3 (32) | zero (15)]
// so that rnn1 reads chunks 8..23, rnn2 chunks 4..23, fc1 chunks 8..27, fc2 chunks 0..3.
constexpr int C_A4 = 0, C_A2 = 32, C_X = 64, C_M = 65, C_A1 = 145, C_A3 = 177, C_END = 209;

// Job types.  A job = one (row tile, fold tile) MMA chain into one TMEM accumulator.  Per step a GRU CTA runs
//   J_R1 (rnn1, conditioning columns only)  -> gates -> J_R2 (rnn2: W_ih2a . h1' + conditioning columns) -> gates,
// then, OFF the sample-to-sample chain while the fc CTAs work, the two recurrent products of the NEXT step,
//   J_S1 (W_hh1 . h1') and J_S2 (W_hh2 . h2'), whose accumulators stay in TMEM until the next step's gates read them;
// an fc CTA runs J_F1 (fc1) -> ReLU -> J_F2 (fc2) -> ReLU.  (loop_res.cu's "shadow products", on tensor cores.)
enum { J_R1 = 0, J_R2, J_S1, J_S2, J_F1, J_F2, NJOB };
enum { ARR_C = 0, ARR_H1NEW, ARR_H2NEW, ARR_S12, ARR_Y1 };
// TMEM accumulator (first column) of each job: 128 lanes = folds, columns = weight rows.  rnn1's gates read X1 + H1,
// rnn2's X2 + H2 (96 columns each: [r | z | n] x 32 units), the fc jobs share the last 128 columns.
__host__ __device__ inline int job_acc(int jt) { return jt == J_R1 ? 0 : jt == J_S1 ? 96 : jt == J_R2 ? 192 : jt == J_S2 ? 288 : 384; }
__host__ __device__ inline int job_stages(int jt) { return jt == J_R1 ? 4 : jt == J_R2 ? 21 : jt == J_F1 ? 21 : jt == J_F2 ? 17 : 16; }
__host__ __device__ inline int job_rows(int jt) { return jt >= J_F1 ? FC_ROWS : GRU_ROWS; }
// grid counter a job's loads wait for (value G * (t + 1)), -1 = ordered behind the previous job of the same CTA
__host__ __device__ inline int job_gate(int jt) { return jt == J_R1 ? 4 : jt == J_R2 ? 0 : jt == J_S1 ? -1 : jt == J_F2 ? 2 : 1; }
struct StageInfo { int arr, g; };
__host__ __device__ inline StageInfo stage_info(int jt, int s) {
    switch (jt) {
        case J_R1: return StageInfo{ARR_C, 2 + s};
        case J_R2: return s < 16 ? StageInfo{ARR_H1NEW, s} : StageInfo{ARR_C, 1 + (s - 16)};
        case J_S1: return StageInfo{ARR_H1NEW, s};
        case J_S2: return StageInfo{ARR_H2NEW, s};
        case J_F1: return s < 16 ? StageInfo{ARR_S12, s} : StageInfo{ARR_C, 2 + (s - 16)};
        default: return s < 16 ? StageInfo{ARR_Y1, s} : StageInfo{ARR_C, 0};
    }
}

struct Params {
    LoopArgs a;
    int NT, NTILES, G, NC;
    int dbg_state;                 // WRNN_MMA_STEPS set: also store the fp32 recurrent state (wrnn_debug_mma_state)
    int64_t steps_run;             // == a.S except under WRNN_MMA_STEPS (bring-up: stop after n steps and inspect the state)
    int ring_bytes;                // shared memory of the stage ring (the staging / fc3 / scratch region follows it)
    const uint8_t *Aimg;           // packed weights: job type j at offA[j], [row tile][stage][hi|lo][KC][rows][8] bf16
    int64_t offA[NJOB];
    const float *bias;             // P1: [16][2][96] | P2: [16][2][96] | P3: [512] | P4: [512]
    const float *Wf3, *bf3;        // [NCP][512], [NCP] fp32
    uint8_t *H1[2], *H2[2], *S12, *Y1, *C;      // B operands, [fold tile][group][hi|lo][KC][NT][8] bf16
    float *h1f, *h2f, *y2f;        // [NTILES*NT][512] fp32 state / fc2 output
    GridSync gs;
    long long *prof;               // debug (WRNN_MMA_PROF=step): clock64 timeline of that step from CTA 0 and CTA G-1, else null
    int64_t prof_step;
};

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
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, void *b) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(b)) : "memory");
}
// generic-proxy writes (st.global by the epilogue threads) are read by the TMA unit (async proxy) of other SMs: EVERY writing
// thread orders its own stores against the async proxy before the CTA publishes the phase, and the producer thread does the
// same after it has acquired the grid counter (the pattern of a TMA store after generic smem writes, applied to global)
__device__ __forceinline__ void fence_proxy_async_global() { asm volatile("fence.proxy.async;" ::: "memory"); }
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
};
__device__ __forceinline__ bool give_up(Watch &w, long long t0, unsigned &spins) {
    if ((++spins & 1023u) == 0u) {
        if (clock64() - t0 > WATCHDOG_CLOCKS) { atomicExch(w.abort_flag, 3); w.dead = true; }
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
// Split-phase grid barrier, ONE MONOTONIC COUNTER PER PHASE of the step (five, each on its own 128-byte line): a CTA
// that has no job in a phase publishes it at once and runs ahead to the next thing it has to wait for, so a single
// counter summed over phases would reach "G arrivals" before every CTA has finished the phase in question.  Per phase,
// counter[p] >= G * (t + 1) does mean that every CTA has published phase p of step t: no CTA can publish phase p of
// step t + 1 before it has passed the fc3 phase of step t, which waits for everybody's fc2 phase of step t.
__device__ __forceinline__ unsigned long long *phase_counter(const GridSync &gs, int p) { return gs.counter + 16 * (p + 1); }
__device__ __forceinline__ void grid_wait(const GridSync &gs, int p, unsigned long long want, Watch &w) {
    if (w.dead) return;
    const long long t0 = clock64();
    unsigned spins = 0;
    const unsigned long long *c = phase_counter(gs, p);
    while (ld_acquire_u64(c) < want)
        if (give_up(w, t0, spins)) return;
}
__device__ __forceinline__ void epi_bar() { bar_sync(1, EPI_THREADS); }

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
    return ((size_t)(nt * NG + (chunk >> 2)) * 2 + hl) * (size_t)(KC * NT * 16) + (size_t)((chunk & 3) * NT + fi) * 16;
}

// MoL draw for one fold by one warp from 3*nmix logits in shared memory: same arithmetic as mol_sample_warp
__device__ __forceinline__ float mol_sample_smem(const float *lg, int nmix, float myu1, float u2, int lane, int *pick) {
    float t = -INFINITY;
    if (lane < nmix) t = lg[lane] - logf(-logf(myu1));
    int idx = lane;
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) {
        const float ot = __shfl_xor_sync(0xffffffffu, t, o);
        const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
        if (ot > t || (ot == t && oi < idx)) { t = ot; idx = oi; }
    }
    const float mean = lg[nmix + idx];
    const float ls = fmaxf(lg[2 * nmix + idx], -32.23619130191664f);
    float x = mean + expf(ls) * (logf(u2) - logf(1.f - u2));
    x = fminf(fmaxf(x, -1.f), 1.f);
    *pick = idx;
    return x;
}

// conditioning operand of step `t` for fold b (one warp), in two halves so that the 832-byte row (HBM, cold) can be
// fetched before the warp has to wait for anything: load_cond_row -> scr[0..207], build_cond_operand reorders, splits
// into bf16 hi/lo and stores the 28 chunks.
__device__ __forceinline__ void load_cond_row(const Params &P, int64_t b, int64_t t, float *scr, int lane) {
    const int64_t S = P.a.S;
    const int RW = FEAT + 4 * AUX;
    const float *row = P.a.cond + ((size_t)b * S + (t < S ? t : 0)) * RW;
#pragma unroll
    for (int i = 0; i < 7; i++) {
        const int e = lane + 32 * i;
        if (e < RW) scr[e] = t < S ? __ldg(row + e) : 0.f;
    }
}
__device__ __forceinline__ void build_cond_operand(const Params &P, int64_t b, float xprev, const float *scr, int lane) {
    __syncwarp();
    if (lane < 4 * CGROUPS) {
        float v[8];
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const int e = lane * 8 + i;
            float x;
            if (e < C_A2) x = scr[FEAT + 3 * AUX + (e - C_A4)];
            else if (e < C_X) x = scr[FEAT + AUX + (e - C_A2)];
            else if (e == C_X) x = xprev;
            else if (e < C_A1) x = scr[e - C_M];
            else if (e < C_A3) x = scr[FEAT + (e - C_A1)];
            else if (e < C_END) x = scr[FEAT + 2 * AUX + (e - C_A3)];
            else x = 0.f;
            v[i] = x;
        }
        const float v0[4] = {v[0], v[1], v[2], v[3]}, v1[4] = {v[4], v[5], v[6], v[7]};
        uint2 h0, l0, h1, l1;
        split4(v0, h0, l0);
        split4(v1, h1, l1);
        const int nt = (int)(b / P.NT), fi = (int)(b % P.NT);
        *reinterpret_cast<uint4 *>(P.C + b_slot(P.NT, CGROUPS, nt, lane, fi, 0)) = make_uint4(h0.x, h0.y, h1.x, h1.y);
        *reinterpret_cast<uint4 *>(P.C + b_slot(P.NT, CGROUPS, nt, lane, fi, 1)) = make_uint4(l0.x, l0.y, l1.x, l1.y);
    }
    __syncwarp();
}

// operand array a job reads at step t (t = -1: the prologue's recurrent products on the zero state)
__device__ __forceinline__ const uint8_t *b_array(const Params &P, int arr, int64_t t) {
    switch (arr) {
        case ARR_C: return P.C;
        case ARR_H1NEW: return P.H1[(t + 1) & 1];
        case ARR_H2NEW: return P.H2[(t + 1) & 1];
        case ARR_S12: return P.S12;
        default: return P.Y1;
    }
}

struct Ring {                   // the stage ring of this CTA and a position in it (producer and MMA issuer keep one each)
    uint32_t nstage, w_slot, stage_bytes;       // depth, bytes reserved for the weight block of a stage, bytes per stage
    uint32_t slot, par;
    __device__ __forceinline__ void advance() {
        if (++slot == nstage) { slot = 0; par ^= 1u; }
    }
};
// A CTA that only ever runs GRU jobs (96-row weight blocks) packs its stages tighter than one that also has an fc job;
// the ring is as deep as its share of shared memory allows (5 stages of 28 KB for a GRU CTA).
__device__ __forceinline__ Ring make_ring(const Params &P, bool has_fc) {
    Ring r;
    r.w_slot = has_fc ? W_FC : W_GRU;
    r.stage_bytes = r.w_slot + ACT_STAGE;
    r.nstage = min((uint32_t)MAXSTAGE, (uint32_t)P.ring_bytes / r.stage_bytes);
    r.slot = 0;
    r.par = 0;
    return r;
}

// One job of the TMA producer lanes: (wait for the grid counter,) stream the stages of (job type, row tile, fold tile).
// A thread completes only about one bulk copy per 700 clocks whatever its size (scripts/mb_umma.cu: 16 B/clk/SM from one
// issuing thread, 52 from four lanes, ~67 = the L2 limit from eight), so the copies are spread over NPROD = 8 lanes of
// the producer warp: stage k of the running count belongs to lane pair (k & 3), the even lane arms the barrier with the
// byte count of both copies and brings the weight block, the odd lane brings the activation block.  Every lane keeps
// the same ring position; a copy may complete its bytes on the barrier before the other lane has armed it (the phase
// cannot end before the arrive).
constexpr int NPROD = 8;
__device__ __forceinline__ void produce_job(const Params &P, int jt, int mt, int nt, int64_t t, Ring &r, uint32_t &count, int lane,
                                            uint8_t *stages, uint64_t *bar_full, uint64_t *bar_empty, Watch &w, long long *pacc = nullptr) {
    const int nst = job_stages(jt), gate = job_gate(jt);
    const uint32_t abytes = 2u * KC * (uint32_t)job_rows(jt) * 16u, bbytes = ACT_STAGE;       // weight block, activation block
    if (gate >= 0) {            // this job's operands were written by that phase of every CTA
        grid_wait(P.gs, gate, (unsigned long long)P.G * (unsigned long long)(t + 1), w);
        fence_proxy_async_global();
    }
    const uint8_t *Abase = P.Aimg + P.offA[jt] + (size_t)mt * nst * abytes;
    for (int s = 0; s < nst; s++, count++) {
        if ((count & 3u) == (uint32_t)(lane >> 1)) {
            const long long c0 = pacc ? clock64() : 0;
            mbar_wait(&bar_empty[r.slot], r.par ^ 1u, w);
            if (pacc && lane == 0) *pacc += clock64() - c0;
            uint8_t *dst = stages + (size_t)r.slot * r.stage_bytes;
            if ((lane & 1) == 0) {
                mbar_expect_tx(&bar_full[r.slot], abytes + bbytes);
                bulk_g2s(dst, Abase + (size_t)s * abytes, abytes, &bar_full[r.slot]);
            } else {
                const StageInfo si = stage_info(jt, s);
                const int NG = si.arr == ARR_C ? CGROUPS : HGROUPS;
                bulk_g2s(dst + r.w_slot, b_array(P, si.arr, t) + (size_t)(nt * NG + si.g) * bbytes, bbytes, &bar_full[r.slot]);
            }
        }
        r.advance();
    }
}

// One job of the MMA issuer thread: three MMAs (hi.hi, hi.lo, lo.hi) per k-step into the job's accumulator.
__device__ __forceinline__ void issue_job(const Params &P, int jt, Ring &r, uint32_t stage0, uint32_t tb,
                                          uint64_t *bar_full, uint64_t *bar_empty, uint64_t *bar_tfull, Watch &w, long long *macc = nullptr) {
    const int nst = job_stages(jt);
    const uint32_t rows = job_rows(jt);
    // A = activations (M = 128 folds), B = weights (N = rows); both K-major, no swizzle, [k-chunk][row][8].  Descriptors
    // differ only in the 14-bit start-address field (16-byte units): constant part once per job.
    const uint64_t dW0 = make_desc(0, rows * 16u, 128), dX0 = make_desc(0, FT * 16u, 128);
    const uint32_t stepW = 2u * rows, stepX = 2u * FT;          // two k-chunks, in 16-byte units
    const uint32_t loW = KC * rows, loX = KC * FT;               // hi -> lo image
    const uint32_t d = tb + (uint32_t)job_acc(jt);
    const uint32_t idesc = make_idesc(FT, (int)rows);
    for (int s = 0; s < nst; s++) {
        const long long c0 = macc ? clock64() : 0;
        mbar_wait(&bar_full[r.slot], r.par, w);
        tc_fence_after();
        const long long c1 = macc ? clock64() : 0;
        const uint32_t aW = (stage0 + r.slot * r.stage_bytes) >> 4, aX = aW + (r.w_slot >> 4);
#pragma unroll
        for (int ks = 0; ks < KC / 2; ks++) {
            const uint64_t wh = dW0 + (aW + ks * stepW), wl = dW0 + (aW + loW + ks * stepW);
            const uint64_t xh = dX0 + (aX + ks * stepX), xl = dX0 + (aX + loX + ks * stepX);
            umma_bf16(d, xh, wh, idesc, (s == 0 && ks == 0) ? 0u : 1u);
            umma_bf16(d, xh, wl, idesc, 1u);
            umma_bf16(d, xl, wh, idesc, 1u);
        }
        const long long c2 = macc ? clock64() : 0;
        umma_commit(&bar_empty[r.slot]);          // the stage is free once these MMAs have read it
        if (macc) { const long long c3 = clock64(); macc[0] += c1 - c0; macc[1] += c2 - c1; macc[2] += c3 - c2; }
        r.advance();
    }
    if (jt != J_S1 && jt != J_S2) umma_commit(bar_tfull);       // accumulator (and every earlier MMA) complete
}

// ---------------------------------------------------------------------------- epilogues (256 threads, warps 2..9)
// Accumulators are [128 folds (TMEM lanes)] x [weight rows (columns)], so a thread owns ONE fold (lane 32 q + lane of the
// warp's lane quarter q = warp % 4) and reads the r, z, n columns of a unit straight from its own lane: no shared-memory
// staging, no barrier inside an epilogue; the two warps of a quarter split the columns (`half`).  Operand stores are one
// 16-byte chunk (8 units) per image per thread, consecutive lanes = consecutive folds = consecutive addresses.
//
// GRU gates (fatchord_version.py:210-216): X (input side) and H (recurrent side, formed one step ahead) hold columns
// [r | z | n] x 32 units; the thread keeps the state of its (fold, 16 units) in registers for the whole run (`hst`).
// rnn2 also emits s12 = h1' + h2' (fc1 reads x + h1 + h2; x is folded into fc1's weights).
template <int LAYER>
__device__ __forceinline__ void epi_gru(const Params &P, int mt, int nt, int64_t t, uint32_t tb, const float *sb, float (&hst)[16],
                                        float (&h1new)[16], int q, int half, int lane) {
    const int fi = q * 32 + lane, u0 = half * 16, gu = mt * UT + u0;
    const size_t fold = (size_t)nt * FT + fi;
    const float *bi = sb + LAYER * 2 * GRU_ROWS, *bh = bi + GRU_ROWS;
    float *hf = LAYER == 0 ? P.h1f : P.h2f;
    uint8_t *Hn = LAYER == 0 ? P.H1[(t + 1) & 1] : P.H2[(t + 1) & 1];
    const uint32_t base = tb + ((uint32_t)(q * 32) << 16), colX = base + (uint32_t)job_acc(LAYER == 0 ? J_R1 : J_R2) + u0,
                   colH = base + (uint32_t)job_acc(LAYER == 0 ? J_S1 : J_S2) + u0;
#pragma unroll
    for (int g = 0; g < 2; g++) {
        uint32_t xr[8], xz[8], xn[8], hr[8], hz[8], hn_[8];
        tmem_ld8_nowait(colX + 8 * g, xr);
        tmem_ld8_nowait(colX + UT + 8 * g, xz);
        tmem_ld8_nowait(colX + 2 * UT + 8 * g, xn);
        tmem_ld8_nowait(colH + 8 * g, hr);
        tmem_ld8_nowait(colH + UT + 8 * g, hz);
        tmem_ld8_nowait(colH + 2 * UT + 8 * g, hn_);
        tmem_ld_wait();
        float hn[8];
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const int u = u0 + 8 * g + i;
            hn[i] = gru_update_fast(__uint_as_float(xr[i]) + bi[u], __uint_as_float(xz[i]) + bi[UT + u], __uint_as_float(xn[i]) + bi[2 * UT + u],
                                    __uint_as_float(hr[i]) + bh[u], __uint_as_float(hz[i]) + bh[UT + u], __uint_as_float(hn_[i]) + bh[2 * UT + u],
                                    hst[8 * g + i]);
            hst[8 * g + i] = hn[i];
        }
        if (P.dbg_state) {                  // fp32 copy of the state for wrnn_debug_mma_state only
            float4 *dst = reinterpret_cast<float4 *>(hf + fold * H + gu + 8 * g);
            dst[0] = make_float4(hn[0], hn[1], hn[2], hn[3]);
            dst[1] = make_float4(hn[4], hn[5], hn[6], hn[7]);
        }
        uint4 hi, lo;
        split8(hn, hi, lo);
        const int chunk = (gu >> 3) + g;
        *reinterpret_cast<uint4 *>(Hn + b_slot(FT, HGROUPS, nt, chunk, fi, 0)) = hi;
        *reinterpret_cast<uint4 *>(Hn + b_slot(FT, HGROUPS, nt, chunk, fi, 1)) = lo;
        if (LAYER == 1) {
            float sm[8];
#pragma unroll
            for (int i = 0; i < 8; i++) sm[i] = h1new[8 * g + i] + hn[i];
            split8(sm, hi, lo);
            *reinterpret_cast<uint4 *>(P.S12 + b_slot(FT, HGROUPS, nt, chunk, fi, 0)) = hi;
            *reinterpret_cast<uint4 *>(P.S12 + b_slot(FT, HGROUPS, nt, chunk, fi, 1)) = lo;
        } else {
#pragma unroll
            for (int i = 0; i < 8; i++) h1new[8 * g + i] = hn[i];
        }
    }
}

// fc1 / fc2 + ReLU (:217-221): 128 rows = 128 columns, 64 per thread.  fc1 -> bf16 hi/lo operand of fc2, fc2 -> fp32 rows for fc3
template <int LAYER>
__device__ __forceinline__ void epi_fc(const Params &P, int mt, int nt, uint32_t tb, const float *sb, int q, int half, int lane) {
    const int fi = q * 32 + lane, r0 = half * 64, gr = mt * FC_ROWS + r0;
    const size_t fold = (size_t)nt * FT + fi;
    const float *bb = sb + 4 * GRU_ROWS + LAYER * FC_ROWS + r0;
    const uint32_t col = tb + ((uint32_t)(q * 32) << 16) + (uint32_t)job_acc(J_F1) + r0;
#pragma unroll 2
    for (int g = 0; g < 8; g += 2) {
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
        if (LAYER == 0) {
            uint4 hi, lo;
            split8(va, hi, lo);
            *reinterpret_cast<uint4 *>(P.Y1 + b_slot(FT, HGROUPS, nt, (gr >> 3) + g, fi, 0)) = hi;
            *reinterpret_cast<uint4 *>(P.Y1 + b_slot(FT, HGROUPS, nt, (gr >> 3) + g, fi, 1)) = lo;
            split8(vb, hi, lo);
            *reinterpret_cast<uint4 *>(P.Y1 + b_slot(FT, HGROUPS, nt, (gr >> 3) + g + 1, fi, 0)) = hi;
            *reinterpret_cast<uint4 *>(P.Y1 + b_slot(FT, HGROUPS, nt, (gr >> 3) + g + 1, fi, 1)) = lo;
        } else {
            float4 *dst = reinterpret_cast<float4 *>(P.y2f + fold * FCD + gr + 8 * g);
            dst[0] = make_float4(va[0], va[1], va[2], va[3]);
            dst[1] = make_float4(va[4], va[5], va[6], va[7]);
            dst[2] = make_float4(vb[0], vb[1], vb[2], vb[3]);
            dst[3] = make_float4(vb[4], vb[5], vb[6], vb[7]);
        }
    }
}

__global__ void __launch_bounds__(THREADS, 1) mma_loop_kernel(const Params P) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t bar_full[MAXSTAGE], bar_empty[MAXSTAGE], bar_tfull;
    __shared__ uint32_t tmem_base_s;
    __shared__ int dead_s;
    uint8_t *stages = smem;
    float *sbias = reinterpret_cast<float *>(smem + P.ring_bytes);      // this CTA's bias rows, layout of SBIAS
    float *wf3 = sbias + SBIAS;                             // [NCP][512]
    float *scr = wf3 + NCP * FCD;                           // per epilogue warp: 256 floats

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, cta = blockIdx.x;
    const int G = P.G;
    const int64_t S = P.a.S, B = P.a.B;
    Watch w{P.gs.abort_flag, false};
    // static roles: GRU (row tile, fold tile) jobs fill CTAs from 0 up, fc jobs from G - 1 down; at most one of each per CTA
    const bool has_gru = cta < MT_GRU * P.NTILES, has_fc = G - 1 - cta < MT_FC * P.NTILES;
    const int mt_g = cta % MT_GRU, nt_g = cta / MT_GRU, mt_f = (G - 1 - cta) % MT_FC, nt_f = (G - 1 - cta) / MT_FC;

    if (tid == 0) {
        for (int i = 0; i < MAXSTAGE; i++) { mbar_init(&bar_full[i], 1); mbar_init(&bar_empty[i], 1); }
        mbar_init(&bar_tfull, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) tmem_alloc(&tmem_base_s, TMEM_COLS);
    for (int i = tid; i < NCP * FCD; i += THREADS) wf3[i] = P.Wf3[i];
    for (int i = tid; i < SBIAS; i += THREADS) {            // P.bias: rnn1 [16][2][96] | rnn2 [16][2][96] | fc1 [512] | fc2 [512]
        float v = 0.f;
        if (i < 4 * GRU_ROWS) {
            if (has_gru) v = P.bias[(size_t)((i / (2 * GRU_ROWS)) * MT_GRU + mt_g) * 2 * GRU_ROWS + i % (2 * GRU_ROWS)];
        } else if (has_fc) {
            const int j = i - 4 * GRU_ROWS;
            v = P.bias[(size_t)2 * MT_GRU * 2 * GRU_ROWS + (size_t)(j / FC_ROWS) * FCD + mt_f * FC_ROWS + j % FC_ROWS];
        }
        sbias[i] = v;
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tb = tmem_base_s;

    if (warp == 0) {
        // ======================================================================= TMA producer
        if (lane < NPROD) {
            Ring r = make_ring(P, has_fc);
            uint32_t count = 0;
            if (has_gru) {      // step "-1": recurrent products of step 0 on the zero state (h = 0, :194-195)
                produce_job(P, J_S1, mt_g, nt_g, -1, r, count, lane, stages, bar_full, bar_empty, w);
                produce_job(P, J_S2, mt_g, nt_g, -1, r, count, lane, stages, bar_full, bar_empty, w);
            }
            for (int64_t t = 0; t < P.steps_run && !w.dead; t++) {
                long long *pa = (P.prof && t == P.prof_step && (cta == 0 || cta == G - 1)) ? P.prof + (cta == 0 ? 0 : 64) + 32 : nullptr;
                if (has_gru) {
                    PROF_MARK(0);
                    produce_job(P, J_R1, mt_g, nt_g, t, r, count, lane, stages, bar_full, bar_empty, w, pa ? pa + 0 : nullptr);
                    PROF_MARK(1);
                    produce_job(P, J_R2, mt_g, nt_g, t, r, count, lane, stages, bar_full, bar_empty, w, pa ? pa + 1 : nullptr);
                    PROF_MARK(2);
                    produce_job(P, J_S1, mt_g, nt_g, t, r, count, lane, stages, bar_full, bar_empty, w, pa ? pa + 2 : nullptr);
                    PROF_MARK(3);
                }
                if (has_fc) {
                    PROF_MARK(4);
                    produce_job(P, J_F1, mt_f, nt_f, t, r, count, lane, stages, bar_full, bar_empty, w, pa ? pa + 4 : nullptr);
                    PROF_MARK(5);
                    produce_job(P, J_F2, mt_f, nt_f, t, r, count, lane, stages, bar_full, bar_empty, w, pa ? pa + 5 : nullptr);
                    PROF_MARK(6);
                }
                if (has_gru) {
                    produce_job(P, J_S2, mt_g, nt_g, t, r, count, lane, stages, bar_full, bar_empty, w, pa ? pa + 3 : nullptr);
                    PROF_MARK(7);
                }
            }
        }
    } else if (warp == 1) {
        // ======================================================================= MMA issuer
        if (lane == 0) {
            Ring r = make_ring(P, has_fc);
            const uint32_t stage0 = smem_u32(stages);
            if (has_gru) {
                issue_job(P, J_S1, r, stage0, tb, bar_full, bar_empty, &bar_tfull, w);
                issue_job(P, J_S2, r, stage0, tb, bar_full, bar_empty, &bar_tfull, w);
            }
            for (int64_t t = 0; t < P.steps_run && !w.dead; t++) {
                long long *ma = (P.prof && t == P.prof_step && (cta == 0 || cta == G - 1)) ? P.prof + (cta == 0 ? 0 : 64) + 40 : nullptr;
                if (has_gru) {
                    issue_job(P, J_R1, r, stage0, tb, bar_full, bar_empty, &bar_tfull, w, ma ? ma + 0 : nullptr);
                    PROF_MARK(10);
                    issue_job(P, J_R2, r, stage0, tb, bar_full, bar_empty, &bar_tfull, w, ma ? ma + 3 : nullptr);
                    PROF_MARK(11);
                    issue_job(P, J_S1, r, stage0, tb, bar_full, bar_empty, &bar_tfull, w, ma ? ma + 6 : nullptr);
                    PROF_MARK(12);
                }
                if (has_fc) {
                    issue_job(P, J_F1, r, stage0, tb, bar_full, bar_empty, &bar_tfull, w, ma ? ma + 12 : nullptr);
                    PROF_MARK(13);
                    issue_job(P, J_F2, r, stage0, tb, bar_full, bar_empty, &bar_tfull, w, ma ? ma + 15 : nullptr);
                    PROF_MARK(14);
                }
                if (has_gru) {
                    issue_job(P, J_S2, r, stage0, tb, bar_full, bar_empty, &bar_tfull, w, ma ? ma + 9 : nullptr);
                    PROF_MARK(15);
                }
            }
        }
        __syncwarp();
    } else {
        // ======================================================================= epilogue warps + fc3 / sampling phase
        const int etid = tid - 64, ewarp = warp - 2;
        const int q = warp & 3, half = ewarp >> 2;          // TMEM lane quarter this warp may read; which accumulator / idle
        float *myscr = scr + ewarp * 256;
        const uint2 key = make_uint2((uint32_t)P.a.seed, (uint32_t)(P.a.seed >> 32));
        const int nmix = P.NC / 3;
        uint32_t tpar = 0;                                   // parity of bar_tfull: one completion per job with an epilogue
        // recurrent state of this thread's (fold, 16 units), both layers, and rnn1's new state for s12: registers for the whole run
        float hst1[16], hst2[16], h1new[16];
#pragma unroll
        for (int i = 0; i < 16; i++) hst1[i] = hst2[i] = h1new[i] = 0.f;        // h1 = h2 = 0 (:194-195)
        auto publish = [&](int p) {                          // every writing thread fences, one thread releases the counter
            fence_proxy_async_global();
            epi_bar();
            if (etid == 0) red_release_add_u64(phase_counter(P.gs, p), 1ull);
        };
        auto wait_acc = [&]() {
            mbar_wait(&bar_tfull, tpar, w);
            tpar ^= 1u;
            __syncwarp();       // lanes leave the spin loop one by one; tcgen05.ld is .sync.aligned: converge first
            tc_fence_after();
        };
        // ---- step "-1": conditioning operand of step 0 (x_prev = 0 or the teacher's first sample)
        for (int64_t b = cta + (int64_t)G * ewarp; b < B; b += (int64_t)G * EPI_WARPS) {
            const float x0 = P.a.teacher ? P.a.teacher[(size_t)b * S] : 0.f;
            load_cond_row(P, b, 0, myscr, lane);
            build_cond_operand(P, b, x0, myscr, lane);
        }
        publish(4);

        for (int64_t t = 0; t < P.steps_run; t++) {
            if (w.dead) break;                               // uniform: set below from a flag read after a barrier
            if (has_gru) {
                wait_acc();
                if (etid == 0) PROF_MARK(20);
                epi_gru<0>(P, mt_g, nt_g, t, tb, sbias, hst1, h1new, q, half, lane);
                if (etid == 0) PROF_MARK(16);
                tc_fence_before();
            }
            publish(0);
            if (etid == 0) PROF_MARK(21);
            if (has_gru) {
                wait_acc();
                if (etid == 0) PROF_MARK(22);
                epi_gru<1>(P, mt_g, nt_g, t, tb, sbias, hst2, h1new, q, half, lane);
                if (etid == 0) PROF_MARK(17);
                tc_fence_before();
            }
            publish(1);
            if (etid == 0) PROF_MARK(23);
            if (has_fc) {
                wait_acc();
                if (etid == 0) PROF_MARK(24);
                epi_fc<0>(P, mt_f, nt_f, tb, sbias, q, half, lane);
                if (etid == 0) PROF_MARK(18);
                tc_fence_before();
            }
            publish(2);
            if (etid == 0) PROF_MARK(25);
            if (has_fc) {
                wait_acc();
                if (etid == 0) PROF_MARK(26);
                epi_fc<1>(P, mt_f, nt_f, tb, sbias, q, half, lane);
                if (etid == 0) PROF_MARK(19);
                tc_fence_before();
            }
            publish(3);
            if (etid == 0) PROF_MARK(27);
            // ---------------- fc3 (:223) in fp32 + MoL draw (utils/distribution.py:87-123) + next conditioning operand.
            // What does not depend on fc2 is done before the wait: next step's conditioning row (HBM) and the noise.
            const int64_t bfirst = cta + (int64_t)G * ewarp;
            float myu1 = 0.5f, u2 = 0.5f;
            auto draw_noise = [&](int64_t b) {
                if (P.a.u1) {
                    if (lane < nmix) myu1 = P.a.u1[((size_t)t * P.a.Btot + P.a.b0 + b) * nmix + lane];
                    u2 = P.a.u2[(size_t)t * P.a.Btot + P.a.b0 + b];
                } else {
                    const uint4 r4 = philox4x32(make_uint4((uint32_t)t, (uint32_t)(P.a.b0 + b), (uint32_t)(lane >> 2), 1u), key);
                    const uint32_t rr[4] = {r4.x, r4.y, r4.z, r4.w};
                    myu1 = u_clip(rr[lane & 3]);
                    const uint4 r2 = philox4x32(make_uint4((uint32_t)t, (uint32_t)(P.a.b0 + b), 0u, 2u), key);
                    u2 = u_clip(r2.x);
                }
            };
            if (bfirst < B) {
                load_cond_row(P, bfirst, t + 1, myscr, lane);
                draw_noise(bfirst);
            }
            if (etid == 0) grid_wait(P.gs, 3, (unsigned long long)G * (unsigned long long)(t + 1), w);
            epi_bar();
            if (etid == 0) PROF_MARK(28);
            for (int64_t b = bfirst; b < B; b += (int64_t)G * EPI_WARPS) {
                if (b != bfirst) {
                    load_cond_row(P, b, t + 1, myscr, lane);
                    draw_noise(b);
                }
                float4 y[4];
#pragma unroll
                for (int jj = 0; jj < 4; jj++) y[jj] = __ldcg(reinterpret_cast<const float4 *>(P.y2f + (size_t)b * FCD + 128 * jj + lane * 4));
                float acc[NCP];
#pragma unroll
                for (int r = 0; r < NCP; r++) {
                    float a = 0.f;
#pragma unroll
                    for (int jj = 0; jj < 4; jj++) {
                        const float4 wv = *reinterpret_cast<const float4 *>(wf3 + r * FCD + 128 * jj + lane * 4);
                        a = fmaf(wv.x, y[jj].x, a);
                        a = fmaf(wv.y, y[jj].y, a);
                        a = fmaf(wv.z, y[jj].z, a);
                        a = fmaf(wv.w, y[jj].w, a);
                    }
                    acc[r] = a;
                }
                warp_reduce_scatter<NCP>(acc, lane);
                const int row = reduce_scatter_base<NCP>(lane);
                const float lg = acc[0] + __ldg(P.bf3 + row);
                __syncwarp();
                myscr[224 + row] = lg;
                __syncwarp();
                const bool live = t < P.a.fold_len[b];
                if (P.a.logits_out && live && lane < P.NC) P.a.logits_out[((size_t)b * S + t) * P.NC + lane] = myscr[224 + lane];
                int pick = 0;
                float xv = mol_sample_smem(myscr + 224, nmix, myu1, u2, lane, &pick);
                if (!live) xv = 0.f;
                if (lane == 0 && live) {
                    P.a.samples[(size_t)b * S + t] = xv;
                    P.a.picks[(size_t)b * S + t] = pick;
                }
                if (P.a.teacher) xv = t + 1 < S ? P.a.teacher[(size_t)b * S + t + 1] : 0.f;
                build_cond_operand(P, b, xv, myscr, lane);
            }
            if (etid == 0) PROF_MARK(29);
            publish(4);
            if (etid == 0) PROF_MARK(30);
            // uniform exit on a fired watchdog
            if (etid == 0) dead_s = (w.dead || ld_volatile_s32(P.gs.abort_flag)) ? 1 : 0;
            epi_bar();
            w.dead = dead_s != 0;
            epi_bar();
        }
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
           d.NC <= mma::NCP;
}
int mma_min_folds() { return 48; }

int mma_plan(wrnn_handle *h) {
    using namespace mma;
    const Dims &d = h->d;
    MmaPlan &pl = h->mma;
    pl.ok = false;
    if (!(d.mode == WRNN_MODE_MOL && d.H == H && d.FC == FCD && d.A == AUX && d.F == FEAT && d.NC <= NCP)) return 0;
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
        const bool in_x = e >= C_X && e < C_A3;        // x_prev, m, a1: column e - 64 of the composed matrix
        if (p == 0) return in_x ? Wc1[(size_t)grow * KI + (e - C_X)] : 0.0;
        if (p == 1) {
            if (e >= C_A2 && e < C_X) return W2i->v[(size_t)grow * (H + AUX) + H + (e - C_A2)];
            return in_x ? Wc2[(size_t)grow * KI + (e - C_X)] : 0.0;
        }
        if (p == 2) {
            if (e >= C_A3 && e < C_END) return F1->v[(size_t)grow * (H + AUX) + H + (e - C_A3)];
            return in_x ? Wc3[(size_t)grow * KI + (e - C_X)] : 0.0;
        }
        return (e >= C_A4 && e < C_A2) ? F2->v[(size_t)grow * (FCD + AUX) + FCD + (e - C_A4)] : 0.0;
    };
    // weight on a 512-wide operand column k for job type jt
    auto main_w = [&](int jt, int grow, int k) -> double {
        switch (jt) {
            case J_R2: return W2i->v[(size_t)grow * (H + AUX) + k];
            case J_S1: return W1h->v[(size_t)grow * H + k];
            case J_S2: return W2h->v[(size_t)grow * H + k];
            case J_F1: return F1->v[(size_t)grow * (H + AUX) + k];
            default: return F2->v[(size_t)grow * (FCD + AUX) + k];
        }
    };
    const int cond_phase[NJOB] = {0, 1, -1, -1, 2, 3};     // which layer's conditioning columns a job carries
    size_t total = 0;
    for (int jt = 0; jt < NJOB; jt++) {
        pl.offA[jt] = (int64_t)total;
        total += (size_t)(jt >= J_F1 ? MT_FC : MT_GRU) * job_stages(jt) * 2 * KC * job_rows(jt) * 16;
    }
    std::vector<uint16_t> img(total / 2 + 2048, 0);       // + slack: the 96-row image is addressed as 128 rows
    for (int jt = 0; jt < NJOB; jt++) {
        const int rows = job_rows(jt), mts = jt >= J_F1 ? MT_FC : MT_GRU, nst = job_stages(jt);
        for (int mt = 0; mt < mts; mt++)
            for (int s = 0; s < nst; s++) {
                const StageInfo si = stage_info(jt, s);
                uint16_t *blk = img.data() + (size_t)pl.offA[jt] / 2 + ((size_t)mt * nst + s) * (2 * KC * rows * 8);
                for (int r = 0; r < rows; r++) {
                    const int grow = jt < J_F1 ? (r / UT) * H + mt * UT + (r % UT) : mt * FC_ROWS + r;
                    for (int kl = 0; kl < KC * 8; kl++) {
                        const int feat = si.g * 32 + kl;
                        const double wv = si.arr == ARR_C ? cond_w(cond_phase[jt], grow, feat) : main_w(jt, grow, feat);
                        const float wf = (float)wv;
                        const uint16_t hi = f2bf(wf), lo = f2bf((float)(wv - (double)bf2f(hi)));
                        const size_t o = ((size_t)(kl / 8) * rows + r) * 8 + kl % 8;
                        blk[o] = hi;
                        blk[(size_t)KC * rows * 8 + o] = lo;
                    }
                }
            }
    }
    std::vector<float> bias((size_t)2 * MT_GRU * 2 * GRU_ROWS + 2 * FCD, 0.f);
    for (int p = 0; p < 2; p++)
        for (int mt = 0; mt < MT_GRU; mt++)
            for (int r = 0; r < GRU_ROWS; r++) {
                const int grow = (r / UT) * H + mt * UT + (r % UT);
                float *b = bias.data() + (size_t)(p * MT_GRU + mt) * 2 * GRU_ROWS;
                b[r] = (float)((p == 0 ? bc1[grow] : bc2[grow]) + (double)(p == 0 ? B1i : B2i)->v[grow]);
                b[GRU_ROWS + r] = (p == 0 ? B1h : B2h)->v[grow];
            }
    for (int r = 0; r < FCD; r++) {
        bias[(size_t)2 * MT_GRU * 2 * GRU_ROWS + r] = (float)(bc3[r] + (double)F1b->v[r]);
        bias[(size_t)2 * MT_GRU * 2 * GRU_ROWS + FCD + r] = F2b->v[r];
    }
    std::vector<float> wf3((size_t)NCP * FCD + NCP, 0.f);
    for (int r = 0; r < d.NC; r++) {
        memcpy(wf3.data() + (size_t)r * FCD, F3->v.data() + (size_t)r * FCD, FCD * sizeof(float));
        wf3[(size_t)NCP * FCD + r] = F3b->v[r];
    }
    if (pl.Aimg) cudaFree(pl.Aimg);
    if (pl.bias) cudaFree(pl.bias);
    if (pl.wf3) cudaFree(pl.wf3);
    pl.Aimg = nullptr; pl.bias = nullptr; pl.wf3 = nullptr;
    WRNN_CUDA(h, cudaMalloc((void **)&pl.Aimg, img.size() * 2));
    WRNN_CUDA(h, cudaMemcpy(pl.Aimg, img.data(), img.size() * 2, cudaMemcpyHostToDevice));
    WRNN_CUDA(h, cudaMalloc((void **)&pl.bias, bias.size() * 4));
    WRNN_CUDA(h, cudaMemcpy(pl.bias, bias.data(), bias.size() * 4, cudaMemcpyHostToDevice));
    WRNN_CUDA(h, cudaMalloc((void **)&pl.wf3, wf3.size() * 4));
    WRNN_CUDA(h, cudaMemcpy(pl.wf3, wf3.data(), wf3.size() * 4, cudaMemcpyHostToDevice));
    pl.img_bytes = total;
    // everything the device offers minus the static part (barriers): the stage ring takes what staging / fc3 rows leave
    pl.smem_bytes = (h->smem_optin - 1024) / 1024 * 1024;
    if (pl.smem_bytes < (size_t)FIXED_SMEM + 2 * (W_FC + ACT_STAGE)) return 0;
    WRNN_CUDA(h, cudaFuncSetAttribute(mma::mma_loop_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem_bytes));
    pl.ok = true;
    return 0;
}

// The fold tile is the M of the MMA: 128 folds = the 128 TMEM lanes, fixed.  16 GRU jobs + 4 fc jobs per fold tile; up to
// num_sms / 16 tiles (9 on a B200: 1152 folds) so that every CTA has at most one GRU job; fc jobs fill CTAs from the top.
int64_t mma_max_folds(const wrnn_handle *h) { return (int64_t)128 * (h->num_sms / mma::MT_GRU); }

int mma_run(wrnn_handle *h, const LoopArgs &a, cudaStream_t st) {
    using namespace mma;
    const MmaPlan &pl = h->mma;
    Params P;
    P.a = a;
    P.G = h->num_sms;
    P.NT = FT;
    P.NTILES = (int)((a.B + P.NT - 1) / P.NT);
    if (MT_GRU * P.NTILES > P.G)
        return fail(h, WRNN_ERR_ARG, "tensor-core loop kernel: %lld folds in tiles of %d need %d CTAs, the device has %d",
                    (long long)a.B, P.NT, MT_GRU * P.NTILES, P.G);
    P.ring_bytes = (int)(pl.smem_bytes - FIXED_SMEM) / 1024 * 1024;
    P.NC = h->d.NC;
    P.steps_run = a.S;
    P.dbg_state = 0;
    if (const char *e = getenv("WRNN_MMA_STEPS")) {
        P.steps_run = std::min<int64_t>(a.S, std::max(1, atoi(e)));
        P.dbg_state = 1;
    }
    P.Aimg = pl.Aimg;
    for (int j = 0; j < NJOB; j++) P.offA[j] = pl.offA[j];
    P.bias = pl.bias;
    P.Wf3 = pl.wf3;
    P.bf3 = pl.wf3 + (size_t)NCP * FCD;
    const size_t NF = (size_t)P.NTILES * P.NT;
    const size_t bH = NF * H * 4, bC = NF * CGROUPS * 128, bF = NF * H * sizeof(float);     // hi + lo bf16 = 4 B per element
    const size_t total = 6 * bH + bC + 3 * bF + 4096;
    WRNN_TRY(ws_reserve(h, h->ws_act, total));
    WRNN_CUDA(h, cudaMemsetAsync(h->ws_act.p, 0, total, st));       // h1 = h2 = 0 (:194-195), padded folds stay zero
    uint8_t *p8 = (uint8_t *)h->ws_act.p;
    P.H1[0] = p8; p8 += bH; P.H1[1] = p8; p8 += bH;
    P.H2[0] = p8; p8 += bH; P.H2[1] = p8; p8 += bH;
    P.S12 = p8; p8 += bH; P.Y1 = p8; p8 += bH;
    P.C = p8; p8 += (bC + 1023) / 1024 * 1024;
    P.h1f = (float *)p8; p8 += bF; P.h2f = (float *)p8; p8 += bF; P.y2f = (float *)p8;
    h->mma_dbg[0] = P.h1f; h->mma_dbg[1] = P.h2f; h->mma_dbg[2] = P.y2f;
    WRNN_TRY(ws_reserve(h, h->ws_mma_sync, 1024));                    // five phase counters, 128 bytes apart, from +128
    WRNN_CUDA(h, cudaMemsetAsync(h->ws_mma_sync.p, 0, 1024, st));
    P.gs.counter = (unsigned long long *)h->ws_mma_sync.p;
    P.gs.abort_flag = (int *)((char *)h->ws_sync.p + 64);             // the call's sticky abort flag (api_run.cu)
    WRNN_CUDA(h, cudaFuncSetAttribute(mma_loop_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem_bytes));
    P.prof = nullptr;
    P.prof_step = -1;
    const char *pe = getenv("WRNN_MMA_PROF");
    if (pe) {
        WRNN_TRY(ws_reserve(h, h->ws_prof, 128 * sizeof(long long)));
        WRNN_CUDA(h, cudaMemsetAsync(h->ws_prof.p, 0, 128 * sizeof(long long), st));
        P.prof = (long long *)h->ws_prof.p;
        P.prof_step = std::min<int64_t>(atoi(pe), P.steps_run - 1);
    }
    void *args[] = {(void *)&P};
    WRNN_CUDA(h, cudaLaunchCooperativeKernel((const void *)mma_loop_kernel, dim3(P.G), dim3(THREADS), args, pl.smem_bytes, st));
    h->launches++;
    if (pe) {     // debug: print the timeline of the profiled step (clocks relative to the CTA's first event of the step)
        std::vector<long long> pr(128);
        WRNN_CUDA(h, cudaStreamSynchronize(st));
        WRNN_CUDA(h, cudaMemcpy(pr.data(), P.prof, 128 * sizeof(long long), cudaMemcpyDeviceToHost));
        static const char *names[32] = {
            "producer: step begin", "producer: rnn1 issued", "producer: rnn2 issued", "producer: recurrent-1 issued", "producer: fc begin",
            "producer: fc1 issued", "producer: fc2 issued", "producer: recurrent-2 issued", "", "",
            "mma: rnn1 issued", "mma: rnn2 issued", "mma: recurrent-1 issued", "mma: fc1 issued", "mma: fc2 issued", "mma: recurrent-2 issued",
            "epilogue: rnn1 gates done (warp 2)", "epilogue: rnn2 gates done (warp 2)", "epilogue: fc1 done (warp 2)", "epilogue: fc2 done (warp 2)",
            "epilogue: rnn1 accumulators ready", "epilogue: rnn1 published", "epilogue: rnn2 accumulators ready", "epilogue: rnn2 published",
            "epilogue: fc1 accumulator ready", "epilogue: fc1 published", "epilogue: fc2 accumulator ready", "epilogue: fc2 published",
            "fc3 phase: fc2 of all CTAs seen", "fc3 phase: folds done", "fc3 phase: published", ""};
        for (int c = 0; c < 2; c++) {
            const long long *q = pr.data() + 64 * c;
            long long t0 = 0;
            for (int i = 0; i < 32; i++)
                if (q[i] && (!t0 || q[i] < t0)) t0 = q[i];
            fprintf(stderr, "[mma prof] CTA %s, step %lld, folds %lld, fold tile %d\n", c == 0 ? "0" : "G-1", (long long)P.prof_step,
                    (long long)a.B, P.NT);
            std::vector<std::pair<long long, int>> ev;
            for (int i = 0; i < 32; i++)
                if (q[i]) ev.push_back({q[i] - t0, i});
            std::sort(ev.begin(), ev.end());
            for (auto &e : ev) fprintf(stderr, "[mma prof]   %8lld  %s\n", e.first, names[e.second]);
            static const char *jn[6] = {"rnn1", "rnn2", "recurrent-1", "recurrent-2", "fc1", "fc2"};
            for (int jt = 0; jt < 6; jt++)
                if (q[40 + 3 * jt + 1])
                    fprintf(stderr, "[mma prof]   %-12s %2d stages: MMA thread full-wait %6lld, issue %6lld, commit %5lld; producer lane 0 empty-wait %6lld\n",
                            jn[jt], job_stages(jt), q[40 + 3 * jt], q[40 + 3 * jt + 1], q[40 + 3 * jt + 2], q[32 + jt]);
        }
    }
    return 0;
}

}  // namespace wrnn

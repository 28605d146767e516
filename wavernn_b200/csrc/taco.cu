// Tacotron decoder loop of `Tacotron.generate` (models/tacotron.py:438-465: encoder projection, then per decoder step
// PreNet -> attention GRU cell -> location-sensitive smoothed attention -> context -> rnn_input -> two residual LSTM
// cells -> mel projection of r frames, until every frame of a step is below the stop threshold) as ONE persistent
// cooperative kernel: every SM keeps its rows of all decoder matrices (20 MB fp32 in total) in shared memory for the
// whole utterance, the layers hand small vectors to one another through L2 and grid barriers.
//
// Seven dependent stages per decoder step instead of one per layer, by algebra on the host (compositions formed in double)
// and in the prologue (per-position projections), handed over as tagged 8-byte words polled by the consumers:
//   A  mel frames of the step before AND prenet.fc1 of this step: fc1(mel[:, r-1]) = (W_fc1 . W_mel[last frame]) . x3 + b
//   B  prenet.fc2 (+ ReLU)
//   C  attention GRU cell (input [context | prenet], hidden)                                             (:236-238)
//   D  processed query W . h_attn + b                                                                     (:190)
//   E  energies v . tanh(query + enc_proj + loc) and sigmoid for the positions a CTA owns                (:196-205)
//   F  rnn_input and LSTM 1 at once: gates = (W_ih1 . W_rnn_input) . [context | h_attn] + W_hh1 . h1; x rows beside (:256-262)
//   G  LSTM 2 on x + h1; publishes x3 = x + h1 + h2 for A                                                (:270-277)
// The context vector (:244) is never formed: it enters F's rows and C's rows only through products with fixed rows, and
// context = sum_l a_l enc[l], so W[row] . context = sum_l a_l (W[row][context columns] . enc[l]) with the bracket a constant
// of the utterance, built once per call.  E therefore hands over L attention weights, not 256-wide partial contexts, and
// the step's one grid barrier carries no data (it only bounds the skew between CTAs; its counter is read off the chain).
// Dropout and zoneout are training-only (eval mode in generate, :423).
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <map>
#include <string>
#include <vector>

#include "internal.h"
#include "loop_common.cuh"

namespace wrnn {

constexpr int TC_THREADS = 512, TC_WARPS = TC_THREADS / 32;
constexpr int TC_FILT = 32, TC_KS = 31, TC_HALO = (TC_KS - 1) / 2, TC_MAXR = 20;

struct TacoParams {
    int D, H, M, P1, P2, r, L, max_iter, cap;      // cap = max_iter * r columns of mel_out
    float thr;
    int NA, rpa, rpb, upc, rpd, upl;               // rows (A, B, D) / units (C: GRU, F/G: LSTM) per CTA
    int GA;                                        // CTAs that share the A rows (the position owners are kept out of it)
    int PP, pc0, PC;                               // positions per owner CTA, first owner CTA, owners in use
    int img_floats;
    int off_A, off_bA, off_B, off_bB, off_C, off_bC, off_D, off_bD, off_F, off_bF, off_G, off_bG, off_E;
    int sm_vec, sm_misc;                           // smem float offsets behind the image
    int Lp, sm_pf, sm_pc;                          // padded text length; per-position projections of the context columns (below)
    const float *image;                            // [G][img_floats]
    const float *common;                           // conv w [32][2][31] (+pad to 2048) | v [D] | L.weight [D][32] | L.bias [D]
    const float *enc, *encp;                       // [L][D]
    uint2 *p1, *p2, *ha, *pq, *x2, *h1, *h2, *x3, *att_g, *cum_g;      // tagged words {value, 1 + producing step}
    uint2 *melf;                                   // frames of the last step, tagged like the rest (read by the stop rule)
    uint2 *s_g;                                    // [L] sigmoid(u_l) of this step, tagged
    float *mel_out, *attn_out;
    int *n_iter;
    int d_first, d_poll;                           // clocks before the first look at a fresh vector / between looks
    long long *prof;                               // debug (WRNN_TACO_PROF): [2 CTAs][32] clocks per site, summed over steps
    GridSync gs;
};

__device__ __forceinline__ float warp_dot(const float *w, const float *x, int k4, int lane) {
    float a0 = 0.f, a1 = 0.f;
    const float4 *w4 = reinterpret_cast<const float4 *>(w), *x4 = reinterpret_cast<const float4 *>(x);
    for (int i0 = lane; i0 < k4; i0 += 128) {                   // four float4 pairs in flight per lane
        float4 a[4], b[4];
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const int i = i0 + 32 * q;
            a[q] = i < k4 ? w4[i] : make_float4(0.f, 0.f, 0.f, 0.f);
            b[q] = i < k4 ? x4[i] : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int q = 0; q < 4; q++) {
            a0 = fmaf(a[q].x, b[q].x, a0);
            a1 = fmaf(a[q].y, b[q].y, a1);
            a0 = fmaf(a[q].z, b[q].z, a0);
            a1 = fmaf(a[q].w, b[q].w, a1);
        }
    }
    a0 += a1;
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) a0 += __shfl_xor_sync(0xffffffffu, a0, o);
    return a0;
}

__device__ __forceinline__ uint2 ld_word(const uint2 *p) {
    // ONE 64-bit scalar access: single-copy atomic in the PTX memory model (a .v2.u32 access is modelled as two), so a
    // word never shows a new tag beside an old value
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return make_uint2((unsigned)v, (unsigned)(v >> 32));
}
__device__ __forceinline__ void st_word(uint2 *p, float value, unsigned tag) {
    const unsigned long long v = (unsigned long long)__float_as_uint(value) | ((unsigned long long)tag << 32);
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// the value of a tagged word once its producer has stored it; a watchdog turns a lost hand-over into the abort flag
__device__ __forceinline__ void spin_clocks(int n) {
    if (n > 0) {
        const long long t0 = clock64();
        while (clock64() - t0 < n) {}
    }
}
__device__ __forceinline__ float wait_word(const uint2 *p, unsigned tag, int *abort_flag, int d_poll = 0) {
    uint2 w = ld_word(p);
    if (w.y != tag) {
        const long long t0 = clock64();
        unsigned spins = 0;
        do {
            spin_clocks(d_poll);
            w = ld_word(p);
            if ((++spins & 255u) == 0u && (clock64() - t0 > 4000000000ll || ld_volatile_s32(abort_flag))) {       // ~2 s
                atomicExch(abort_flag, 1);
                break;
            }
        } while (w.y != tag);
    }
    return __uint_as_float(w.x);
}

// the grid barrier in two halves: a CTA arrives as soon as it has nothing more to publish and waits after the work that
// needs nothing from the others
// `release`: the CTA has plain stores the others read after the barrier (the position owners' partial sums).  Everything
// else travels as self-validating tagged words, so a CTA without such stores arrives with a relaxed add and spares itself
// the fence (measured: ~2 k clocks of thread 0, which the whole CTA then waits for at its next __syncthreads)
__device__ __forceinline__ void grid_arrive(const GridSync &gs, unsigned long long &gen, bool release) {
    __syncthreads();
    gen++;
    if (threadIdx.x == 0) {
        if (release) red_release_add_u64(gs.counter, 1ull);
        else asm volatile("red.relaxed.gpu.global.add.u64 [%0], %1;" ::"l"(gs.counter), "l"(1ull) : "memory");
    }
}
__device__ __forceinline__ void grid_wait(const GridSync &gs, unsigned long long gen) {
    if (threadIdx.x == 0) {
        const unsigned long long want = gen * gridDim.x;
        const long long t0 = clock64();
        unsigned spins = 0;
        while (ld_acquire_u64(gs.counter) < want) {
            if ((++spins & 63u) == 0u && (clock64() - t0 > 4000000000ll || ld_volatile_s32(gs.abort_flag))) {   // ~2 s
                atomicExch(gs.abort_flag, 1);
                break;
            }
        }
    }
    __syncthreads();
}

__global__ void __launch_bounds__(TC_THREADS, 1) taco_loop_kernel(const TacoParams P) {
    extern __shared__ __align__(16) float smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, cta = blockIdx.x;
    const int D = P.D, H = P.H, r = P.r, L = P.L;
    float *img = smem, *vec = smem + P.sm_vec, *misc = smem + P.sm_misc;
    float *PF = smem + P.sm_pf, *PCt = smem + P.sm_pc;      // [5 upl][Lp], [3 upc][Lp]: W[row][context columns] . enc[l]
    float *gsum = misc;                 // [<= 32] dot products of a stage
    float *c1 = misc + 32, *c2 = c1 + P.upl;            // LSTM cell states of the owned units
    for (int i = tid; i < P.img_floats / 4; i += TC_THREADS)
        reinterpret_cast<float4 *>(img)[i] = reinterpret_cast<const float4 *>(P.image + (size_t)cta * P.img_floats)[i];
    __syncthreads();                    // the owners' region below overlays image space their CTAs do not use

    // ---- attention positions this CTA owns
    const int pi = cta - P.pc0;
    const bool owner = pi >= 0 && pi < P.PC;
    const int l0 = owner ? pi * P.PP : 0, nl = owner ? max(0, min(P.PP, L - l0)) : 0;
    float *encp_s = smem + P.off_E, *ploc = encp_s + (size_t)P.PP * D;
    float *cw = ploc + (size_t)P.PP * D;                 // [32][2][31] padded to 2048
    float *loc = cw + 2048;                              // [2][PP + 30]
    const int LS = (P.PP + 2 * TC_HALO + 3) & ~3;        // row stride of loc: what follows stays 16-byte aligned
    float *conv = loc + 2 * LS;                          // [PP][32]
    float *sv = conv + TC_FILT * P.PP;                   // [PP] sigmoid(u)
    float *cumv = sv + P.PP;                             // [PP] cumulative attention
    float *red = cumv + P.PP;                            // [PP][8]
    float *wsum_s = misc + 32 + 2 * P.upl;               // [16] per-warp sums of sigmoid(u)
    const int a = tid & 255, half = tid >> 8;            // energy thread = (attention dim a, position parity); D == 256
    float lw[TC_FILT], lb = 0.f, va = 0.f;
    if (owner && nl > 0) {
        for (int i = tid; i < nl * D; i += TC_THREADS) encp_s[i] = P.encp[(size_t)l0 * D + i];
        for (int i = tid; i < TC_FILT * 2 * TC_KS; i += TC_THREADS) cw[i] = P.common[i];
        for (int i = tid; i < P.PP; i += TC_THREADS) cumv[i] = 0.f;
        va = P.common[2048 + a];
#pragma unroll
        for (int f = 0; f < TC_FILT; f++) lw[f] = P.common[2048 + D + a * TC_FILT + f];
        lb = P.common[2048 + D + D * TC_FILT + a];
        for (int p = half; p < nl; p += 2) ploc[(size_t)p * D + a] = lb;      // step 0: cumulative = attention = 0
    }
    for (int i = tid; i < 2 * P.upl; i += TC_THREADS) c1[i] = 0.f;
    __syncthreads();

    const float *WA = img + P.off_A, *bA = img + P.off_bA, *WB = img + P.off_B, *bB = img + P.off_bB;
    const float *WC = img + P.off_C, *bC = img + P.off_bC, *WD = img + P.off_D, *bD = img + P.off_bD;
    const float *WF = img + P.off_F, *bF = img + P.off_bF, *WG = img + P.off_G, *bG = img + P.off_bG;
    const int a0 = cta * P.rpa, na = cta < P.GA ? max(0, min(P.rpa, P.NA - a0)) : 0;
    const int b0 = cta * P.rpb, nb = max(0, min(P.rpb, P.P2 - b0));
    const int g0 = cta * P.upc, ng = max(0, min(P.upc, D - g0));
    const int d0 = cta * P.rpd, nd = max(0, min(P.rpd, D - d0));
    const int u0 = cta * P.upl, nu = max(0, min(P.upl, H - u0));
    const int KC = 2 * D + P.P2;        // GRU: [context | prenet | hidden]
    const int KF = 2 * D + H;           // LSTM 1: [context | h_attn | h1]
    const int nmel = P.M * r;
    // The context vector is never formed.  It enters LSTM 1, the x rows and the attention GRU only through products with
    // fixed rows, and context = sum_l a_l enc[l], so W[row] . context = sum_l a_l (W[row][context columns] . enc[l]): the
    // bracket is a per-utterance constant.  Built here once per call (26 rows x L dots of 256), it turns the hand-over
    // "256-wide partial contexts of 20 owners, behind a grid barrier" into "L attention weights as tagged words".
    const int Lp = P.Lp, unitC_ = 2 * KC + (D + P.P2) + D;
    for (int i = warp; i < (5 * nu + 3 * ng) * L; i += TC_WARPS) {
        const int row = i / L, l = i - row * L;
        const float *wrow;
        float *dst;
        if (row < 4 * nu) { wrow = WF + (size_t)row * KF; dst = PF + (size_t)row * Lp; }
        else if (row < 5 * nu) { wrow = WF + (size_t)4 * P.upl * KF + (size_t)(row - 4 * nu) * 2 * D; dst = PF + (size_t)row * Lp; }
        else {
            const int g = row - 5 * nu, j = g / 3, q = g - j * 3;
            wrow = WC + (size_t)j * unitC_ + (q < 2 ? (size_t)q * KC : (size_t)2 * KC);
            dst = PCt + (size_t)g * Lp;
        }
        const float d = warp_dot(wrow, P.enc + (size_t)l * D, D >> 2, lane);
        if (lane == 0) dst[l] = d;
    }
    for (int i = tid; i < 5 * P.upl * (Lp - L); i += TC_THREADS) PF[(size_t)(i / (Lp - L)) * Lp + L + i % (Lp - L)] = 0.f;
    for (int i = tid; i < 3 * P.upc * (Lp - L); i += TC_THREADS) PCt[(size_t)(i / (Lp - L)) * Lp + L + i % (Lp - L)] = 0.f;
    float *gctx = wsum_s + 16;          // [3 upc] context part of the GRU rows (r, z, n_i) for the NEXT step
    for (int i = tid; i < 3 * P.upc; i += TC_THREADS) gctx[i] = 0.f;
    __syncthreads();
    unsigned long long gen = 0;
    int n_done = 0;
    __shared__ long long prof_acc[32];
    long long prof_last = 0;
    const bool prof_on = P.prof != nullptr && tid == 0 && (cta == 0 || cta == P.pc0);
    if (tid < 32) prof_acc[tid] = 0;
#define TMARK(i)                                      \
    if (prof_on) {                                    \
        const long long c__ = clock64();              \
        prof_acc[i] += c__ - prof_last;               \
        prof_last = c__;                              \
    }

    // Hand-over between stages: every exchanged value travels as an 8-byte word {value, tag} written with one store;
    // tag = 1 + the decoder step that produced it.  A consumer thread polls the words it stages itself, so a stage
    // starts one L2 round trip after its last producer stored: no fence, no counter, no second trip for the data.
    // One full grid barrier per step remains (E -> F: the partial contexts are too many words to poll); it is also what
    // makes single buffers safe (every buffer is rewritten only after a barrier that all of its readers have passed).
    auto fetch = [&](const uint2 *src, int n, unsigned tag, float *dst, bool delay = true) {
        if (n > 0 && delay) spin_clocks(P.d_first);
        for (int i = tid; i < n; i += TC_THREADS) dst[i] = wait_word(src + i, tag, P.gs.abort_flag, P.d_poll);
    };
    // activation of LSTM gate row w = unit * 4 + gate, applied by the warp that formed the row (gates i, f, g, o)
    auto lstm_act = [&](int w, float x) { return (w & 3) == 2 ? tanhf(x) : sigmoidf_acc(x); };
    float *vec2 = misc + ((32 + 2 * P.upl + 160 + D + 3) & ~3);      // [H] second staging vector (the products formed ahead)
    float *hh1s = vec2 + H, *hh2s = hh1s + 4 * P.upl;      // W_hh1 . h1 and W_hh2 . h2 of the owned gate rows
    float *fha = hh2s + 4 * P.upl;                         // [5 upl] h_attn columns of the LSTM 1 gate rows and of the x rows
    float *ghh = fha + 5 * P.upl;                          // [4 upc] W_hh . h_attn of the GRU rows (r, z, -, n) for the NEXT step
    float *hown = ghh + 4 * P.upc;                         // [upc] h_attn of the owned GRU units
    for (int i = tid; i < 5 * P.upc; i += TC_THREADS) ghh[i] = 0.f;
    // up to three vectors of <= 512 words staged in ONE L2 round trip: first looks issued together, stragglers polled
    auto fetch3 = [&](const uint2 *s1, int n1, unsigned t1, float *d1, const uint2 *s2, int n2, unsigned t2, float *d2,
                      const uint2 *s3, int n3, unsigned t3, float *d3) {
        uint2 w1 = make_uint2(0u, t1), w2 = make_uint2(0u, t2), w3 = make_uint2(0u, t3);
        if (tid < n1) w1 = ld_word(s1 + tid);
        if (tid < n2) w2 = ld_word(s2 + tid);
        if (tid < n3) w3 = ld_word(s3 + tid);
        if (tid < n1) d1[tid] = w1.y == t1 ? __uint_as_float(w1.x) : wait_word(s1 + tid, t1, P.gs.abort_flag, P.d_poll);
        if (tid < n2) d2[tid] = w2.y == t2 ? __uint_as_float(w2.x) : wait_word(s2 + tid, t2, P.gs.abort_flag, P.d_poll);
        if (tid < n3) d3[tid] = w3.y == t3 ? __uint_as_float(w3.x) : wait_word(s3 + tid, t3, P.gs.abort_flag, P.d_poll);
    };

    for (int it = 0;; it++) {
        const int cur = it & 1, nxt = cur ^ 1;
        const unsigned tag_prev = (unsigned)it, tag_now = (unsigned)it + 1u;
        // ---- A: mel frames of step it-1, prenet.fc1 of step it (x3 = 0 before the first step: fc1(go frame) = relu(b));
        //         position owners finish the location features of this step (L . conv + b; the conv was formed in G)
        if (prof_on) prof_last = clock64();
        fetch(P.x3, H, tag_prev, vec);
        __syncthreads();
        TMARK(0);
        for (int j = warp; j < na; j += TC_WARPS) {
            const float d = warp_dot(WA + (size_t)j * H, vec, H >> 2, lane);
            if (lane == 0) {
                const int row = a0 + j;
                if (row < nmel) {
                    if (it > 0) {
                        const int m = row / r, jj = row - m * r;
                        P.mel_out[(size_t)m * P.cap + (size_t)(it - 1) * r + jj] = d;
                        st_word(P.melf + row, d, tag_now);
                    }
                } else {
                    st_word(P.p1 + (row - nmel), fmaxf(d + bA[j], 0.f), tag_now);
                }
            }
        }
        if (nl > 0 && it > 0) {
            for (int p = half; p < nl; p += 2) {
                float acc = lb;
                const float4 *c4 = reinterpret_cast<const float4 *>(conv + p * TC_FILT);
#pragma unroll
                for (int f4 = 0; f4 < TC_FILT / 4; f4++) {
                    const float4 c = c4[f4];
                    acc = fmaf(lw[4 * f4 + 0], c.x, acc);
                    acc = fmaf(lw[4 * f4 + 1], c.y, acc);
                    acc = fmaf(lw[4 * f4 + 2], c.z, acc);
                    acc = fmaf(lw[4 * f4 + 3], c.w, acc);
                }
                ploc[(size_t)p * D + a] = acc;
            }
        }
        n_done = it;
        if (it == P.max_iter) break;
        __syncthreads();
        TMARK(1);
        // ---- B: prenet.fc2
        fetch(P.p1, nb > 0 ? P.P1 : 0, tag_now, vec);
        __syncthreads();
        TMARK(2);
        for (int j = warp; j < nb; j += TC_WARPS) {
            const float d = warp_dot(WB + (size_t)j * P.P1, vec, P.P1 >> 2, lane);
            if (lane == 0) st_word(P.p2 + b0 + j, fmaxf(d + bB[j], 0.f), tag_now);
        }
        __syncthreads();
        TMARK(3);
        // ---- C: attention GRU cell
        if (ng > 0) fetch(P.p2, P.P2, tag_now, vec);
        __syncthreads();
        TMARK(4);
        for (int w = warp; w < 4 * ng; w += TC_WARPS) {           // hidden-state and context columns were formed ahead (ghh, gctx)
            const int j = w >> 2, q = w & 3;
            const float *row = WC + (size_t)j * (2 * KC + (D + P.P2) + D);
            if (q == 3) {
                if (lane == 0) gsum[w] = ghh[w] + bC[w];
            } else {
                const float d = warp_dot(row + (size_t)q * KC + D, vec, P.P2 >> 2, lane) + gctx[j * 3 + q];      // prenet columns
                if (lane == 0) gsum[w] = q < 2 ? sigmoidf_acc(d + ghh[w] + bC[w]) : d + bC[w];
            }
        }
        __syncthreads();
        if (tid < ng) {
            const float rr = gsum[tid * 4 + 0], zz = gsum[tid * 4 + 1];
            const float nn = tanhf(gsum[tid * 4 + 2] + rr * gsum[tid * 4 + 3]);
            const float hn = (1.f - zz) * nn + zz * hown[tid];
            hown[tid] = hn;
            st_word(P.ha + (size_t)nxt * D + g0 + tid, hn, tag_now);
        }
        __syncthreads();
        TMARK(5);
        // ---- D: processed query
        fetch(P.ha + (size_t)nxt * D, nd > 0 ? D : 0, tag_now, vec);
        __syncthreads();
        TMARK(6);
        for (int j = warp; j < nd; j += TC_WARPS) {
            const float d = warp_dot(WD + (size_t)j * D, vec, D >> 2, lane);
            if (lane == 0) st_word(P.pq + d0 + j, d + bD[j], tag_now);
        }
        __syncthreads();
        TMARK(7);
        // ---- in the shadow of E (position owners only) and of the barrier: every product whose input is known by now.
        //      W_hh1 . h1 and W_hh2 . h2 (states of the step before), the h_attn columns of LSTM 1's composed gate rows
        //      and of the x rows, and W_hh . h_attn of the attention GRU for the NEXT step
        uint2 early_s = make_uint2(0u, 0u);
        auto shadow = [&]() {
            if (nu == 0 && ng == 0) return;
            fetch3(P.h1 + (size_t)cur * H, nu > 0 ? H : 0, tag_prev, vec, P.h2 + (size_t)cur * H, nu > 0 ? H : 0, tag_prev, vec + H,
                   P.ha + (size_t)nxt * D, D, tag_now, vec2);
            __syncthreads();
            const int j1 = 4 * nu, j2 = 8 * nu, j3 = 12 * nu;
            for (int w = warp; w < j3; w += TC_WARPS) {
                // a first look at this step's attention weights goes out ahead of the last round of jobs: F validates it
                if (w + TC_WARPS >= j3 && tid < L) early_s = ld_word(P.s_g + tid);
                if (w < j1) {
                    const float d = warp_dot(WF + (size_t)w * KF + 2 * D, vec, H >> 2, lane);
                    if (lane == 0) hh1s[w] = d;
                } else if (w < j2) {
                    const float d = warp_dot(WG + (size_t)(w - j1) * 2 * H + H, vec + H, H >> 2, lane);
                    if (lane == 0) hh2s[w - j1] = d;
                } else {
                    const float d = warp_dot(WF + (size_t)(w - j2) * KF + D, vec2, D >> 2, lane);
                    if (lane == 0) fha[w - j2] = d;
                }
            }
            __syncthreads();
        };
        if (nl == 0) {
            grid_arrive(P.gs, gen, false);
            shadow();
        }
        // ---- E: energies, sigmoid and the partial context of the owned positions
        if (nl > 0) {
            fetch(P.pq, D, tag_now, vec);
            __syncthreads();
            for (int p0 = half; p0 < nl; p0 += 8) {              // four positions per pass: their shuffle trees overlap
                float v[4];
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    const int p = p0 + 2 * q;
                    v[q] = p < nl ? va * tanhf(vec[a] + encp_s[(size_t)p * D + a] + ploc[(size_t)p * D + a]) : 0.f;
                }
#pragma unroll
                for (int o = 16; o >= 1; o >>= 1) {
#pragma unroll
                    for (int q = 0; q < 4; q++) v[q] += __shfl_xor_sync(0xffffffffu, v[q], o);
                }
                if (lane == 0) {
#pragma unroll
                    for (int q = 0; q < 4; q++)
                        if (p0 + 2 * q < nl) red[(p0 + 2 * q) * 8 + (warp & 7)] = v[q];
                }
            }
            __syncthreads();
            if (tid < nl) {
                float e = 0.f;
                for (int w = 0; w < 8; w++) e += red[tid * 8 + w];
                const float sg = sigmoidf_acc(e);
                sv[tid] = sg;
                st_word(P.s_g + l0 + tid, sg, tag_now);
            }
        }
        if (nl > 0) {
            grid_arrive(P.gs, gen, false);
            shadow();
        }
        TMARK(8);
        // ---- F: the stop rule on the frames of step it-1 (:455: all below the threshold and t > 10); the attention
        //         weights arrive; rnn_input + LSTM 1 through the per-position projections; owners normalise their scores
        float inv_total = 0.f;
        TMARK(9);
        {
            const bool check = it > 0 && (it - 1) * r > 10;
            uint2 mw[4];                                        // first looks now, validated after the other loads are out
#pragma unroll
            for (int q = 0; q < 4; q++)
                mw[q] = (check && tid + q * TC_THREADS < nmel) ? ld_word(P.melf + tid + q * TC_THREADS) : make_uint2(0xff800000u, tag_now);
            // the attention weights before normalisation: one tagged word per position, one look per thread
            float sg = 0.f;
            if (tid < L) sg = early_s.y == tag_now ? __uint_as_float(early_s.x) : wait_word(P.s_g + tid, tag_now, P.gs.abort_flag, P.d_poll);
            if (tid < Lp) vec[tid] = sg;
            // their sum: per warp by the xor tree, then across the warps in order (the same bits in every CTA)
            float wsum = sg;
#pragma unroll
            for (int o = 16; o >= 1; o >>= 1) wsum += __shfl_xor_sync(0xffffffffu, wsum, o);
            if (lane == 0) wsum_s[warp] = wsum;
            int ok = check ? 1 : 0;
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const float mv = mw[q].y == tag_now ? __uint_as_float(mw[q].x)
                                                    : wait_word(P.melf + tid + q * TC_THREADS, tag_now, P.gs.abort_flag);
                ok &= (mv < P.thr) ? 1 : 0;
            }
            for (int i = tid + 4 * TC_THREADS; check && i < nmel; i += TC_THREADS) ok &= (wait_word(P.melf + i, tag_now, P.gs.abort_flag) < P.thr) ? 1 : 0;
            if (__syncthreads_and(ok)) break;
            float total = 0.f;
            for (int w = 0; w < (L + 31) / 32; w++) total += wsum_s[w];
            inv_total = 1.f / total;
            if (tid < nl) {
                const float att = sv[tid] / total;
                const float cu = cumv[tid] + att;
                cumv[tid] = cu;
                st_word(P.att_g + l0 + tid, att, tag_now);
                st_word(P.cum_g + l0 + tid, cu, tag_now);
                P.attn_out[(size_t)it * L + l0 + tid] = att;
            }
        }
        __syncthreads();
        TMARK(10);
        for (int w = warp; w < 5 * nu + 6 * ng; w += TC_WARPS) {
            if (w >= 5 * nu + 3 * ng) {                                         // W_hh . h_attn of the GRU rows (r, z, n) for the next step
                const int g = w - 5 * nu - 3 * ng, j = g / 3, q = g - j * 3;
                const float *row = WC + (size_t)j * (2 * KC + (D + P.P2) + D);
                const float d = warp_dot(q < 2 ? row + (size_t)q * KC + D + P.P2 : row + 2 * KC + D + P.P2, vec2, D >> 2, lane);
                if (lane == 0) ghh[j * 4 + (q < 2 ? q : 3)] = d;
                continue;
            }
            if (w >= 5 * nu) {                                                  // context part of the GRU rows for the next step
                const float d = warp_dot(PCt + (size_t)(w - 5 * nu) * Lp, vec, Lp >> 2, lane) * inv_total;
                if (lane == 0) gctx[w - 5 * nu] = d;
                continue;
            }
            const float d = warp_dot(PF + (size_t)w * Lp, vec, Lp >> 2, lane) * inv_total;       // W[row] . context
            if (w < 4 * nu) {                                                   // w = unit * 4 + gate
                if (lane == 0) gsum[w] = lstm_act(w, d + fha[w] + hh1s[w] + bF[(w >> 2) * 5 + (w & 3)]);
            } else {                                                            // x row: its h_attn columns here too
                const float dh = warp_dot(WF + (size_t)4 * P.upl * KF + (size_t)(w - 4 * nu) * 2 * D + D, vec2, D >> 2, lane);
                if (lane == 0) gsum[w] = d + dh + bF[(w - 4 * nu) * 5 + 4];
            }
        }
        __syncthreads();
        if (tid < nu) {
            const float c = gsum[tid * 4 + 1] * c1[tid] + gsum[tid * 4 + 0] * gsum[tid * 4 + 2], h = gsum[tid * 4 + 3] * tanhf(c);
            c1[tid] = c;
            st_word(P.x2 + u0 + tid, gsum[4 * nu + tid] + h, tag_now);
            st_word(P.h1 + (size_t)nxt * H + u0 + tid, h, tag_now);
        }
        __syncthreads();
        TMARK(11);
        // ---- G: LSTM 2 on x + h1; position owners convolve [cumulative | attention] for the next step
        // the step's grid barrier carries no data any more (every hand-over is self-validating); it only bounds the skew
        // between CTAs.  Its counter is read here, the read in flight while x2 arrives, and is complete by now in practice
        unsigned long long arrived = 0;
        if (tid == 0) arrived = ld_acquire_u64(P.gs.counter);
        fetch(P.x2, nu > 0 ? H : 0, tag_now, vec);
        if (tid == 0 && arrived < gen * gridDim.x) {
            const long long t0 = clock64();
            unsigned spins = 0;
            while (ld_acquire_u64(P.gs.counter) < gen * gridDim.x) {
                if ((++spins & 63u) == 0u && (clock64() - t0 > 4000000000ll || ld_volatile_s32(P.gs.abort_flag))) {
                    atomicExch(P.gs.abort_flag, 1);
                    break;
                }
            }
        }
        if (nl > 0) {
            const int span = nl + 2 * TC_HALO;
            for (int i = tid; i < 2 * span; i += TC_THREADS) {
                const int c = i / span, k = i - c * span, l = l0 - TC_HALO + k;
                loc[c * LS + k] = (l >= 0 && l < L) ? wait_word((c == 0 ? P.cum_g : P.att_g) + l, tag_now, P.gs.abort_flag) : 0.f;
            }
        }
        __syncthreads();
        TMARK(12);
        for (int w = warp; w < 4 * nu; w += TC_WARPS) {
            const float d = warp_dot(WG + (size_t)w * 2 * H, vec, H >> 2, lane);
            if (lane == 0) gsum[w] = lstm_act(w, d + hh2s[w] + bG[w]);
        }
        if (nl > 0) {
            for (int i = tid; i < TC_FILT * nl; i += TC_THREADS) {
                const int p = i / TC_FILT, f = i - p * TC_FILT;
                float acc = 0.f;
#pragma unroll
                for (int c = 0; c < 2; c++) {
                    const float *w = cw + (f * 2 + c) * TC_KS, *x = loc + c * LS + p;
#pragma unroll
                    for (int k = 0; k < TC_KS; k++) acc = fmaf(w[k], x[k], acc);
                }
                conv[p * TC_FILT + f] = acc;
            }
        }
        __syncthreads();
        if (tid < nu) {
            const float c = gsum[tid * 4 + 1] * c2[tid] + gsum[tid * 4 + 0] * gsum[tid * 4 + 2], h = gsum[tid * 4 + 3] * tanhf(c);
            c2[tid] = c;
            st_word(P.x3 + u0 + tid, vec[u0 + tid] + h, tag_now);
            st_word(P.h2 + (size_t)nxt * H + u0 + tid, h, tag_now);
        }
        __syncthreads();
        TMARK(13);
    }
    if (prof_on)
        for (int i = 0; i < 32; i++) P.prof[(cta == 0 ? 0 : 32) + i] = prof_acc[i];
    if (cta == 0 && tid == 0) *P.n_iter = n_done;
}

// enc_proj = enc_seq . W^T (models/tacotron.py:440, `encoder_proj`, no bias): one CTA per position
__global__ void taco_encproj_kernel(const float *enc, const float *Wt /*[D in][D out]*/, float *encp, int D) {
    extern __shared__ float row[];
    const int l = blockIdx.x;
    for (int i = threadIdx.x; i < D; i += blockDim.x) row[i] = enc[(size_t)l * D + i];
    __syncthreads();
    for (int o = threadIdx.x; o < D; o += blockDim.x) {
        float acc = 0.f;
        for (int k = 0; k < D; k++) acc = fmaf(row[k], Wt[(size_t)k * D + o], acc);
        encp[(size_t)l * D + o] = acc;
    }
}

}  // namespace wrnn

// ============================================================================ host
struct wrnn_taco_handle {
    int D = 0, H = 0, M = 0, device = 0, num_sms = 0;
    size_t smem_optin = 0;
    std::string err;
    std::map<std::string, wrnn::HostTensor> w;
    int packed_r = 0;            // the reduction factor the image was packed for (0: none)
    wrnn::TacoParams plan;       // layout fields only
    float *image = nullptr, *common = nullptr, *projT = nullptr;
    std::vector<double> WcF;     // (W_ih1 . W_rnn_input) [4H][2D] and its bias [4H]: independent of r, formed once
    std::vector<double> bcF;
    void *ws = nullptr;
    size_t ws_bytes = 0;
    cudaEvent_t ev[2] = {nullptr, nullptr};
    int64_t launches = 0;
};

static thread_local std::string g_taco_err;

namespace wrnn {

static int taco_fail(wrnn_taco_handle *h, int code, const char *msg) {
    if (h) h->err = msg;
    else g_taco_err = msg;
    if (code == WRNN_ERR_CUDA) cudaGetLastError();
    return code;
}
#define TC_CUDA(h, expr)                                                                          \
    do {                                                                                          \
        cudaError_t e__ = (expr);                                                                 \
        if (e__ != cudaSuccess) {                                                                 \
            char b__[512];                                                                        \
            snprintf(b__, sizeof b__, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__); \
            return wrnn::taco_fail((h), WRNN_ERR_CUDA, b__);                                      \
        }                                                                                         \
    } while (0)

static int taco_pack(wrnn_taco_handle *h, int r) {
    const int D = h->D, H = h->H, M = h->M, G = h->num_sms, P1 = 256, P2 = 128;
    auto get = [&](const char *k, std::vector<int64_t> shape) -> const HostTensor * {
        auto it = h->w.find(k);
        if (it == h->w.end() || it->second.shape != shape) {
            h->err = std::string("checkpoint key missing or misshapen: ") + k;
            return nullptr;
        }
        return &it->second;
    };
    const HostTensor *Wp = get("encoder_proj.weight", {D, D}),
                     *W1 = get("decoder.prenet.fc1.weight", {P1, M}), *b1 = get("decoder.prenet.fc1.bias", {P1}),
                     *W2 = get("decoder.prenet.fc2.weight", {P2, P1}), *b2 = get("decoder.prenet.fc2.bias", {P2}),
                     *Cw = get("decoder.attn_net.conv.weight", {TC_FILT, 2, TC_KS}),
                     *Lw = get("decoder.attn_net.L.weight", {D, TC_FILT}), *Lb = get("decoder.attn_net.L.bias", {D}),
                     *Wq = get("decoder.attn_net.W.weight", {D, D}), *bq = get("decoder.attn_net.W.bias", {D}),
                     *Vw = get("decoder.attn_net.v.weight", {1, D}),
                     *Gi = get("decoder.attn_rnn.weight_ih", {3 * D, D + P2}), *Gh = get("decoder.attn_rnn.weight_hh", {3 * D, D}),
                     *Gbi = get("decoder.attn_rnn.bias_ih", {3 * D}), *Gbh = get("decoder.attn_rnn.bias_hh", {3 * D}),
                     *Wr = get("decoder.rnn_input.weight", {H, 2 * D}), *br = get("decoder.rnn_input.bias", {H}),
                     *Ai = get("decoder.res_rnn1.weight_ih", {4 * H, H}), *Ah = get("decoder.res_rnn1.weight_hh", {4 * H, H}),
                     *Abi = get("decoder.res_rnn1.bias_ih", {4 * H}), *Abh = get("decoder.res_rnn1.bias_hh", {4 * H}),
                     *Bi = get("decoder.res_rnn2.weight_ih", {4 * H, H}), *Bh = get("decoder.res_rnn2.weight_hh", {4 * H, H}),
                     *Bbi = get("decoder.res_rnn2.bias_ih", {4 * H}), *Bbh = get("decoder.res_rnn2.bias_hh", {4 * H}),
                     *Wm = get("decoder.mel_proj.weight", {M * TC_MAXR, H});
    if (!Wp || !W1 || !b1 || !W2 || !b2 || !Cw || !Lw || !Lb || !Wq || !bq || !Vw || !Gi || !Gh || !Gbi || !Gbh || !Wr || !br ||
        !Ai || !Ah || !Abi || !Abh || !Bi || !Bh || !Bbi || !Bbh || !Wm)
        return WRNN_ERR_STATE;
    TacoParams &p = h->plan;
    p.D = D; p.H = H; p.M = M; p.P1 = P1; p.P2 = P2; p.r = r;
    p.NA = M * r + P1;
    p.rpb = (P2 + G - 1) / G;
    p.upc = (D + G - 1) / G;
    p.rpd = (D + G - 1) / G;
    p.upl = (H + G - 1) / G;
    {
        const int n_lstm = (H + p.upl - 1) / p.upl;
        p.GA = G - n_lstm >= 8 ? n_lstm : G;           // the same split as the position owners' in wrnn_taco_decode
    }
    p.rpa = (p.NA + p.GA - 1) / p.GA;
    const int KC = 2 * D + P2, KF = 2 * D + H, unitC = 2 * KC + (D + P2) + D;
    auto r4 = [](int x) { return (x + 3) & ~3; };
    int o = 0;
    p.off_A = o;  o += p.rpa * H;
    p.off_bA = o; o += r4(p.rpa);
    p.off_B = o;  o += p.rpb * P1;
    p.off_bB = o; o += r4(p.rpb);
    p.off_C = o;  o += p.upc * unitC;
    p.off_bC = o; o += r4(p.upc * 4);
    p.off_D = o;  o += p.rpd * D;
    p.off_bD = o; o += r4(p.rpd);
    p.off_F = o;  o += p.upl * (4 * KF + 2 * D);
    p.off_bF = o; o += r4(p.upl * 5);
    p.off_G = o;  o += p.upl * 4 * 2 * H;
    p.off_bG = o; o += r4(p.upl * 4);
    p.img_floats = r4(o);
    // (W_ih1 . W_rnn_input) and its bias, in double (once per checkpoint)
    if (h->WcF.empty()) {
        h->WcF.assign((size_t)4 * H * 2 * D, 0.0);
        h->bcF.assign((size_t)4 * H, 0.0);
        for (int g = 0; g < 4 * H; g++) {
            double *dst = h->WcF.data() + (size_t)g * 2 * D;
            double bacc = (double)Abi->v[g] + (double)Abh->v[g];
            for (int j = 0; j < H; j++) {
                const double wij = Ai->v[(size_t)g * H + j];
                const float *src = Wr->v.data() + (size_t)j * 2 * D;
                for (int k = 0; k < 2 * D; k++) dst[k] += wij * (double)src[k];
                bacc += wij * (double)br->v[j];
            }
            h->bcF[g] = bacc;
        }
    }
    // prenet.fc1 composed with the mel projection of the last frame of a step: [P1][H]
    std::vector<double> W1c((size_t)P1 * H, 0.0);
    for (int k = 0; k < P1; k++)
        for (int m = 0; m < M; m++) {
            const double w = W1->v[(size_t)k * M + m];
            const float *src = Wm->v.data() + (size_t)(m * TC_MAXR + r - 1) * H;
            double *dst = W1c.data() + (size_t)k * H;
            for (int x = 0; x < H; x++) dst[x] += w * (double)src[x];
        }
    std::vector<float> img((size_t)G * p.img_floats, 0.f);
    for (int c = 0; c < G; c++) {
        float *im = img.data() + (size_t)c * p.img_floats;
        for (int j = 0; j < p.rpa; j++) {
            const int row = c * p.rpa + j;
            if (row >= p.NA || c >= p.GA) break;
            float *dst = im + p.off_A + (size_t)j * H;
            if (row < M * r) {
                const int m = row / r, jj = row - m * r;
                memcpy(dst, Wm->v.data() + (size_t)(m * TC_MAXR + jj) * H, (size_t)H * sizeof(float));
            } else {
                const int k = row - M * r;
                for (int x = 0; x < H; x++) dst[x] = (float)W1c[(size_t)k * H + x];
                im[p.off_bA + j] = b1->v[k];
            }
        }
        for (int j = 0; j < p.rpb; j++) {
            const int row = c * p.rpb + j;
            if (row >= P2) break;
            memcpy(im + p.off_B + (size_t)j * P1, W2->v.data() + (size_t)row * P1, (size_t)P1 * sizeof(float));
            im[p.off_bB + j] = b2->v[row];
        }
        for (int j = 0; j < p.upc; j++) {                   // GRU gate order r, z, n (torch.nn.GRUCell)
            const int u = c * p.upc + j;
            if (u >= D) break;
            float *dst = im + p.off_C + (size_t)j * unitC;
            for (int g = 0; g < 2; g++) {
                memcpy(dst + (size_t)g * KC, Gi->v.data() + (size_t)(g * D + u) * (D + P2), (size_t)(D + P2) * sizeof(float));
                memcpy(dst + (size_t)g * KC + D + P2, Gh->v.data() + (size_t)(g * D + u) * D, (size_t)D * sizeof(float));
                im[p.off_bC + j * 4 + g] = (float)((double)Gbi->v[g * D + u] + (double)Gbh->v[g * D + u]);
            }
            memcpy(dst + 2 * KC, Gi->v.data() + (size_t)(2 * D + u) * (D + P2), (size_t)(D + P2) * sizeof(float));
            memcpy(dst + 2 * KC + D + P2, Gh->v.data() + (size_t)(2 * D + u) * D, (size_t)D * sizeof(float));
            im[p.off_bC + j * 4 + 2] = Gbi->v[2 * D + u];
            im[p.off_bC + j * 4 + 3] = Gbh->v[2 * D + u];
        }
        for (int j = 0; j < p.rpd; j++) {
            const int row = c * p.rpd + j;
            if (row >= D) break;
            memcpy(im + p.off_D + (size_t)j * D, Wq->v.data() + (size_t)row * D, (size_t)D * sizeof(float));
            im[p.off_bD + j] = bq->v[row];
        }
        for (int j = 0; j < p.upl; j++) {                   // LSTM gate order i, f, g, o (torch.nn.LSTMCell)
            const int u = c * p.upl + j;
            if (u >= H) break;
            for (int g = 0; g < 4; g++) {
                float *dst = im + p.off_F + (size_t)(j * 4 + g) * KF;
                const double *wc = h->WcF.data() + (size_t)(g * H + u) * 2 * D;
                for (int k = 0; k < 2 * D; k++) dst[k] = (float)wc[k];
                memcpy(dst + 2 * D, Ah->v.data() + (size_t)(g * H + u) * H, (size_t)H * sizeof(float));
                im[p.off_bF + j * 5 + g] = (float)h->bcF[g * H + u];
                float *dg = im + p.off_G + (size_t)(j * 4 + g) * 2 * H;
                memcpy(dg, Bi->v.data() + (size_t)(g * H + u) * H, (size_t)H * sizeof(float));
                memcpy(dg + H, Bh->v.data() + (size_t)(g * H + u) * H, (size_t)H * sizeof(float));
                im[p.off_bG + j * 4 + g] = (float)((double)Bbi->v[g * H + u] + (double)Bbh->v[g * H + u]);
            }
            memcpy(im + p.off_F + (size_t)4 * p.upl * KF + (size_t)j * 2 * D, Wr->v.data() + (size_t)u * 2 * D, (size_t)2 * D * sizeof(float));
            im[p.off_bF + j * 5 + 4] = br->v[u];
        }
    }
    std::vector<float> common((size_t)2048 + D + (size_t)D * TC_FILT + D, 0.f);
    memcpy(common.data(), Cw->v.data(), (size_t)TC_FILT * 2 * TC_KS * sizeof(float));
    memcpy(common.data() + 2048, Vw->v.data(), (size_t)D * sizeof(float));
    memcpy(common.data() + 2048 + D, Lw->v.data(), (size_t)D * TC_FILT * sizeof(float));
    memcpy(common.data() + 2048 + D + (size_t)D * TC_FILT, Lb->v.data(), (size_t)D * sizeof(float));
    std::vector<float> projT((size_t)D * D);
    for (int oo = 0; oo < D; oo++)
        for (int k = 0; k < D; k++) projT[(size_t)k * D + oo] = Wp->v[(size_t)oo * D + k];
    for (float **q : {&h->image, &h->common, &h->projT}) {
        if (*q) cudaFree(*q);
        *q = nullptr;
    }
    TC_CUDA(h, cudaMalloc((void **)&h->image, img.size() * sizeof(float)));
    TC_CUDA(h, cudaMemcpy(h->image, img.data(), img.size() * sizeof(float), cudaMemcpyHostToDevice));
    TC_CUDA(h, cudaMalloc((void **)&h->common, common.size() * sizeof(float)));
    TC_CUDA(h, cudaMemcpy(h->common, common.data(), common.size() * sizeof(float), cudaMemcpyHostToDevice));
    TC_CUDA(h, cudaMalloc((void **)&h->projT, projT.size() * sizeof(float)));
    TC_CUDA(h, cudaMemcpy(h->projT, projT.data(), projT.size() * sizeof(float), cudaMemcpyHostToDevice));
    h->packed_r = r;
    return 0;
}

}  // namespace wrnn

using namespace wrnn;

extern "C" {

int wrnn_taco_create(int32_t decoder_dims, int32_t lstm_dims, int32_t n_mels, int device, wrnn_taco_handle **out) {
    if (!out) return taco_fail(nullptr, WRNN_ERR_ARG, "null argument");
    *out = nullptr;
    // PreNet's widths are fixed at 256 / 128 (models/tacotron.py:137) and the attention GRU's input is
    // decoder_dims + decoder_dims // 2 = context + prenet (:218), so the reference itself only runs at decoder_dims 256
    if (decoder_dims != 256) return taco_fail(nullptr, WRNN_ERR_ARG, "decoder_dims must be 256 (PreNet's output is 128 = decoder_dims // 2)");
    if (lstm_dims < 8 || lstm_dims % 4 || lstm_dims > 512 || n_mels < 1 || n_mels > 128)
        return taco_fail(nullptr, WRNN_ERR_ARG, "lstm_dims must be a multiple of 4 up to 512, n_mels 1 .. 128");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) return taco_fail(nullptr, WRNN_ERR_CUDA, "no CUDA device: this library has no CPU fallback");
    if (device < 0 || device >= ndev) return taco_fail(nullptr, WRNN_ERR_ARG, "device out of range");
    cudaDeviceProp prop;
    wrnn::DeviceGuard guard(device);
    if (guard.err != cudaSuccess || cudaGetDeviceProperties(&prop, device) != cudaSuccess)
        return taco_fail(nullptr, WRNN_ERR_CUDA, "cannot open device");
    wrnn_taco_handle *h = new wrnn_taco_handle();
    h->D = decoder_dims; h->H = lstm_dims; h->M = n_mels; h->device = device;
    h->num_sms = prop.multiProcessorCount;
    h->smem_optin = prop.sharedMemPerBlockOptin;
    cudaEventCreate(&h->ev[0]);
    cudaEventCreate(&h->ev[1]);
    *out = h;
    return 0;
}

void wrnn_taco_destroy(wrnn_taco_handle *h) {
    if (!h) return;
    wrnn::DeviceGuard guard(h->device);
    for (float *q : {h->image, h->common, h->projT})
        if (q) cudaFree(q);
    if (h->ws) cudaFree(h->ws);
    for (int i = 0; i < 2; i++)
        if (h->ev[i]) cudaEventDestroy(h->ev[i]);
    delete h;
}

const char *wrnn_taco_last_error(const wrnn_taco_handle *h) { return h ? h->err.c_str() : g_taco_err.c_str(); }

int wrnn_taco_load_weights(wrnn_taco_handle *h, const wrnn_tensor *tensors, int32_t n) {
    if (!h || (!tensors && n > 0)) return taco_fail(h, WRNN_ERR_ARG, "null argument");
    for (int i = 0; i < n; i++) {
        const wrnn_tensor &t = tensors[i];
        if (!t.name || !t.data || t.ndim < 0 || t.ndim > 4) return taco_fail(h, WRNN_ERR_ARG, "bad tensor");
        if (t.dtype != 0) continue;
        HostTensor ht;
        int64_t numel = 1;
        for (int k = 0; k < t.ndim; k++) { ht.shape.push_back(t.shape[k]); numel *= t.shape[k]; }
        ht.v.assign((const float *)t.data, (const float *)t.data + numel);
        h->w[t.name] = std::move(ht);
    }
    h->packed_r = 0;
    h->WcF.clear();
    h->bcF.clear();
    return 0;
}

int wrnn_taco_decode(wrnn_taco_handle *h, const float *enc_seq, int32_t L, int32_t r, int32_t steps, float stop_threshold,
                     float *mel, float *attn, int32_t *n_iter, void *cuda_stream) {
    if (!h || !enc_seq || !mel || !attn || !n_iter) return taco_fail(h, WRNN_ERR_ARG, "null argument");
    if (L < 1 || r < 1 || r > TC_MAXR || steps < 1) return taco_fail(h, WRNN_ERR_ARG, "need L >= 1, 1 <= r <= 20 (Decoder.max_r), steps >= 1");
    cudaStream_t st = (cudaStream_t)cuda_stream;
    wrnn::DeviceGuard guard(h->device);
    TC_CUDA(h, guard.err);
    if (h->packed_r != r) {
        if (h->w.empty()) return taco_fail(h, WRNN_ERR_STATE, "weights not loaded: call wrnn_taco_load_weights first");
        const int rc = taco_pack(h, r);
        if (rc) return rc;
    }
    const int D = h->D, H = h->H, M = h->M, G = h->num_sms;
    TacoParams P = h->plan;
    P.L = L;
    P.max_iter = (steps + r - 1) / r;                    // `for t in range(0, steps, r)` (:446)
    P.cap = P.max_iter * r;
    P.thr = stop_threshold;
    // attention positions go to the CTAs that own no LSTM units when there are at least 8 of them (their F / G image
    // space is free for the position data), else to all CTAs
    const int n_lstm = (H + P.upl - 1) / P.upl;
    const bool disjoint = G - n_lstm >= 8;
    P.pc0 = disjoint ? n_lstm : 0;
    const int owners = disjoint ? G - n_lstm : G;
    P.PP = (L + owners - 1) / owners;
    P.PC = (L + P.PP - 1) / P.PP;
    const int e_floats = 2 * P.PP * D + 2048 + 2 * ((P.PP + 2 * TC_HALO + 3) & ~3) + TC_FILT * P.PP + 2 * P.PP + 8 * P.PP + 8;
    P.off_E = disjoint ? P.off_F : P.img_floats;
    P.Lp = (L + 3) & ~3;
    P.sm_pf = std::max(P.img_floats, (P.off_E + e_floats + 3) & ~3);
    P.sm_pc = P.sm_pf + 5 * P.upl * P.Lp;
    P.sm_vec = P.sm_pc + 3 * P.upc * P.Lp;
    P.sm_misc = P.sm_vec + std::max(std::max(2 * D + H, 2 * H), P.Lp);
    if (L > TC_THREADS) return taco_fail(h, WRNN_ERR_ARG, "more than 512 characters: the attention weights of a step are one word per thread");
    const size_t smem_bytes = ((size_t)P.sm_misc + 32 + 2 * P.upl + 160 + D + H + 13 * P.upl + 5 * P.upc + 12) * sizeof(float);
    if (smem_bytes > h->smem_optin) {
        char b[256];
        snprintf(b, sizeof b, "decoder loop needs %zu bytes of shared memory per SM (r = %d, %d characters), the device allows %zu",
                 smem_bytes, r, L, h->smem_optin);
        return taco_fail(h, WRNN_ERR_ARG, b);
    }
    // workspace: sync[128 B] | state vectors | n_iter | enc, enc_proj | mel_out [M][cap] | attn_out [max_iter][L]
    size_t off = 128;
    auto take = [&](size_t floats) { const size_t o = off; off += ((floats * 4 + 127) & ~(size_t)127); return o; };
    const size_t o_p1 = take(2 * P.P1), o_p2 = take(2 * P.P2), o_ha = take(4 * D), o_pq = take(2 * D), o_sg = take(2 * (size_t)L),
                 o_x2 = take(2 * H), o_h1 = take(4 * H), o_h2 = take(4 * H), o_x3 = take(2 * H),
                 o_melf = take(2 * (size_t)M * TC_MAXR), o_att = take(2 * (size_t)L), o_cum = take(2 * (size_t)L), o_n = take(1);
    const size_t state_end = off;
    const size_t o_enc = take((size_t)L * D), o_encp = take((size_t)L * D), o_mel = take((size_t)M * P.cap),
                 o_attn = take((size_t)P.max_iter * L);
    if (off > h->ws_bytes) {
        if (h->ws) cudaFree(h->ws);
        h->ws = nullptr; h->ws_bytes = 0;
        TC_CUDA(h, cudaMalloc(&h->ws, off + off / 8));
        h->ws_bytes = off + off / 8;
    }
    char *ws = (char *)h->ws;
    auto F = [&](size_t o) { return (float *)(ws + o); };
    TC_CUDA(h, cudaMemsetAsync(ws, 0, state_end, st));
    TC_CUDA(h, cudaMemcpyAsync(ws + o_enc, enc_seq, (size_t)L * D * 4, cudaMemcpyHostToDevice, st));
    P.image = h->image; P.common = h->common; P.enc = F(o_enc); P.encp = F(o_encp);
    auto U = [&](size_t o) { return (uint2 *)(ws + o); };
    P.p1 = U(o_p1); P.p2 = U(o_p2); P.ha = U(o_ha); P.pq = U(o_pq); P.s_g = U(o_sg);
    P.x2 = U(o_x2); P.h1 = U(o_h1); P.h2 = U(o_h2); P.x3 = U(o_x3); P.melf = U(o_melf); P.att_g = U(o_att); P.cum_g = U(o_cum);
    P.mel_out = F(o_mel); P.attn_out = F(o_attn); P.n_iter = (int *)(ws + o_n);
    P.gs.counter = (unsigned long long *)ws;
    P.gs.abort_flag = (int *)(ws + 64);
    // measured on B200 (scripts/taco_delay_sweep.py): a first look 300 .. 600 clocks after the own publish is 12 % faster
    // than an immediate one (a look that fails costs a second L2 round trip and queues ahead of the late stores)
    P.d_first = getenv("WRNN_TACO_D1") ? atoi(getenv("WRNN_TACO_D1")) : 450;
    P.d_poll = getenv("WRNN_TACO_D2") ? atoi(getenv("WRNN_TACO_D2")) : 0;
    P.prof = nullptr;
    long long *d_prof = nullptr;
    if (getenv("WRNN_TACO_PROF")) {
        TC_CUDA(h, cudaMalloc((void **)&d_prof, 64 * sizeof(long long)));
        TC_CUDA(h, cudaMemsetAsync(d_prof, 0, 64 * sizeof(long long), st));
        P.prof = d_prof;
    }
    TC_CUDA(h, cudaFuncSetAttribute(taco_loop_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
    TC_CUDA(h, cudaEventRecord(h->ev[0], st));
    taco_encproj_kernel<<<L, 256, D * sizeof(float), st>>>(P.enc, h->projT, F(o_encp), D);
    void *args[] = {(void *)&P};
    TC_CUDA(h, cudaLaunchCooperativeKernel((const void *)taco_loop_kernel, dim3(G), dim3(TC_THREADS), args, smem_bytes, st));
    TC_CUDA(h, cudaEventRecord(h->ev[1], st));
    h->launches = 2;
    int head[2] = {0, 0};           // abort flag, n_iter
    TC_CUDA(h, cudaMemcpyAsync(&head[0], ws + 64, sizeof(int), cudaMemcpyDeviceToHost, st));
    TC_CUDA(h, cudaMemcpyAsync(&head[1], ws + o_n, sizeof(int), cudaMemcpyDeviceToHost, st));
    TC_CUDA(h, cudaStreamSynchronize(st));
    TC_CUDA(h, cudaGetLastError());
    if (head[0]) return taco_fail(h, WRNN_ERR_CUDA, "Tacotron decoder loop kernel aborted: grid barrier watchdog fired");
    if (d_prof) {
        long long pr[64];
        cudaMemcpy(pr, d_prof, sizeof pr, cudaMemcpyDeviceToHost);
        cudaFree(d_prof);
        static const char *site[20] = {"A wait x3", "A dots", "B wait p1", "B dots", "C wait p2", "C dots+gates", "D wait h_attn", "D dots", "E / ahead",
                                       "barrier", "F stop+ctx", "F dots+gates", "G wait x2", "G dots+gates", "", "", "", "", "", ""};
        const int nn = std::max(1, head[1]);
        fprintf(stderr, "taco prof (clk per decoder step; CTA 0 | first position owner)\n");
        for (int i = 0; i < 14; i++) fprintf(stderr, "  %-14s %7.0f | %7.0f\n", site[i], (double)pr[i] / nn, (double)pr[32 + i] / nn);
    }
    const int n = head[1];
    if (n < 1 || n > P.max_iter) return taco_fail(h, WRNN_ERR_CUDA, "Tacotron decoder loop kernel returned an impossible step count");
    // mel: (n_mels, n * r) row-major as `torch.cat(mel_outputs, dim=2)[0]`; attn: (n, L)
    TC_CUDA(h, cudaMemcpy2DAsync(mel, (size_t)n * r * 4, ws + o_mel, (size_t)P.cap * 4, (size_t)n * r * 4, M, cudaMemcpyDeviceToHost, st));
    TC_CUDA(h, cudaMemcpyAsync(attn, ws + o_attn, (size_t)n * L * 4, cudaMemcpyDeviceToHost, st));
    TC_CUDA(h, cudaStreamSynchronize(st));
    *n_iter = n;
    return 0;
}

int wrnn_taco_last_timing(wrnn_taco_handle *h, float *ms_loop, int64_t *kernel_launches) {
    if (!h) return WRNN_ERR_ARG;
    float ms = 0;
    TC_CUDA(h, cudaEventSynchronize(h->ev[1]));
    TC_CUDA(h, cudaEventElapsedTime(&ms, h->ev[0], h->ev[1]));
    if (ms_loop) *ms_loop = ms;
    if (kernel_launches) *kernel_launches = h->launches;
    return 0;
}

}  // extern "C"

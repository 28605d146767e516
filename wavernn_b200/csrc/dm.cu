// deepmind coarse/fine dual-softmax WaveRNN generation (models/deepmind_version.py:75-165) as one
// persistent cooperative kernel: every SM keeps its rows of R, O1..O4 in shared memory for all steps;
// the layers exchange through small L2-resident buffers separated by grid barriers (6 per step).
//   A  R.hidden for the units this CTA owns (all three gates), coarse gates -> hidden_coarse       (:116-125)
//   B  O1 + ReLU      C  O2 -> coarse logits                                                        (:128-129)
//   D  softmax + Categorical sample of the coarse value (every CTA, redundantly), fine gates       (:130-145)
//   E  O3 + ReLU      F  O4 -> fine logits; the fine sample is drawn at the start of the next step (:148-152)
// The reference runs one sequence; `n_seq` independent sequences (<= 8 per launch) share the weights.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <map>
#include <string>
#include <vector>

#include "internal.h"
#include "loop_common.cuh"

namespace wrnn {

constexpr int DM_THREADS = 512, DM_WARPS = DM_THREADS / 32, DM_MAXN = 8;

struct DmParams {
    int H, S2, Q, N;             // hidden, hidden/2, quantisation, sequences in this launch
    int upc, r1, r2;             // hidden units / O1,O3 rows / O2,O4 rows per CTA
    int64_t T;                   // steps
    int img_floats;
    int off_R, off_O1, off_O2, off_O3, off_O4, off_I, off_b;
    const float *image;          // [G][img_floats]
    float *hidden;               // [2][N][H]
    float *o1, *o3;              // [N][S2]
    float *lc, *lf;              // [N][Q] coarse / fine logits
    const float *qc, *qf;        // injected Exp(1) noise [T][Ntot][Q] or null
    int64_t Ntot, n0;
    int32_t *coarse, *fine;      // [N][T]
    uint64_t seed;
    long long *prof;             // debug (WRNN_DM_PROF): clocks per site of CTA 0, summed over steps
    GridSync gs;
};

// dot products of `nrows` smem rows (stride ld) with N smem vectors (stride ldx), one warp per row, K % 4 == 0
template <int NMAX>
__device__ __forceinline__ void warp_rows(const float *W, int ld, int K, const float *X, int ldx, int N, int lane, float (&out)[NMAX]) {
    float acc[NMAX];
#pragma unroll
    for (int n = 0; n < NMAX; n++) acc[n] = 0.f;
    const float4 *w4 = reinterpret_cast<const float4 *>(W);
    const int k4n = K >> 2;
    if (NMAX <= 2) {
        // four chunks per pass, their loads issued before the first FMA (a lane still adds its chunks lane, lane + 32, ...
        // in ascending order: the bits of the simple loop below)
        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int k0 = lane; k0 < k4n; k0 += 128) {
            float4 w[4], x[NMAX][4];
#pragma unroll
            for (int c = 0; c < 4; c++) {
                const int k4 = k0 + 32 * c;
                w[c] = k4 < k4n ? w4[k4] : z;
#pragma unroll
                for (int n = 0; n < NMAX; n++) x[n][c] = k4 < k4n ? reinterpret_cast<const float4 *>(X + (size_t)n * ldx)[k4] : z;
            }
#pragma unroll
            for (int c = 0; c < 4; c++)
                if (k0 + 32 * c < k4n) {
#pragma unroll
                    for (int n = 0; n < NMAX; n++) {
                        acc[n] = fmaf(w[c].x, x[n][c].x, acc[n]);
                        acc[n] = fmaf(w[c].y, x[n][c].y, acc[n]);
                        acc[n] = fmaf(w[c].z, x[n][c].z, acc[n]);
                        acc[n] = fmaf(w[c].w, x[n][c].w, acc[n]);
                    }
                }
        }
    } else {
        for (int k4 = lane; k4 < k4n; k4 += 32) {
            const float4 w = w4[k4];
#pragma unroll
            for (int n = 0; n < NMAX; n++) {        // all NMAX sets, absent sequences included (their rows of X are shared memory
                const float4 x = reinterpret_cast<const float4 *>(X + (size_t)n * ldx)[k4];      // nobody reads the result of)
                acc[n] = fmaf(w.x, x.x, acc[n]);
                acc[n] = fmaf(w.y, x.y, acc[n]);
                acc[n] = fmaf(w.z, x.z, acc[n]);
                acc[n] = fmaf(w.w, x.w, acc[n]);
            }
        }
    }
    // the NMAX xor trees level by level, so that their shuffles overlap (each tree's own order is unchanged)
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) {
#pragma unroll
        for (int n = 0; n < NMAX; n++) acc[n] += __shfl_xor_sync(0xffffffffu, acc[n], o);
    }
#pragma unroll
    for (int n = 0; n < NMAX; n++) out[n] = acc[n];
    (void)ld;
    (void)N;
}

// `nrows` rows x NS sequences as jobs of (one row, GS consecutive sequences) spread over all warps: with 4 rows and 8
// sequences one warp per row leaves 12 of 16 warps idle; a job keeps `warp_rows`' per-lane order for every (row, sequence),
// so the sums are the same bits.  emit(row, sequence, value) is called by lane 0.
template <int NS, int GS, typename F>
__device__ __forceinline__ void rows_split(const float *W, int ldw, int K, const float *X, int ldx, int nrows, int N, int warp, int lane, F emit,
                                           int part = 0, int parts = 1) {
    constexpr int NG = NS / GS;
    float out[GS];
    const int jobs = nrows * NG, jb = (int)((long long)jobs * part / parts), je = (int)((long long)jobs * (part + 1) / parts);
    for (int job = jb + warp; job < je; job += DM_WARPS) {
        const int j = job / NG, n0 = (job - j * NG) * GS;
        if (n0 >= N) continue;
        warp_rows<GS>(W + (size_t)j * ldw, ldw, K, X + (size_t)n0 * ldx, ldx, N - n0, lane, out);
        if (lane == 0) {
#pragma unroll
            for (int g = 0; g < GS; g++)
                if (n0 + g < N) emit(j, n0 + g, out[g]);
        }
    }
}

// R . hidden: a job = the THREE gate rows of one unit x GS sequences, so a staged activation chunk is read once for three
// rows (the product is bound by shared-memory reads of the activations: one row per job re-reads them 21 times per CTA).
// Per (row, sequence) the chunks are still taken lane, lane + 32, ... in ascending order and reduced by the same xor tree.
template <int NS, int GS, typename F>
__device__ __forceinline__ void unit_rows_split(const float *W, int K, const float *X, int ldx, int nunits, int N, int warp, int lane, F emit,
                                                int part, int parts) {
    constexpr int NG = NS / GS;
    const int jobs = nunits * NG, jb = (int)((long long)jobs * part / parts), je = (int)((long long)jobs * (part + 1) / parts);
    for (int job = jb + warp; job < je; job += DM_WARPS) {
        const int u = job / NG, n0 = (job - u * NG) * GS;
        if (n0 >= N) continue;
        float acc[3][GS];
#pragma unroll
        for (int g = 0; g < 3; g++)
#pragma unroll
            for (int n = 0; n < GS; n++) acc[g][n] = 0.f;
        const float4 *w4 = reinterpret_cast<const float4 *>(W + (size_t)u * 3 * K);
        const int k4n = K >> 2;
        for (int k4 = lane; k4 < k4n; k4 += 32) {
            float4 w[3];
#pragma unroll
            for (int g = 0; g < 3; g++) w[g] = w4[g * k4n + k4];
#pragma unroll
            for (int n = 0; n < GS; n++) {
                const float4 x = reinterpret_cast<const float4 *>(X + (size_t)(n0 + n) * ldx)[k4];
#pragma unroll
                for (int g = 0; g < 3; g++) {
                    acc[g][n] = fmaf(w[g].x, x.x, acc[g][n]);
                    acc[g][n] = fmaf(w[g].y, x.y, acc[g][n]);
                    acc[g][n] = fmaf(w[g].z, x.z, acc[g][n]);
                    acc[g][n] = fmaf(w[g].w, x.w, acc[g][n]);
                }
            }
        }
#pragma unroll
        for (int g = 0; g < 3; g++)
#pragma unroll
            for (int n = 0; n < GS; n++) {
#pragma unroll
                for (int o = 16; o >= 1; o >>= 1) acc[g][n] += __shfl_xor_sync(0xffffffffu, acc[g][n], o);
                if (lane == 0 && n0 + n < N) emit(u * 3 + g, n0 + n, acc[g][n]);
            }
    }
}

// One (unit, GS sequences) job of R . hidden over the float4 chunks [c0, c1) of the rows, CONTINUING the per-lane
// accumulators: lane l takes its chunks l, l + 32, ... in ascending order, so ranges visited in ascending order over
// several calls give the bits of one pass (`unit_rows_split`).
template <int GS>
__device__ __forceinline__ void unit_rows_range(const float *W, int K, const float *X, int ldx, int u, int n0, int c0, int c1, int lane,
                                                float (&acc)[3][GS]) {
    const float4 *w4 = reinterpret_cast<const float4 *>(W + (size_t)u * 3 * K);
    const int k4n = K >> 2;
    for (int k4 = c0 + ((lane - c0) & 31); k4 < c1; k4 += 32) {
        float4 w[3];
#pragma unroll
        for (int g = 0; g < 3; g++) w[g] = w4[g * k4n + k4];
#pragma unroll
        for (int n = 0; n < GS; n++) {
            const float4 x = reinterpret_cast<const float4 *>(X + (size_t)(n0 + n) * ldx)[k4];
#pragma unroll
            for (int g = 0; g < 3; g++) {
                acc[g][n] = fmaf(w[g].x, x.x, acc[g][n]);
                acc[g][n] = fmaf(w[g].y, x.y, acc[g][n]);
                acc[g][n] = fmaf(w[g].z, x.z, acc[g][n]);
                acc[g][n] = fmaf(w[g].w, x.w, acc[g][n]);
            }
        }
    }
}

// rows x cols floats, row stride ld (global) / lds (shared), cols % 4 == 0
__device__ __forceinline__ void stage_cg_2d(float *dst, int lds, const float *src, int ld, int rows, int cols, int tid) {
    const int c4 = cols >> 2, n4 = rows * c4;
    for (int base = 0; base < n4; base += 2 * DM_THREADS) {          // two loads in flight per thread: one L2 round trip for 8 x 448
        const int i0 = base + tid, i1 = i0 + DM_THREADS;
        const int r0 = i0 / c4, q0 = i0 - r0 * c4, r1 = i1 / c4, q1 = i1 - r1 * c4;
        float4 v0, v1;
        if (i0 < n4) v0 = __ldcg(reinterpret_cast<const float4 *>(src + (size_t)r0 * ld) + q0);
        if (i1 < n4) v1 = __ldcg(reinterpret_cast<const float4 *>(src + (size_t)r1 * ld) + q1);
        if (i0 < n4) reinterpret_cast<float4 *>(dst + (size_t)r0 * lds)[q0] = v0;
        if (i1 < n4) reinterpret_cast<float4 *>(dst + (size_t)r1 * lds)[q1] = v1;
    }
}
// the grid barrier in two halves: work that needs nothing from the other CTAs goes between arrive and wait
__device__ __forceinline__ void dm_arrive(const GridSync &gs, unsigned long long &gen) {
    __syncthreads();
    gen++;
    if (threadIdx.x == 0) red_release_add_u64(gs.counter, 1ull);
}
__device__ __forceinline__ void dm_wait(const GridSync &gs, unsigned long long gen) {
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

__device__ __forceinline__ void stage_cg(float *dst, const float *src, int n, int tid) {     // written by other CTAs
    for (int i = tid; i < (n >> 2); i += DM_THREADS) reinterpret_cast<float4 *>(dst)[i] = __ldcg(reinterpret_cast<const float4 *>(src) + i);
}


// The Exp(1) draws a thread will need in `dm_sample_block` (same thread -> (sequence, class) mapping, same Philox call and
// component, or the injected value): they depend on the step only, so they are formed in a barrier's slack, off the chain.
__device__ __forceinline__ void dm_draw_noise(float (&q)[4], int C, int N, const float *qsrc, size_t qstride, uint2 key, uint32_t c0,
                                              uint32_t c1base, uint32_t stream, int tid) {
    const int warp = tid >> 5, lane = tid & 31;
    const int WPS = N <= 2 ? 8 : N <= 4 ? 4 : 2, JW = 8 / WPS;
    const int n = warp / WPS, g = warp - n * WPS;
#pragma unroll
    for (int jj = 0; jj < 4; jj++) q[jj] = 1.f;
    if (n >= N) return;
    const int call = (g * JW) >> 2;          // the Philox call of categorical_warp that covers this warp's j range
    uint4 r = make_uint4(0, 0, 0, 0);
    if (!qsrc) r = philox4x32(make_uint4(c0, c1base + (uint32_t)n, (uint32_t)(call * 32 + lane), stream), key);
    const uint32_t rr[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int jj = 0; jj < 4; jj++) {
        const int j = g * JW + jj, i = lane + 32 * j;
        if (jj < JW && i < C) q[jj] = qsrc ? qsrc[(size_t)n * qstride + i] : exp1(rr[j & 3]);
    }
}

// Categorical draw of N <= 8 sequences by the whole block.  `categorical_warp` (loop_common.cuh) gives a lane the classes
// lane + 32 j and runs three dependent passes of expf over them in ONE warp: 9.6 k clocks per draw, two draws per step = 45 %
// of the step (measured, WRNN_DM_PROF).  Here a sequence gets 8 / 4 / 2 warps (1-2 / 3-4 / 5-8 sequences), expf is evaluated
// once, and the softmax is never normalised: the pick is argmax_i (p_i / sum p) / q_i with p_i = e_i / s, where s and
// sum p are common positive factors of a sequence, i.e. argmax_i e_i / q_i (ties to the lower index, as before).  The
// reference fixtures and 2 x 44 100 + 2 x 64 000 draws against the oracle are reproduced exactly
// (tests/manual/dm_long_check.py).  C <= 256.
__device__ __forceinline__ void dm_sample_block(const float *lrow, int C, int N, const float (&qpre)[4], float *scr, int *res, int tid) {
    const int warp = tid >> 5, lane = tid & 31;
    const int WPS = N <= 2 ? 8 : N <= 4 ? 4 : 2, JW = 8 / WPS;
    const int n = warp / WPS, g = warp - n * WPS;
    const bool act = n < N;
    float *vbuf = scr + DM_MAXN * 256;       // [N][8][32] best value per (sequence, warp of the sequence, lane)
    int *ibuf = reinterpret_cast<int *>(vbuf + DM_MAXN * 256);
    float *wmax = reinterpret_cast<float *>(ibuf + DM_MAXN * 256);       // [16]
    const float *lg = lrow + (size_t)(act ? n : 0) * C;
    float l[4], e[4];
    float m = -INFINITY;
#pragma unroll
    for (int jj = 0; jj < 4; jj++) {
        const int i = lane + 32 * (g * JW + jj);
        l[jj] = (act && jj < JW && i < C) ? lg[i] : -INFINITY;
        m = fmaxf(m, l[jj]);
    }
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (lane == 0) wmax[warp] = m;
    __syncthreads();
    if (act) {
        m = wmax[n * WPS];
        for (int k = 1; k < WPS; k++) m = fmaxf(m, wmax[n * WPS + k]);
    }
    // argmax_i (p_i / sum p) / q_i with p_i = e_i / s: s and sum p are common positive factors of a sequence, so the pick is
    // argmax_i e_i / q_i and neither sum is formed (two block barriers and two sequential 8-term sums per draw less)
    float best = -INFINITY;
    int bi = 0x7fffffff;
    if (act) {
#pragma unroll
        for (int jj = 0; jj < 4; jj++) {
            const int j = g * JW + jj, i = lane + 32 * j;
            if (jj < JW && i < C) {
                e[jj] = expf(l[jj] - m);
                const float v = e[jj] / qpre[jj];
                if (v > best) { best = v; bi = i; }    // ascending i per lane: first max kept
            }
        }
        vbuf[(n * 8 + g) * 32 + lane] = best;
        ibuf[(n * 8 + g) * 32 + lane] = bi;
    }
    __syncthreads();
    if (act && g == 0) {
        best = -INFINITY;
        bi = 0x7fffffff;
        for (int k = 0; k < WPS; k++) {
            const float v = vbuf[(n * 8 + k) * 32 + lane];
            if (v > best) { best = v; bi = ibuf[(n * 8 + k) * 32 + lane]; }
        }
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
        }
        if (lane == 0) res[n] = bi;
    }
    __syncthreads();
}

// NS = accumulator sets compiled in (1, 2, 4 or 8 >= N): the products carry no predicated sets for absent sequences
template <int NS>
__global__ void __launch_bounds__(DM_THREADS, 1) dm_loop_kernel(const DmParams P) {
    extern __shared__ __align__(16) float smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, cta = blockIdx.x;
    const int H = P.H, S2 = P.S2, Q = P.Q, N = P.N;
    float *wimg = smem;
    float *xbuf = wimg + P.img_floats;               // [N][H] staged inputs
    float *xo = xbuf + (size_t)DM_MAXN * H;          // [N][S2] o1 / o3 as staged
    float *rsum = xo + (size_t)DM_MAXN * S2;         // [upc][3][N] R.hidden of the owned units, formed one step ahead
    float *lrow = rsum + (size_t)P.upc * 3 * DM_MAXN;   // [N][Q] staged logits
    float *sscr = lrow + (size_t)DM_MAXN * Q;           // sampler scratch (dm_sample_block)
    __shared__ int picks[DM_MAXN];
    const bool fast_draw = Q <= 256;
    __shared__ int prev_c[DM_MAXN], prev_f[DM_MAXN], cur_c[DM_MAXN];
    for (int i = tid; i < P.img_floats / 4; i += DM_THREADS)
        reinterpret_cast<float4 *>(wimg)[i] = reinterpret_cast<const float4 *>(P.image + (size_t)cta * P.img_floats)[i];
    if (tid < DM_MAXN) prev_c[tid] = prev_f[tid] = cur_c[tid] = 0;
    for (int i = tid; i < DM_MAXN * H; i += DM_THREADS) xbuf[i] = 0.f;                 // hidden = 0 (:167-169)
    for (int i = tid; i < P.upc * 3 * DM_MAXN; i += DM_THREADS) rsum[i] = 0.f;         // R . 0
    __syncthreads();
    const float *R = wimg + P.off_R, *O1 = wimg + P.off_O1, *O2 = wimg + P.off_O2, *O3 = wimg + P.off_O3, *O4 = wimg + P.off_O4;
    const float *Iw = wimg + P.off_I;      // per unit: 3 gates x 3 input weights (coarse units use 2)
    const float *bb = wimg + P.off_b;      // per unit: b_u, b_r, b_e | then O1 bias[r1] O2[r2] O3[r1] O4[r2]
    const float *bO1 = bb + 3 * P.upc, *bO2 = bO1 + P.r1, *bO3 = bO2 + P.r2, *bO4 = bO3 + P.r1;
    const int u0 = cta * P.upc, nu = max(0, min(P.upc, H - u0));
    const int a0 = cta * P.r1, na = max(0, min(P.r1, S2 - a0));
    const int c0 = cta * P.r2, nc = max(0, min(P.r2, Q - c0));
    const uint2 key = make_uint2((uint32_t)P.seed, (uint32_t)(P.seed >> 32));
    unsigned long long gen = 0;
    // R . hidden of the next step, spread over the slack of four barriers when every (unit, sequence group) job has a warp
    // of its own: the accumulators then live in registers from C to F (quarter of the columns per barrier)
    constexpr int GSR = NS >= 4 ? 4 : NS, NGR = NS / GSR;
    const bool spread = nu * NGR <= DM_WARPS;
    const int ru = warp / NGR, rn0 = (warp - ru * NGR) * GSR;
    const bool rjob = spread && warp < nu * NGR && rn0 < N;
    const int rq = (H >> 2) / 4;        // float4 chunks per quarter of a row
    float racc[3][GSR];
    float q_coarse[4] = {1.f, 1.f, 1.f, 1.f}, q_fine[4] = {1.f, 1.f, 1.f, 1.f};      // Exp(1) noise of the two draws, formed ahead
    __shared__ long long prof_acc[16];
    long long prof_last = 0;
    const bool prof_on = P.prof != nullptr && tid == 0 && cta == 0;
    if (tid < 16) prof_acc[tid] = 0;
#define DMARK(i)                                      \
    if (prof_on) {                                    \
        const long long c__ = clock64();              \
        prof_acc[i] += c__ - prof_last;               \
        prof_last = c__;                              \
    }

    for (int64_t t = 0; t <= P.T; t++) {
        float *hnxt = P.hidden + (size_t)((t + 1) & 1) * N * H;
        // ---- fine sample of step t-1 (every CTA draws the same value; CTA 0 records it)
        if (prof_on) prof_last = clock64();
        if (t > 0) {
            stage_cg(lrow, P.lf, N * Q, tid);
            __syncthreads();
            DMARK(0);
            if (fast_draw) {
                dm_sample_block(lrow, Q, N, q_fine, sscr, picks, tid);
                if (tid < N) {
                    prev_f[tid] = picks[tid];
                    prev_c[tid] = cur_c[tid];
                    if (cta == 0) P.fine[(size_t)tid * P.T + t - 1] = picks[tid];
                }
            } else {
                for (int n = warp; n < N; n += DM_WARPS) {
                    const float *qs = P.qf ? P.qf + ((size_t)(t - 1) * P.Ntot + P.n0 + n) * Q : nullptr;
                    const int k = categorical_warp(lrow + (size_t)n * Q, Q, qs, key, (uint32_t)(t - 1), (uint32_t)(P.n0 + n), 5u, lane);
                    if (lane == 0) {
                        prev_f[n] = k;
                        prev_c[n] = cur_c[n];
                        if (cta == 0) P.fine[(size_t)n * P.T + t - 1] = k;
                    }
                }
            }
            __syncthreads();
        }
        DMARK(1);
        if (t == P.T) break;
        // ---- A: coarse units update (R . hidden was formed in the barrier waits of E and F of the step before; xbuf still
        //         holds that hidden state)
        for (int i = tid; i < nu * N; i += DM_THREADS) {
            const int u = i / N, n = i - u * N, unit = u0 + u;
            if (unit < S2) {
                const float pc = (float)prev_c[n] / 127.5f - 1.f, pf = (float)prev_f[n] / 127.5f - 1.f;      // :106-108
                const float *iw = Iw + u * 9;
                const float Iu = iw[0] * pc + iw[1] * pf, Ir = iw[3] * pc + iw[4] * pf, Ie = iw[6] * pc + iw[7] * pf;
                const float uu = sigmoidf_acc(rsum[(u * 3 + 0) * DM_MAXN + n] + Iu + bb[u * 3 + 0]);
                const float rr = sigmoidf_acc(rsum[(u * 3 + 1) * DM_MAXN + n] + Ir + bb[u * 3 + 1]);
                const float ee = tanhf(rr * rsum[(u * 3 + 2) * DM_MAXN + n] + Ie + bb[u * 3 + 2]);
                hnxt[(size_t)n * H + unit] = uu * xbuf[n * H + unit] + (1.f - uu) * ee;                    // :125
            }
        }
        DMARK(2);
        dm_arrive(P.gs, gen);                   // in this barrier's slack: the Exp(1) noise of the coarse draw of this step (D)
        if (fast_draw) dm_draw_noise(q_coarse, Q, N, P.qc ? P.qc + ((size_t)t * P.Ntot + P.n0) * Q : nullptr, (size_t)Q, key, (uint32_t)t,
                                     (uint32_t)P.n0, 4u, tid);
        dm_wait(P.gs, gen);
        DMARK(3);
        // ---- B: O1 + ReLU on hidden_coarse
        stage_cg_2d(xbuf, H, hnxt, H, N, S2, tid);        // the coarse half; the fine half stays the previous step's for D
        __syncthreads();
        DMARK(14);
        rows_split<NS, (NS >= 4 ? 2 : 1)>(O1, S2, S2, xbuf, H, na, N, warp, lane,
                       [&](int j, int n, float v) { P.o1[(size_t)n * S2 + a0 + j] = fmaxf(v + bO1[j], 0.f); });
        DMARK(4);
        dm_arrive(P.gs, gen);                   // and here the noise of the fine draw, made at the top of the next step
        if (fast_draw) dm_draw_noise(q_fine, Q, N, P.qf ? P.qf + ((size_t)t * P.Ntot + P.n0) * Q : nullptr, (size_t)Q, key, (uint32_t)t,
                                     (uint32_t)P.n0, 5u, tid);
        dm_wait(P.gs, gen);
        DMARK(5);
        // ---- C: O2 -> coarse logits
        stage_cg_2d(xo, S2, P.o1, S2, N, S2, tid);
        __syncthreads();
        rows_split<NS, (NS >= 4 ? 2 : 1)>(O2, S2, S2, xo, S2, nc, N, warp, lane, [&](int j, int n, float v) { P.lc[(size_t)n * Q + c0 + j] = v + bO2[j]; });
        DMARK(6);
        dm_arrive(P.gs, gen);
        if (rjob) {
#pragma unroll
            for (int g = 0; g < 3; g++)
#pragma unroll
                for (int n = 0; n < GSR; n++) racc[g][n] = 0.f;
            unit_rows_range<GSR>(R, H, xbuf, H, ru, rn0, 0, rq, lane, racc);                 // coarse columns, first half
        }
        dm_wait(P.gs, gen);
        DMARK(7);
        // ---- D: coarse sample (redundant), fine gates
        stage_cg(lrow, P.lc, N * Q, tid);
        __syncthreads();
        if (fast_draw) {
            dm_sample_block(lrow, Q, N, q_coarse, sscr, picks, tid);
            if (tid < N) {
                cur_c[tid] = picks[tid];
                if (cta == 0) P.coarse[(size_t)tid * P.T + t] = picks[tid];
            }
        } else {
            for (int n = warp; n < N; n += DM_WARPS) {
                const float *qs = P.qc ? P.qc + ((size_t)t * P.Ntot + P.n0 + n) * Q : nullptr;
                const int k = categorical_warp(lrow + (size_t)n * Q, Q, qs, key, (uint32_t)t, (uint32_t)(P.n0 + n), 4u, lane);
                if (lane == 0) {
                    cur_c[n] = k;
                    if (cta == 0) P.coarse[(size_t)n * P.T + t] = k;
                }
            }
        }
        __syncthreads();
        for (int i = tid; i < nu * N; i += DM_THREADS) {
            const int u = i / N, n = i - u * N, unit = u0 + u;
            if (unit >= S2) {
                const float pc = (float)prev_c[n] / 127.5f - 1.f, pf = (float)prev_f[n] / 127.5f - 1.f;
                const float cc = (float)cur_c[n] / 127.5f - 1.f;                                          // :135
                const float *iw = Iw + u * 9;
                const float Iu = iw[0] * pc + iw[1] * pf + iw[2] * cc, Ir = iw[3] * pc + iw[4] * pf + iw[5] * cc,
                            Ie = iw[6] * pc + iw[7] * pf + iw[8] * cc;
                const float uu = sigmoidf_acc(rsum[(u * 3 + 0) * DM_MAXN + n] + Iu + bb[u * 3 + 0]);
                const float rr = sigmoidf_acc(rsum[(u * 3 + 1) * DM_MAXN + n] + Ir + bb[u * 3 + 1]);
                const float ee = tanhf(rr * rsum[(u * 3 + 2) * DM_MAXN + n] + Ie + bb[u * 3 + 2]);
                hnxt[(size_t)n * H + unit] = uu * xbuf[n * H + unit] + (1.f - uu) * ee;                    // :145
            }
        }
        DMARK(8);
        dm_arrive(P.gs, gen);
        if (rjob) unit_rows_range<GSR>(R, H, xbuf, H, ru, rn0, rq, 2 * rq, lane, racc);             // coarse columns, second half
        dm_wait(P.gs, gen);
        DMARK(9);
        // ---- E: O3 + ReLU on hidden_fine
        stage_cg_2d(xbuf + S2, H, hnxt + S2, H, N, S2, tid);     // xbuf = the whole new hidden state from here on
        __syncthreads();
        rows_split<NS, (NS >= 4 ? 2 : 1)>(O3, S2, S2, xbuf + S2, H, na, N, warp, lane,
                       [&](int j, int n, float v) { P.o3[(size_t)n * S2 + a0 + j] = fmaxf(v + bO3[j], 0.f); });
        DMARK(10);
        // R . hidden for the NEXT step's gates, first half of the jobs, between this barrier's arrive and wait
        dm_arrive(P.gs, gen);
        if (spread) {
            if (rjob) unit_rows_range<GSR>(R, H, xbuf, H, ru, rn0, 2 * rq, 3 * rq, lane, racc);         // fine columns, first half
        } else {
            unit_rows_split<NS, GSR>(R, H, xbuf, H, nu, N, warp, lane, [&](int j, int n, float v) { rsum[j * DM_MAXN + n] = v; }, 0, 2);
        }
        dm_wait(P.gs, gen);
        DMARK(11);
        // ---- F: O4 -> fine logits
        stage_cg_2d(xo, S2, P.o3, S2, N, S2, tid);
        __syncthreads();
        rows_split<NS, (NS >= 4 ? 2 : 1)>(O4, S2, S2, xo, S2, nc, N, warp, lane, [&](int j, int n, float v) { P.lf[(size_t)n * Q + c0 + j] = v + bO4[j]; });
        DMARK(12);
        dm_arrive(P.gs, gen);                   // second half (j = u*3 + gate)
        if (spread) {
            if (rjob) {
                unit_rows_range<GSR>(R, H, xbuf, H, ru, rn0, 3 * rq, H >> 2, lane, racc);               // fine columns, second half
#pragma unroll
                for (int g = 0; g < 3; g++)
#pragma unroll
                    for (int n = 0; n < GSR; n++) {
#pragma unroll
                        for (int o = 16; o >= 1; o >>= 1) racc[g][n] += __shfl_xor_sync(0xffffffffu, racc[g][n], o);
                        if (lane == 0 && rn0 + n < N) rsum[(ru * 3 + g) * DM_MAXN + rn0 + n] = racc[g][n];
                    }
            }
        } else {
            unit_rows_split<NS, GSR>(R, H, xbuf, H, nu, N, warp, lane, [&](int j, int n, float v) { rsum[j * DM_MAXN + n] = v; }, 1, 2);
        }
        dm_wait(P.gs, gen);
        DMARK(13);
    }
    if (prof_on)
        for (int i = 0; i < 16; i++) P.prof[i] = prof_acc[i];
}

}  // namespace wrnn

// ============================================================================ host
struct wrnn_dm_handle {
    int hidden = 0, quant = 0, device = 0, num_sms = 0;
    size_t smem_optin = 0;
    std::string err;
    std::map<std::string, wrnn::HostTensor> w;
    bool packed = false;
    wrnn::DmParams plan;         // layout fields only
    float *image = nullptr;
    size_t smem_bytes = 0;
    void *ws = nullptr;
    size_t ws_bytes = 0;
    cudaEvent_t ev[2] = {nullptr, nullptr};
    int64_t launches = 0;
};

static thread_local std::string g_dm_err;

namespace wrnn {

static int dm_fail(wrnn_dm_handle *h, int code, const char *msg) {
    if (h) h->err = msg;
    else g_dm_err = msg;
    if (code == WRNN_ERR_CUDA) cudaGetLastError();
    return code;
}
#define DM_CUDA(h, expr)                                                                          \
    do {                                                                                          \
        cudaError_t e__ = (expr);                                                                 \
        if (e__ != cudaSuccess) {                                                                 \
            char b__[512];                                                                        \
            snprintf(b__, sizeof b__, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__); \
            return wrnn::dm_fail((h), WRNN_ERR_CUDA, b__);                                        \
        }                                                                                         \
    } while (0)

static int dm_pack(wrnn_dm_handle *h) {
    const int H = h->hidden, S2 = H / 2, Q = h->quant, G = h->num_sms;
    auto get = [&](const char *k, std::vector<int64_t> shape) -> const HostTensor * {
        auto it = h->w.find(k);
        if (it == h->w.end() || it->second.shape != shape) {
            h->err = std::string("checkpoint key missing or misshapen: ") + k;
            return nullptr;
        }
        return &it->second;
    };
    const HostTensor *R = get("R.weight", {3 * H, H}), *O1 = get("O1.weight", {S2, S2}), *O1b = get("O1.bias", {S2}),
                     *O2 = get("O2.weight", {Q, S2}), *O2b = get("O2.bias", {Q}), *O3 = get("O3.weight", {S2, S2}),
                     *O3b = get("O3.bias", {S2}), *O4 = get("O4.weight", {Q, S2}), *O4b = get("O4.bias", {Q}),
                     *Ic = get("I_coarse.weight", {3 * S2, 2}), *If = get("I_fine.weight", {3 * S2, 3}),
                     *bu = get("bias_u", {H}), *br = get("bias_r", {H}), *be = get("bias_e", {H});
    if (!R || !O1 || !O1b || !O2 || !O2b || !O3 || !O3b || !O4 || !O4b || !Ic || !If || !bu || !br || !be) return WRNN_ERR_STATE;
    DmParams &p = h->plan;
    p.H = H; p.S2 = S2; p.Q = Q;
    p.upc = (H + G - 1) / G;
    p.r1 = (S2 + G - 1) / G;
    p.r2 = (Q + G - 1) / G;
    int o = 0;
    p.off_R = o;  o += p.upc * 3 * H;
    p.off_O1 = o; o += p.r1 * S2;
    p.off_O2 = o; o += p.r2 * S2;
    p.off_O3 = o; o += p.r1 * S2;
    p.off_O4 = o; o += p.r2 * S2;
    p.off_I = o;  o += p.upc * 9;
    p.off_b = o;  o += p.upc * 3 + 2 * p.r1 + 2 * p.r2;
    p.img_floats = (o + 3) & ~3;
    h->smem_bytes = ((size_t)p.img_floats + (size_t)DM_MAXN * H + (size_t)DM_MAXN * S2 + (size_t)p.upc * 3 * DM_MAXN + (size_t)DM_MAXN * Q + 3 * (size_t)DM_MAXN * 256 + 64) * sizeof(float);
    if (h->smem_bytes > h->smem_optin) return dm_fail(h, WRNN_ERR_ARG, "deepmind model too large for the shared-memory resident kernel");
    std::vector<float> img((size_t)G * p.img_floats, 0.f);
    for (int c = 0; c < G; c++) {
        float *im = img.data() + (size_t)c * p.img_floats;
        float *bb = im + p.off_b, *bO1 = bb + 3 * p.upc, *bO2 = bO1 + p.r1, *bO3 = bO2 + p.r2, *bO4 = bO3 + p.r1;
        for (int u = 0; u < p.upc; u++) {
            const int unit = c * p.upc + u;
            if (unit >= H) break;
            for (int g = 0; g < 3; g++) {
                // R.hidden splits into [u_c, u_f, r_c, r_f, e_c, e_f] (:117-119): gate g of unit j is row g*H + j
                memcpy(im + p.off_R + (size_t)(u * 3 + g) * H, R->v.data() + (size_t)(g * H + unit) * H, (size_t)H * sizeof(float));
                float *iw = im + p.off_I + u * 9 + g * 3;
                if (unit < S2) { iw[0] = Ic->v[(size_t)(g * S2 + unit) * 2]; iw[1] = Ic->v[(size_t)(g * S2 + unit) * 2 + 1]; }
                else for (int k = 0; k < 3; k++) iw[k] = If->v[(size_t)(g * S2 + unit - S2) * 3 + k];
            }
            bb[u * 3 + 0] = bu->v[unit]; bb[u * 3 + 1] = br->v[unit]; bb[u * 3 + 2] = be->v[unit];
        }
        for (int r = 0; r < p.r1; r++) {
            const int row = c * p.r1 + r;
            if (row >= S2) break;
            memcpy(im + p.off_O1 + (size_t)r * S2, O1->v.data() + (size_t)row * S2, (size_t)S2 * sizeof(float));
            memcpy(im + p.off_O3 + (size_t)r * S2, O3->v.data() + (size_t)row * S2, (size_t)S2 * sizeof(float));
            bO1[r] = O1b->v[row]; bO3[r] = O3b->v[row];
        }
        for (int r = 0; r < p.r2; r++) {
            const int row = c * p.r2 + r;
            if (row >= Q) break;
            memcpy(im + p.off_O2 + (size_t)r * S2, O2->v.data() + (size_t)row * S2, (size_t)S2 * sizeof(float));
            memcpy(im + p.off_O4 + (size_t)r * S2, O4->v.data() + (size_t)row * S2, (size_t)S2 * sizeof(float));
            bO2[r] = O2b->v[row]; bO4[r] = O4b->v[row];
        }
    }
    if (h->image) cudaFree(h->image);
    h->image = nullptr;
    DM_CUDA(h, cudaMalloc((void **)&h->image, img.size() * sizeof(float)));
    DM_CUDA(h, cudaMemcpy(h->image, img.data(), img.size() * sizeof(float), cudaMemcpyHostToDevice));
    h->packed = true;
    return 0;
}

}  // namespace wrnn

using namespace wrnn;

extern "C" {

int wrnn_dm_create(int32_t hidden_size, int32_t quantisation, int device, wrnn_dm_handle **out) {
    if (!out) return dm_fail(nullptr, WRNN_ERR_ARG, "null argument");
    *out = nullptr;
    if (hidden_size < 8 || hidden_size % 8 || quantisation < 4 || quantisation % 4)
        return dm_fail(nullptr, WRNN_ERR_ARG, "hidden_size must be a multiple of 8 and quantisation a multiple of 4");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) return dm_fail(nullptr, WRNN_ERR_CUDA, "no CUDA device: this library has no CPU fallback");
    if (device < 0 || device >= ndev) return dm_fail(nullptr, WRNN_ERR_ARG, "device out of range");
    cudaDeviceProp prop;
    wrnn::DeviceGuard guard(device);
    if (guard.err != cudaSuccess || cudaGetDeviceProperties(&prop, device) != cudaSuccess)
        return dm_fail(nullptr, WRNN_ERR_CUDA, "cannot open device");
    wrnn_dm_handle *h = new wrnn_dm_handle();
    h->hidden = hidden_size; h->quant = quantisation; h->device = device;
    h->num_sms = prop.multiProcessorCount;
    h->smem_optin = prop.sharedMemPerBlockOptin;
    cudaEventCreate(&h->ev[0]);
    cudaEventCreate(&h->ev[1]);
    *out = h;
    return 0;
}

void wrnn_dm_destroy(wrnn_dm_handle *h) {
    if (!h) return;
    wrnn::DeviceGuard guard(h->device);
    if (h->image) cudaFree(h->image);
    if (h->ws) cudaFree(h->ws);
    for (int i = 0; i < 2; i++)
        if (h->ev[i]) cudaEventDestroy(h->ev[i]);
    delete h;
}

const char *wrnn_dm_last_error(const wrnn_dm_handle *h) { return h ? h->err.c_str() : g_dm_err.c_str(); }

int wrnn_dm_load_weights(wrnn_dm_handle *h, const wrnn_tensor *tensors, int32_t n) {
    if (!h || (!tensors && n > 0)) return dm_fail(h, WRNN_ERR_ARG, "null argument");
    for (int i = 0; i < n; i++) {
        const wrnn_tensor &t = tensors[i];
        if (!t.name || !t.data || t.ndim < 0 || t.ndim > 4) return dm_fail(h, WRNN_ERR_ARG, "bad tensor");
        if (t.dtype != 0) continue;
        HostTensor ht;
        int64_t numel = 1;
        for (int k = 0; k < t.ndim; k++) { ht.shape.push_back(t.shape[k]); numel *= t.shape[k]; }
        ht.v.assign((const float *)t.data, (const float *)t.data + numel);
        h->w[t.name] = std::move(ht);
    }
    h->packed = false;
    return 0;
}

}  // extern "C"

extern "C" {

int wrnn_dm_generate(wrnn_dm_handle *h, int64_t seq_len, int32_t n_seq, const float *q_coarse, const float *q_fine,
                     uint64_t seed, int64_t *output, int64_t *coarse, int64_t *fine, void *cuda_stream) {
    if (!h || !output || !coarse || !fine || seq_len < 1 || n_seq < 1) return dm_fail(h, WRNN_ERR_ARG, "bad argument");
    if ((q_coarse == nullptr) != (q_fine == nullptr)) return dm_fail(h, WRNN_ERR_ARG, "give both noise tensors or none");
    cudaStream_t st = (cudaStream_t)cuda_stream;
    wrnn::DeviceGuard guard(h->device);
    DM_CUDA(h, guard.err);
    if (!h->packed) {
        if (h->w.empty()) return dm_fail(h, WRNN_ERR_STATE, "weights not loaded: call wrnn_dm_load_weights first");
        const int rc = dm_pack(h);
        if (rc) return rc;
    }
    const int H = h->hidden, S2 = H / 2, Q = h->quant;
    const size_t nq = q_coarse ? (size_t)seq_len * n_seq * Q : 0;
    // workspace: sync[128 B] | hidden[2][8][H] | o1, o3 [8][S2] | lc, lf [8][Q] | coarse, fine int32 [n_seq][T] | noise
    size_t off = 128;
    const size_t o_hidden = off; off += (size_t)2 * DM_MAXN * H * 4;
    const size_t o_o1 = off; off += (size_t)DM_MAXN * S2 * 4;
    const size_t o_o3 = off; off += (size_t)DM_MAXN * S2 * 4;
    const size_t o_lc = off; off += (size_t)DM_MAXN * Q * 4;
    const size_t o_lf = off; off += (size_t)DM_MAXN * Q * 4;
    const size_t act_end = off;
    const size_t o_c = off; off += (size_t)n_seq * seq_len * 4;
    const size_t o_f = off; off += (size_t)n_seq * seq_len * 4;
    off = (off + 255) & ~(size_t)255;
    const size_t o_qc = off; off += nq * 4;
    const size_t o_qf = off; off += nq * 4;
    if (off > h->ws_bytes) {
        if (h->ws) cudaFree(h->ws);
        h->ws = nullptr; h->ws_bytes = 0;
        DM_CUDA(h, cudaMalloc(&h->ws, off + off / 8));
        h->ws_bytes = off + off / 8;
    }
    char *ws = (char *)h->ws;
    if (nq) {
        DM_CUDA(h, cudaMemcpyAsync(ws + o_qc, q_coarse, nq * 4, cudaMemcpyHostToDevice, st));
        DM_CUDA(h, cudaMemcpyAsync(ws + o_qf, q_fine, nq * 4, cudaMemcpyHostToDevice, st));
    }
    const void *kernels[4] = {(const void *)dm_loop_kernel<1>, (const void *)dm_loop_kernel<2>, (const void *)dm_loop_kernel<4>,
                              (const void *)dm_loop_kernel<8>};
    for (const void *k : kernels) DM_CUDA(h, cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->smem_bytes));
    h->launches = 0;
    DM_CUDA(h, cudaEventRecord(h->ev[0], st));
    for (int n0 = 0; n0 < n_seq; n0 += DM_MAXN) {          // sequences are independent: groups of <= 8 per launch
        DmParams P = h->plan;
        P.N = std::min(DM_MAXN, n_seq - n0);
        P.T = seq_len;
        P.image = h->image;
        DM_CUDA(h, cudaMemsetAsync(ws, 0, act_end, st));     // barrier counter, abort flag, hidden = 0 (:167-169)
        P.gs.counter = (unsigned long long *)ws;
        P.gs.abort_flag = (int *)(ws + 64);
        P.hidden = (float *)(ws + o_hidden);
        P.o1 = (float *)(ws + o_o1); P.o3 = (float *)(ws + o_o3);
        P.lc = (float *)(ws + o_lc); P.lf = (float *)(ws + o_lf);
        P.qc = nq ? (const float *)(ws + o_qc) : nullptr;
        P.qf = nq ? (const float *)(ws + o_qf) : nullptr;
        P.Ntot = n_seq; P.n0 = n0;
        P.coarse = (int32_t *)(ws + o_c) + (size_t)n0 * seq_len;
        P.fine = (int32_t *)(ws + o_f) + (size_t)n0 * seq_len;
        P.seed = seed;
        P.prof = nullptr;
        long long *d_prof = nullptr;
        if (getenv("WRNN_DM_PROF")) {
            DM_CUDA(h, cudaMalloc((void **)&d_prof, 16 * sizeof(long long)));
            DM_CUDA(h, cudaMemsetAsync(d_prof, 0, 16 * sizeof(long long), st));
            P.prof = d_prof;
        }
        void *args[] = {(void *)&P};
        DM_CUDA(h, cudaLaunchCooperativeKernel(kernels[P.N <= 1 ? 0 : P.N <= 2 ? 1 : P.N <= 4 ? 2 : 3], dim3(h->num_sms), dim3(DM_THREADS), args, h->smem_bytes, st));
        h->launches++;
        int flag = 0;
        DM_CUDA(h, cudaMemcpyAsync(&flag, ws + 64, sizeof(int), cudaMemcpyDeviceToHost, st));
        DM_CUDA(h, cudaStreamSynchronize(st));
        if (d_prof) {
            long long pr[16];
            cudaMemcpy(pr, d_prof, sizeof pr, cudaMemcpyDeviceToHost);
            cudaFree(d_prof);
            static const char *site[14] = {"stage lf", "sample f", "A gates", "A barrier", "B O1 products", "B barrier", "C O2", "C barrier",
                                           "D stage + sample c + gates", "D barrier", "E O3", "E barrier + R.h 1/2", "F O4", "F barrier + R.h 2/2"};
            fprintf(stderr, "dm prof, %d sequences (clk per step, CTA 0)\n", P.N);
            for (int i = 0; i < 14; i++) fprintf(stderr, "  %-28s %7.0f\n", site[i], (double)pr[i] / (double)seq_len);
            fprintf(stderr, "  B load of hidden_coarse      %7.0f\n", (double)pr[14] / (double)seq_len);
        }
        if (flag) return dm_fail(h, WRNN_ERR_CUDA, "deepmind loop kernel aborted: grid barrier watchdog fired");
    }
    DM_CUDA(h, cudaEventRecord(h->ev[1], st));
    std::vector<int32_t> hc((size_t)n_seq * seq_len), hf((size_t)n_seq * seq_len);
    DM_CUDA(h, cudaMemcpyAsync(hc.data(), ws + o_c, hc.size() * 4, cudaMemcpyDeviceToHost, st));
    DM_CUDA(h, cudaMemcpyAsync(hf.data(), ws + o_f, hf.size() * 4, cudaMemcpyDeviceToHost, st));
    DM_CUDA(h, cudaStreamSynchronize(st));
    DM_CUDA(h, cudaGetLastError());
    for (size_t i = 0; i < hc.size(); i++) {
        coarse[i] = hc[i];
        fine[i] = hf[i];
        output[i] = (int64_t)hc[i] * 256 + hf[i] - 32768;      // combine_signal (utils/dsp.py:33-34)
    }
    return 0;
}

int wrnn_dm_last_timing(wrnn_dm_handle *h, float *ms_loop, int64_t *kernel_launches) {
    if (!h) return WRNN_ERR_ARG;
    float ms = 0;
    DM_CUDA(h, cudaEventSynchronize(h->ev[1]));
    DM_CUDA(h, cudaEventElapsedTime(&ms, h->ev[0], h->ev[1]));
    if (ms_loop) *ms_loop = ms;
    if (kernel_launches) *kernel_launches = h->launches;
    return 0;
}

}  // extern "C"

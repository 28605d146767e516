// Per-sample loop, phase-pipelined fp32 kernel ("pipe"): the fast path for fold groups of <= 20 folds.
// One persistent cooperative launch runs all S steps of `generate`'s hot loop
// (models/fatchord_version.py:201-241) with every weight resident in shared memory.
//
// Roles.  "Row" CTAs (NR of them, 128 at the default dims) own `upc` GRU units of both GRUs and `rF`
// rows of fc1/fc2.  "Tail" CTAs (the other SMs, 20 at the default dims) each own whole folds: they
// hold all of fc3 and do fc3 + sampling for their folds.
//
// What leaves the dependency chain (compared with the reference's order of operations):
//   * I followed by rnn1's input product is one matrix: W' = W_ih1 . W_I (3H x (1+F+A)), composed on the
//     host in double.  Only its column 0 multiplies the previous sample, so everything else of rnn1's
//     input side -- and of xI = I([x, m_t, a1_t]) for the residual -- is computed one step ahead.
//   * Both W_hh . h products only need the previous step's hidden state: they are computed in the
//     latency shadow of the exchanges ("shadow" jobs below).
// Chain per step: x_prev -> rnn1 gates -> x1 -> rnn2 (W_ih2 product, gates) -> x2 -> fc1 -> y1 -> fc2 -> y2
// -> (tail) fc3 + sample -> x_prev : five exchanges.
//
// Exchange.  Activations travel through L2-resident global buffers [2][B][K] (double-buffered by step
// parity).  There is no barrier and no flag: every 4-byte word validates itself.  The producer rounds
// the fp32 value to 22 mantissa bits (nearest) and stores (t+1)&3 in the two freed low bits; a consumer
// polls the words it needs (ld.relaxed.gpu) until all carry the tag of the step it is in, then clears
// the bits.  A slot of parity p last held step t-2, whose tag differs, and zero-initialised buffers carry
// tag 0 which step 0 / step 1 never use.  The previous sample is exchanged exactly: an 8-byte word
// {fp32 bits, step + 1}.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "internal.h"
#include "loop_common.cuh"
#include "exchange.cuh"

namespace wrnn {

constexpr int PT = 512;          // threads per CTA
constexpr int PW = PT / 32;
constexpr int MAXB = 20;         // folds per launch
constexpr int FT = 5;            // folds per thread tile
constexpr int MAXU = 4;          // GRU units / fc rows per CTA
constexpr int PART_FLOATS = PW * 16 * 20;   // [warp][tile<=16][<=20 accumulators]

struct PipeParams {
    Dims d;
    int B;                       // folds in this launch (<= MAXB)
    int64_t S;
    int G, NR, NT, upc, rF;
    int Kc, ldc;                 // F + A conditioning inputs of the I layer (column 0 = x_prev is separate)
    int KH, ldh;                 // H
    int KX1, ldx1;               // H + A  (rnn2 input, fc1 input)
    int KX2, ldx2;               // FC + A (fc2 input)
    int ldx;                     // max(ldx1, ldx2)
    int KF, ldf3;                // FC (fc3 input)
    int off_Wc, off_w0, off_hh1, off_ih2, off_hh2, off_f1, off_f2, off_bias, img_floats;
    int tail_floats;
    const float *row_image;      // [NR][img_floats]
    const float *tail_image;     // [NC][ldf3] | bias[NC]
    LoopArgs a;
    float *x1, *h1, *x2, *h2, *y1, *y2;      // [2][B][H] / [2][B][FC], tagged words
    unsigned long long *xprev;               // [2][MAXB] {bits, step + 1}
    int *abort_flag;
    long long *prof;             // optional [2][16] clock accumulators (WRNN_PROF=1), else null
};
#define PROF_MARK(slot)                                                     \
    do {                                                                    \
        if (prof_on) {                                                      \
            const long long now__ = clock64();                              \
            prof_acc[slot] += now__ - prof_t;                               \
            prof_t = now__;                                                 \
        }                                                                   \
    } while (0)


// Gather a tagged vector src[nrow][K] (K % 4 == 0) into smem dst[row][0..K) with row stride ld.
// Every thread owns float4 slots tid, tid + PT, ...; it probes its first slot until valid, then reads
// the rest (at the default dims all slots of a thread come from the same producer CTA).
// Returns false if the watchdog fired.
template <int MAXSLOT>
__device__ __forceinline__ bool gather_tagged(const float *src, int nrow, int K, float *dst, int ld, uint32_t tag,
                                              const int *abort_flag, int tid) {
    const int K4 = K >> 2, n4 = nrow * K4;
    if (tid >= n4) return true;
    const uint4 *s4 = reinterpret_cast<const uint4 *>(src);
    uint4 v[MAXSLOT];
    bool ok = true;
    v[0] = ld_relaxed_v4(s4 + tid);
    if (!tags_ok(v[0], tag)) {
        const long long t0 = clock64();
        unsigned spins = 0;
        do {
            v[0] = ld_relaxed_v4(s4 + tid);
            if ((++spins & 255u) == 0u && (clock64() - t0 > WATCHDOG_CLOCKS || ld_volatile_s32(abort_flag))) { ok = false; break; }
        } while (!tags_ok(v[0], tag));
    }
#pragma unroll
    for (int j = 1; j < MAXSLOT; j++) {
        const int i = tid + j * PT;
        if (i < n4) v[j] = ld_relaxed_v4(s4 + i);
    }
#pragma unroll
    for (int j = 1; j < MAXSLOT; j++) {
        const int i = tid + j * PT;
        if (i < n4 && ok && !tags_ok(v[j], tag)) {
            const long long t0 = clock64();
            unsigned spins = 0;
            do {
                v[j] = ld_relaxed_v4(s4 + i);
                if ((++spins & 255u) == 0u && (clock64() - t0 > WATCHDOG_CLOCKS || ld_volatile_s32(abort_flag))) { ok = false; break; }
            } while (!tags_ok(v[j], tag));
        }
    }
#pragma unroll
    for (int j = 0; j < MAXSLOT; j++) {
        const int i = tid + j * PT;
        if (i < n4) {
            const int r = i / K4, k4 = i - r * K4;
            float4 o;
            o.x = untag(v[j].x); o.y = untag(v[j].y); o.z = untag(v[j].z); o.w = untag(v[j].w);
            *reinterpret_cast<float4 *>(dst + (size_t)r * ld + 4 * k4) = o;
        }
    }
    return ok;
}

// Copy `n` floats (n % 4 == 0) per fold from the conditioning rows (read-only) into smem columns.
__device__ __forceinline__ void load_cond_cols(const float *cond, int64_t fold_stride, int B, int n, float *dst,
                                               int ld, int col0, int tid) {
    const int n4 = n >> 2;
    for (int i = tid; i < B * n4; i += PT) {
        const int f = i / n4, k4 = i - f * n4;
        const float4 v = __ldg(reinterpret_cast<const float4 *>(cond + (size_t)f * fold_stride) + k4);
        *reinterpret_cast<float4 *>(dst + (size_t)f * ld + col0 + 4 * k4) = v;
    }
}

// ---------------------------------------------------------------------------- dense job
// part[warp][tile][RT*FT] += W[rows of tile][k] * X[folds of tile][k] over this thread's K share.
// Lanes of a warp = tiles x k-sub-splits (so smem reads are broadcasts), warps = further K splits.
// Rows beyond the CTA's real row count are zero rows of the image; fold rows >= B are zero rows of X.
struct JobMap {
    int ntiles, tiles_p2, nksub, FG;
};
__device__ __forceinline__ JobMap job_map(int RG, int FG) {
    JobMap m;
    m.FG = FG;
    m.ntiles = RG * FG;
    int p2 = 1;
    while (p2 < m.ntiles) p2 <<= 1;
    m.tiles_p2 = p2;
    m.nksub = 32 / p2;
    return m;
}

template <int RT>
__device__ __forceinline__ void dense_job(const float *W, int ldw, const float *X, int ldx, int K4, const JobMap m,
                                          float *part, int warp, int lane) {
    const int tile = lane & (m.tiles_p2 - 1), ksub = lane / m.tiles_p2;
    const bool active = tile < m.ntiles;
    const int rg = tile / m.FG, fg = tile - rg * m.FG;
    const int ks = warp * m.nksub + ksub, nks = PW * m.nksub;
    float acc[RT][FT];
#pragma unroll
    for (int i = 0; i < RT; i++)
#pragma unroll
        for (int j = 0; j < FT; j++) acc[i][j] = 0.f;
    if (active) {
        const float4 *W4 = reinterpret_cast<const float4 *>(W + (size_t)(rg * RT) * ldw);
        const float4 *X4 = reinterpret_cast<const float4 *>(X + (size_t)(fg * FT) * ldx);
        const int ldw4 = ldw >> 2, ldx4 = ldx >> 2;
        for (int k4 = ks; k4 < K4; k4 += nks) {
            float4 w[RT], x[FT];
#pragma unroll
            for (int i = 0; i < RT; i++) w[i] = W4[i * ldw4 + k4];
#pragma unroll
            for (int j = 0; j < FT; j++) x[j] = X4[j * ldx4 + k4];
#pragma unroll
            for (int i = 0; i < RT; i++)
#pragma unroll
                for (int j = 0; j < FT; j++) {
                    float a = acc[i][j];
                    a = fmaf(w[i].x, x[j].x, a);
                    a = fmaf(w[i].y, x[j].y, a);
                    a = fmaf(w[i].z, x[j].z, a);
                    a = fmaf(w[i].w, x[j].w, a);
                    acc[i][j] = a;
                }
        }
    }
    // fold the k-sub-splits of this warp (fixed order), then lane group 0 stores the warp's partial
    for (int o = m.tiles_p2; o < 32; o <<= 1) {
#pragma unroll
        for (int i = 0; i < RT; i++)
#pragma unroll
            for (int j = 0; j < FT; j++) acc[i][j] += __shfl_xor_sync(0xffffffffu, acc[i][j], o);
    }
    if (active && ksub == 0) {
        float *dst = part + ((size_t)warp * 16 + tile) * (RT * FT);
#pragma unroll
        for (int i = 0; i < RT; i++)
#pragma unroll
            for (int j = 0; j < FT; j++) dst[i * FT + j] = acc[i][j];
    }
}

// Sum over warps (ascending: fixed order) of output (row r, fold f) of a job with row tiles of RT.
template <int RT>
__device__ __forceinline__ float job_out(const float *part, const JobMap m, int r, int f) {
    const int rg = r / RT, i = r - rg * RT, fg = f / FT, j = f - fg * FT;
    const float *p = part + (size_t)(rg * m.FG + fg) * (RT * FT) + i * FT + j;
    float s = p[0];
#pragma unroll
    for (int w = 1; w < PW; w++) s += p[(size_t)w * 16 * (RT * FT)];
    return s;
}

// ---------------------------------------------------------------------------- tail CTA: fc3 + sampling
__device__ __forceinline__ float mol_sample_smem(const float *logits, int nmix, float myu1, float u2, int lane, int *pick) {
    float t = -INFINITY;
    if (lane < nmix) t = logits[lane] - logf(-logf(myu1));
    int idx = lane;
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) {
        const float ot = __shfl_xor_sync(0xffffffffu, t, o);
        const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
        if (ot > t || (ot == t && oi < idx)) { t = ot; idx = oi; }
    }
    const float mean = logits[nmix + idx];
    const float ls = fmaxf(logits[2 * nmix + idx], -32.23619130191664f);   // log(1e-14)
    float x = mean + expf(ls) * (logf(u2) - logf(1.f - u2));
    x = fminf(fmaxf(x, -1.f), 1.f);
    *pick = idx;
    return x;
}

__device__ void tail_role(const PipeParams &P, int j, float *smem) {
    const Dims &d = P.d;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int NC = d.NC, KF = P.KF, ldf3 = P.ldf3, B = P.B;
    const int64_t S = P.a.S;
    if (j >= B) return;                               // no fold to own
    float *w3 = smem;                                 // [NC][ldf3]
    float *b3 = w3 + (size_t)NC * ldf3;               // [NC] (+pad)
    float *ybuf = b3 + ((NC + 3) & ~3);               // [ldf3]
    float *lg = ybuf + ldf3;                          // [NC]
    for (int i = tid; i < P.tail_floats / 4; i += PT)
        reinterpret_cast<float4 *>(w3)[i] = __ldg(reinterpret_cast<const float4 *>(P.tail_image) + i);
    __syncthreads();
    const uint2 key = make_uint2((uint32_t)P.a.seed, (uint32_t)(P.a.seed >> 32));
    const int nmix = NC / 3;
    const int64_t Bt = P.a.Btot, b0 = P.a.b0;
    const bool prof_on = P.prof != nullptr && j == 0 && tid == 0;
    long long prof_acc[16] = {0}, prof_t = clock64();

    for (int64_t t = 0; t <= S; t++) {
        for (int f = j; f < B; f += P.NT) {
            float xv = 0.f;
            int pick = 0;
            if (t > 0) {
                const int64_t tp = t - 1;
                const uint32_t tag = (uint32_t)((tp + 1) & 3);
                const float *src = P.y2 + ((size_t)(tp & 1) * B + f) * KF;
                const bool ok = gather_tagged<1>(src, 1, KF, ybuf, ldf3, tag, P.abort_flag, tid);
                if (__syncthreads_or(!ok)) {
                    if (tid == 0) atomicExch(P.abort_flag, 1);
                    return;
                }
                PROF_MARK(0);
                // fc3: 16 lanes per row, two rows per warp, row blocks of 32
                for (int rb = 0; rb < NC; rb += 2 * PW) {
                    const int r = rb + (tid >> 4), ks = tid & 15;
                    float acc = 0.f;
                    if (r < NC) {
                        const float4 *W4 = reinterpret_cast<const float4 *>(w3 + (size_t)r * ldf3);
                        const float4 *Y4 = reinterpret_cast<const float4 *>(ybuf);
                        for (int k4 = ks; k4 < (KF >> 2); k4 += 16) {
                            const float4 w = W4[k4], y = Y4[k4];
                            acc = fmaf(w.x, y.x, acc);
                            acc = fmaf(w.y, y.y, acc);
                            acc = fmaf(w.z, y.z, acc);
                            acc = fmaf(w.w, y.w, acc);
                        }
                    }
#pragma unroll
                    for (int o = 8; o >= 1; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
                    if (r < NC && ks == 0) {
                        const float v = acc + b3[r];
                        lg[r] = v;
                        if (P.a.logits_out && tp < P.a.fold_len[f])
                            P.a.logits_out[((size_t)f * S + tp) * NC + r] = v;
                    }
                }
                __syncthreads();
                PROF_MARK(1);
                if (warp == 0) {
                    const int64_t b = b0 + f;           // global fold index: noise / Philox addressing
                    if (d.mode == WRNN_MODE_MOL) {
                        float myu1 = 0.5f, u2;
                        if (P.a.u1) {
                            if (lane < nmix) myu1 = P.a.u1[((size_t)tp * Bt + b) * nmix + lane];
                            u2 = P.a.u2[(size_t)tp * Bt + b];
                        } else {
                            const uint4 r = philox4x32(make_uint4((uint32_t)tp, (uint32_t)b, (uint32_t)(lane >> 2), 1u), key);
                            const uint32_t rr[4] = {r.x, r.y, r.z, r.w};
                            myu1 = u_clip(rr[lane & 3]);
                            const uint4 r2 = philox4x32(make_uint4((uint32_t)tp, (uint32_t)b, 0u, 2u), key);
                            u2 = u_clip(r2.x);
                        }
                        xv = mol_sample_smem(lg, nmix, myu1, u2, lane, &pick);
                    } else {
                        const float *qsrc = P.a.q ? P.a.q + ((size_t)tp * Bt + b) * NC : nullptr;
                        pick = categorical_warp(lg, NC, qsrc, key, (uint32_t)tp, (uint32_t)b, 3u, lane);
                        xv = raw_value(pick, NC);
                    }
                    if (tp >= P.a.fold_len[f]) xv = 0.f;
                    else if (lane == 0) {
                        P.a.samples[(size_t)f * S + tp] = xv;
                        P.a.picks[(size_t)f * S + tp] = pick;
                    }
                }
            }
            if (warp == 0 && lane == 0 && t < S) {
                if (P.a.teacher) xv = P.a.teacher[(size_t)f * S + t];
                st_relaxed_u64(P.xprev + (size_t)(t & 1) * MAXB + f,
                               ((unsigned long long)(uint32_t)(t + 1) << 32) | (unsigned long long)__float_as_uint(xv));
            }
            __syncthreads();       // ybuf / lg are reused by the next fold
            PROF_MARK(2);
        }
    }
    if (prof_on)
        for (int i = 0; i < 16; i++) P.prof[16 + i] = prof_acc[i];
}

// ---------------------------------------------------------------------------- row CTA
__device__ void row_role(const PipeParams &P, int cta, float *smem) {
    const Dims &d = P.d;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int H = d.H, FC = d.FC, F = d.F, A = d.A, B = P.B, upc = P.upc, rF = P.rF;
    const int64_t S = P.a.S;
    const int CW = F + d.R;                               // conditioning row width
    const int64_t cstride = S * CW;                       // between folds

    float *wimg = smem;
    float *xbuf = wimg + P.img_floats;                    // [MAXB][ldx]
    float *hbuf = xbuf + (size_t)MAXB * P.ldx;            // [MAXB][ldh]
    float *cbuf = hbuf + (size_t)MAXB * P.ldh;            // [MAXB][ldc]
    float *part = cbuf + (size_t)MAXB * P.ldc;            // [PW][16][20]
    float *pre_i1 = part + PART_FLOATS;                   // [4*MAXU][MAXB]  rnn1 input side + xI, conditioning part
    float *pre_h1 = pre_i1 + 4 * MAXU * MAXB;             // [3*MAXU][MAXB]  W_hh1 . h1 + b_hh1
    float *pre_h2 = pre_h1 + 3 * MAXU * MAXB;             // [3*MAXU][MAXB]
    float *hst1 = pre_h2 + 3 * MAXU * MAXB;               // [MAXU][MAXB]    own units' hidden state
    float *hst2 = hst1 + MAXU * MAXB;
    float *xp = hst2 + MAXU * MAXB;                       // [32] previous sample per fold

    {   // weight image + zeroed buffers
        const float4 *src = reinterpret_cast<const float4 *>(P.row_image + (size_t)cta * P.img_floats);
        for (int i = tid; i < P.img_floats / 4; i += PT) reinterpret_cast<float4 *>(wimg)[i] = __ldg(src + i);
        const int nz = (int)(xp + 32 - xbuf);
        for (int i = tid; i < nz; i += PT) xbuf[i] = 0.f;
    }
    __syncthreads();
    const float *Wc = wimg + P.off_Wc, *w0 = wimg + P.off_w0, *Whh1 = wimg + P.off_hh1, *Wih2 = wimg + P.off_ih2,
                *Whh2 = wimg + P.off_hh2, *Wf1 = wimg + P.off_f1, *Wf2 = wimg + P.off_f2;
    const float *b1c = wimg + P.off_bias, *bhh1 = b1c + 4 * upc, *bih2 = bhh1 + 3 * upc, *bhh2 = bih2 + 3 * upc,
                *bf1 = bhh2 + 3 * upc, *bf2 = bf1 + rF;
    const int u0 = cta * upc, nu = max(0, min(upc, H - u0));
    const int r0 = cta * rF, nf = max(0, min(rF, FC - r0));
    const int FG = (B + FT - 1) / FT;
    const JobMap mg = job_map((3 * upc + 2) / 3, FG);         // GRU matrices: row tiles of 3
    const JobMap mc = job_map((4 * upc + 3) / 4, FG);         // conditioning matrix: row tiles of 4
    const JobMap mf = job_map(rF, FG);                        // fc1 / fc2: row tiles of 1
    const bool vec_u = (upc == 4), vec_f = (rF == 4);

    // pre_h1 / pre_h2 at t = 0: h = 0, so W_hh . h + b_hh = b_hh (fatchord_version.py:194-195)
    for (int i = tid; i < 3 * upc * B; i += PT) {
        const int r = i / B, f = i - r * B;
        pre_h1[r * MAXB + f] = bhh1[r];
        pre_h2[r * MAXB + f] = bhh2[r];
    }
    // conditioning part of rnn1's input side for step 0
    load_cond_cols(P.a.cond, cstride, B, P.Kc, cbuf, P.ldc, 0, tid);
    __syncthreads();
    dense_job<4>(Wc, P.ldc, cbuf, P.ldc, P.Kc >> 2, mc, part, warp, lane);
    __syncthreads();
    for (int i = tid; i < 4 * upc * B; i += PT) {
        const int r = i / B, f = i - r * B;
        pre_i1[r * MAXB + f] = job_out<4>(part, mc, r, f) + b1c[r];
    }
    __syncthreads();

    bool aborted = false;
    const bool prof_on = P.prof != nullptr && cta == 0 && tid == 0;
    long long prof_acc[16] = {0}, prof_t = clock64();
    for (int64_t t = 0; t < S; t++) {
        const uint32_t tag = (uint32_t)((t + 1) & 3);
        const size_t par = (size_t)(t & 1);
        const float *cond_t = P.a.cond + (size_t)t * CW;
        bool ok = true;

        // ================================================================ R1: previous sample -> rnn1 gates
        if (tid < B) {
            const unsigned long long *slot = P.xprev + par * MAXB + tid;
            unsigned long long v = ld_relaxed_u64(slot);
            if ((uint32_t)(v >> 32) != (uint32_t)(t + 1)) {
                const long long t0 = clock64();
                unsigned spins = 0;
                do {
                    v = ld_relaxed_u64(slot);
                    if ((++spins & 255u) == 0u && (clock64() - t0 > WATCHDOG_CLOCKS || ld_volatile_s32(P.abort_flag))) { ok = false; break; }
                } while ((uint32_t)(v >> 32) != (uint32_t)(t + 1));
            }
            xp[tid] = __uint_as_float((uint32_t)v);
        }
        if (__syncthreads_or(!ok)) { aborted = true; break; }
        PROF_MARK(0);
        if (tid < B && nu > 0) {
            const int f = tid;
            const float xpv = xp[f];
            uint32_t hw[MAXU], xw[MAXU];
#pragma unroll
            for (int u = 0; u < MAXU; u++) {
                hw[u] = xw[u] = tag;
                if (u < nu) {
                    const float ir = fmaf(w0[0 * upc + u], xpv, pre_i1[(0 * upc + u) * MAXB + f]);
                    const float iz = fmaf(w0[1 * upc + u], xpv, pre_i1[(1 * upc + u) * MAXB + f]);
                    const float in_ = fmaf(w0[2 * upc + u], xpv, pre_i1[(2 * upc + u) * MAXB + f]);
                    const float xI = fmaf(w0[3 * upc + u], xpv, pre_i1[(3 * upc + u) * MAXB + f]);
                    const float hn = gru_update(ir, iz, in_, pre_h1[(0 * upc + u) * MAXB + f], pre_h1[(1 * upc + u) * MAXB + f],
                                                pre_h1[(2 * upc + u) * MAXB + f], hst1[u * MAXB + f]);
                    hw[u] = tag_word(hn, tag);
                    const float hr = untag(hw[u]);
                    hst1[u * MAXB + f] = hr;
                    xw[u] = tag_word(xI + hr, tag);            // x = x + h1 (:212)
                }
            }
            float *ph = P.h1 + (par * B + f) * H + u0, *px = P.x1 + (par * B + f) * H + u0;
            if (vec_u) {
                st_relaxed_v4(px, make_uint4(xw[0], xw[1], xw[2], xw[3]));
                st_relaxed_v4(ph, make_uint4(hw[0], hw[1], hw[2], hw[3]));
            } else {
                for (int u = 0; u < nu; u++) { st_relaxed_u32(px + u, xw[u]); st_relaxed_u32(ph + u, hw[u]); }
            }
        }
        // ---- shadow A: conditioning part of rnn1's input side for step t + 1
        if (t + 1 < S) {
            load_cond_cols(cond_t + CW, cstride, B, P.Kc, cbuf, P.ldc, 0, tid);
            __syncthreads();                                   // also orders the gate reads of pre_i1 above
            dense_job<4>(Wc, P.ldc, cbuf, P.ldc, P.Kc >> 2, mc, part, warp, lane);
            __syncthreads();
            for (int i = tid; i < 4 * upc * B; i += PT) {
                const int r = i / B, f = i - r * B;
                pre_i1[r * MAXB + f] = job_out<4>(part, mc, r, f) + b1c[r];
            }
        }

        PROF_MARK(1);
        // ================================================================ R2: x1 -> rnn2
        load_cond_cols(cond_t + F + A, cstride, B, A, xbuf, P.ldx1, H, tid);            // a2_t
        ok = gather_tagged<5>(P.x1 + par * B * H, B, H, xbuf, P.ldx1, tag, P.abort_flag, tid);
        if (__syncthreads_or(!ok)) { aborted = true; break; }
        PROF_MARK(2);
        dense_job<3>(Wih2, P.ldx1, xbuf, P.ldx1, P.KX1 >> 2, mg, part, warp, lane);
        __syncthreads();
        PROF_MARK(3);
        if (tid < B && nu > 0) {
            const int f = tid;
            uint32_t hw[MAXU], xw[MAXU];
#pragma unroll
            for (int u = 0; u < MAXU; u++) {
                hw[u] = xw[u] = tag;
                if (u < nu) {
                    const float ir = job_out<3>(part, mg, 0 * upc + u, f) + bih2[0 * upc + u];
                    const float iz = job_out<3>(part, mg, 1 * upc + u, f) + bih2[1 * upc + u];
                    const float in_ = job_out<3>(part, mg, 2 * upc + u, f) + bih2[2 * upc + u];
                    const float hn = gru_update(ir, iz, in_, pre_h2[(0 * upc + u) * MAXB + f], pre_h2[(1 * upc + u) * MAXB + f],
                                                pre_h2[(2 * upc + u) * MAXB + f], hst2[u * MAXB + f]);
                    hw[u] = tag_word(hn, tag);
                    const float hr = untag(hw[u]);
                    hst2[u * MAXB + f] = hr;
                    xw[u] = tag_word(xbuf[(size_t)f * P.ldx1 + u0 + u] + hr, tag);      // x = x + h2 (:216)
                }
            }
            float *ph = P.h2 + (par * B + f) * H + u0, *px = P.x2 + (par * B + f) * H + u0;
            if (vec_u) {
                st_relaxed_v4(px, make_uint4(xw[0], xw[1], xw[2], xw[3]));
                st_relaxed_v4(ph, make_uint4(hw[0], hw[1], hw[2], hw[3]));
            } else {
                for (int u = 0; u < nu; u++) { st_relaxed_u32(px + u, xw[u]); st_relaxed_u32(ph + u, hw[u]); }
            }
        }
        PROF_MARK(4);
        // ---- shadow B: W_hh1 . h1(t) for step t + 1
        ok = gather_tagged<5>(P.h1 + par * B * H, B, H, hbuf, P.ldh, tag, P.abort_flag, tid);
        if (__syncthreads_or(!ok)) { aborted = true; break; }                       // also: part is free again
        PROF_MARK(5);
        dense_job<3>(Whh1, P.ldh, hbuf, P.ldh, P.KH >> 2, mg, part, warp, lane);
        __syncthreads();
        for (int i = tid; i < 3 * upc * B; i += PT) {
            const int r = i / B, f = i - r * B;
            pre_h1[r * MAXB + f] = job_out<3>(part, mg, r, f) + bhh1[r];
        }

        PROF_MARK(6);
        // ================================================================ F1: x2 -> fc1 + ReLU
        load_cond_cols(cond_t + F + 2 * A, cstride, B, A, xbuf, P.ldx1, H, tid);        // a3_t
        ok = gather_tagged<5>(P.x2 + par * B * H, B, H, xbuf, P.ldx1, tag, P.abort_flag, tid);
        if (__syncthreads_or(!ok)) { aborted = true; break; }
        PROF_MARK(7);
        dense_job<1>(Wf1, P.ldx1, xbuf, P.ldx1, P.KX1 >> 2, mf, part, warp, lane);
        __syncthreads();
        if (tid < B && nf > 0) {
            const int f = tid;
            uint32_t yw[MAXU];
#pragma unroll
            for (int r = 0; r < MAXU; r++) {
                yw[r] = tag;
                if (r < nf) yw[r] = tag_word(fmaxf(job_out<1>(part, mf, r, f) + bf1[r], 0.f), tag);
            }
            float *py = P.y1 + (par * B + f) * FC + r0;
            if (vec_f) st_relaxed_v4(py, make_uint4(yw[0], yw[1], yw[2], yw[3]));
            else for (int r = 0; r < nf; r++) st_relaxed_u32(py + r, yw[r]);
        }
        __syncthreads();

        PROF_MARK(8);
        // ================================================================ F2: y1 -> fc2 + ReLU
        load_cond_cols(cond_t + F + 3 * A, cstride, B, A, xbuf, P.ldx2, FC, tid);       // a4_t
        ok = gather_tagged<5>(P.y1 + par * B * FC, B, FC, xbuf, P.ldx2, tag, P.abort_flag, tid);
        if (__syncthreads_or(!ok)) { aborted = true; break; }
        PROF_MARK(9);
        dense_job<1>(Wf2, P.ldx2, xbuf, P.ldx2, P.KX2 >> 2, mf, part, warp, lane);
        __syncthreads();
        if (tid < B && nf > 0) {
            const int f = tid;
            uint32_t yw[MAXU];
#pragma unroll
            for (int r = 0; r < MAXU; r++) {
                yw[r] = tag;
                if (r < nf) yw[r] = tag_word(fmaxf(job_out<1>(part, mf, r, f) + bf2[r], 0.f), tag);
            }
            float *py = P.y2 + (par * B + f) * FC + r0;
            if (vec_f) st_relaxed_v4(py, make_uint4(yw[0], yw[1], yw[2], yw[3]));
            else for (int r = 0; r < nf; r++) st_relaxed_u32(py + r, yw[r]);
        }
        PROF_MARK(10);
        // ---- shadow C: W_hh2 . h2(t) for step t + 1 (in the shadow of fc3 + sampling)
        ok = gather_tagged<5>(P.h2 + par * B * H, B, H, hbuf, P.ldh, tag, P.abort_flag, tid);
        if (__syncthreads_or(!ok)) { aborted = true; break; }
        PROF_MARK(11);
        dense_job<3>(Whh2, P.ldh, hbuf, P.ldh, P.KH >> 2, mg, part, warp, lane);
        __syncthreads();
        for (int i = tid; i < 3 * upc * B; i += PT) {
            const int r = i / B, f = i - r * B;
            pre_h2[r * MAXB + f] = job_out<3>(part, mg, r, f) + bhh2[r];
        }
        __syncthreads();
        PROF_MARK(12);
    }
    if (prof_on)
        for (int i = 0; i < 16; i++) P.prof[i] = prof_acc[i];
    if (aborted && tid == 0) atomicExch(P.abort_flag, 1);        // make every other CTA leave its poll loop
}

__global__ void __launch_bounds__(PT, 1) pipe_loop_kernel(const PipeParams P) {
    extern __shared__ __align__(16) float smem[];
    const int cta = blockIdx.x;
    if (cta < P.NR) row_role(P, cta, smem);
    else tail_role(P, cta - P.NR, smem);
}

// ============================================================================ host: plan + launch
static inline int ld_pad(int k) {        // smallest ld >= k with ld % 16 == 8: conflict-free row tiles
    int ld = (k + 3) & ~3;
    while (ld % 16 != 8) ld += 4;
    return ld;
}

bool pipe_supported(const wrnn_handle *h) {
    const Dims &d = h->d;
    if (d.mode != WRNN_MODE_MOL) return false;                       // fc3 must fit one CTA (RAW: fma kernel)
    if (d.H % 4 || d.FC % 4 || d.A % 4 || (d.F + d.A) % 4 || d.NC > 64) return false;
    return h->pipe.ok;
}

int pipe_plan(wrnn_handle *h) {
    const Dims &d = h->d;
    PipePlan &p = h->pipe;
    p.ok = false;
    if (d.mode != WRNN_MODE_MOL || d.H % 4 || d.FC % 4 || d.A % 4 || (d.F + d.A) % 4 || d.NC > 64) return 0;
    const int G = h->num_sms;
    const int nt_min = G / 8 > 1 ? G / 8 : 1;
    if (G - nt_min < 1) return 0;
    p.G = G;
    p.upc = (d.H + (G - nt_min) - 1) / (G - nt_min);
    if (p.upc > MAXU || MAXB * (d.H / 4) > 5 * PT || MAXB * (d.FC / 4) > 5 * PT) return 0;
    p.NR = (d.H + p.upc - 1) / p.upc;
    p.rF = (d.FC + p.NR - 1) / p.NR;
    if (p.rF > MAXU) return 0;
    p.NT = G - p.NR;
    p.Kc = d.F + d.A;  p.ldc = ld_pad(p.Kc);
    p.KH = d.H;        p.ldh = ld_pad(p.KH);
    p.KX1 = d.H + d.A; p.ldx1 = ld_pad(p.KX1);
    p.KX2 = d.FC + d.A; p.ldx2 = ld_pad(p.KX2);
    p.ldx = p.ldx1 > p.ldx2 ? p.ldx1 : p.ldx2;
    p.KF = d.FC;       p.ldf3 = ld_pad(p.KF);
    const int upc = p.upc, rF = p.rF;
    int o = 0;
    p.off_Wc = o;  o += 4 * upc * p.ldc;
    p.off_w0 = o;  o += 4 * MAXU;
    p.off_hh1 = o; o += 3 * upc * p.ldh;
    p.off_ih2 = o; o += 3 * upc * p.ldx1;
    p.off_hh2 = o; o += 3 * upc * p.ldh;
    p.off_f1 = o;  o += rF * p.ldx1;
    p.off_f2 = o;  o += rF * p.ldx2;
    p.off_bias = o; o += 4 * upc + 9 * upc + 2 * rF;
    p.img_floats = (o + 3) & ~3;
    p.tail_floats = (d.NC * p.ldf3 + ((d.NC + 3) & ~3) + 3) & ~3;
    const size_t row_smem = ((size_t)p.img_floats + (size_t)MAXB * (p.ldx + p.ldh + p.ldc) + PART_FLOATS +
                             (4 + 3 + 3 + 2) * MAXU * MAXB + 32) * sizeof(float);
    const size_t tail_smem = ((size_t)p.tail_floats + p.ldf3 + ((d.NC + 3) & ~3) + 64) * sizeof(float);
    p.smem_bytes = row_smem > tail_smem ? row_smem : tail_smem;
    if (p.smem_bytes > h->smem_optin) return 0;

    auto get = [&](const char *k, std::vector<int64_t> shape) -> const HostTensor * {
        auto it = h->w.find(k);
        if (it == h->w.end()) { fail(h, WRNN_ERR_STATE, "checkpoint key missing: %s", k); return nullptr; }
        if (it->second.shape != shape) { fail(h, WRNN_ERR_ARG, "checkpoint key %s has the wrong shape", k); return nullptr; }
        return &it->second;
    };
    const int H = d.H, FC = d.FC, NC = d.NC, A = d.A, KinI = 1 + d.F + d.A;
    const HostTensor *WI = get("I.weight", {H, KinI}), *BI = get("I.bias", {H});
    const HostTensor *W1i = get("rnn1.weight_ih_l0", {3 * H, H}), *W1h = get("rnn1.weight_hh_l0", {3 * H, H});
    const HostTensor *B1i = get("rnn1.bias_ih_l0", {3 * H}), *B1h = get("rnn1.bias_hh_l0", {3 * H});
    const HostTensor *W2i = get("rnn2.weight_ih_l0", {3 * H, H + A}), *W2h = get("rnn2.weight_hh_l0", {3 * H, H});
    const HostTensor *B2i = get("rnn2.bias_ih_l0", {3 * H}), *B2h = get("rnn2.bias_hh_l0", {3 * H});
    const HostTensor *F1 = get("fc1.weight", {FC, H + A}), *F1b = get("fc1.bias", {FC});
    const HostTensor *F2 = get("fc2.weight", {FC, FC + A}), *F2b = get("fc2.bias", {FC});
    const HostTensor *F3 = get("fc3.weight", {NC, FC}), *F3b = get("fc3.bias", {NC});
    if (!WI || !BI || !W1i || !W1h || !B1i || !B1h || !W2i || !W2h || !B2i || !B2h || !F1 || !F1b || !F2 || !F2b ||
        !F3 || !F3b)
        return WRNN_ERR_STATE;

    // W' = W_ih1 . W_I and b' = W_ih1 . b_I + b_ih1, accumulated in double (I then rnn1's input product,
    // fatchord_version.py:209-210)
    std::vector<float> Wp((size_t)3 * H * KinI), bp((size_t)3 * H);
    {
        std::vector<double> acc(KinI);
        for (int r = 0; r < 3 * H; r++) {
            std::fill(acc.begin(), acc.end(), 0.0);
            double bacc = B1i->v[r];
            const float *wr = W1i->v.data() + (size_t)r * H;
            for (int k = 0; k < H; k++) {
                const double w = wr[k];
                const float *ir = WI->v.data() + (size_t)k * KinI;
                for (int c = 0; c < KinI; c++) acc[c] += w * (double)ir[c];
                bacc += w * (double)BI->v[k];
            }
            for (int c = 0; c < KinI; c++) Wp[(size_t)r * KinI + c] = (float)acc[c];
            bp[r] = (float)bacc;
        }
    }

    std::vector<float> img((size_t)p.NR * p.img_floats, 0.f);
    for (int c = 0; c < p.NR; c++) {
        float *im = img.data() + (size_t)c * p.img_floats;
        float *bias = im + p.off_bias;
        float *b1c = bias, *bhh1 = b1c + 4 * upc, *bih2 = bhh1 + 3 * upc, *bhh2 = bih2 + 3 * upc, *bf1 = bhh2 + 3 * upc,
              *bf2 = bf1 + rF;
        for (int u = 0; u < upc; u++) {
            const int unit = c * upc + u;
            if (unit >= H) break;
            // xI row (residual input of :212): W_I[unit], conditioning columns + column 0
            memcpy(im + p.off_Wc + (size_t)(3 * upc + u) * p.ldc, WI->v.data() + (size_t)unit * KinI + 1, (size_t)(KinI - 1) * sizeof(float));
            im[p.off_w0 + 3 * upc + u] = WI->v[(size_t)unit * KinI];
            b1c[3 * upc + u] = BI->v[unit];
            for (int g = 0; g < 3; g++) {
                const int row = g * H + unit, lr = g * upc + u;
                memcpy(im + p.off_Wc + (size_t)lr * p.ldc, Wp.data() + (size_t)row * KinI + 1, (size_t)(KinI - 1) * sizeof(float));
                im[p.off_w0 + lr] = Wp[(size_t)row * KinI];
                b1c[lr] = bp[row];
                memcpy(im + p.off_hh1 + (size_t)lr * p.ldh, W1h->v.data() + (size_t)row * H, (size_t)H * sizeof(float));
                memcpy(im + p.off_ih2 + (size_t)lr * p.ldx1, W2i->v.data() + (size_t)row * (H + A), (size_t)(H + A) * sizeof(float));
                memcpy(im + p.off_hh2 + (size_t)lr * p.ldh, W2h->v.data() + (size_t)row * H, (size_t)H * sizeof(float));
                bhh1[lr] = B1h->v[row];
                bih2[lr] = B2i->v[row];
                bhh2[lr] = B2h->v[row];
            }
        }
        for (int r = 0; r < rF; r++) {
            const int row = c * rF + r;
            if (row >= FC) break;
            memcpy(im + p.off_f1 + (size_t)r * p.ldx1, F1->v.data() + (size_t)row * (H + A), (size_t)(H + A) * sizeof(float));
            memcpy(im + p.off_f2 + (size_t)r * p.ldx2, F2->v.data() + (size_t)row * (FC + A), (size_t)(FC + A) * sizeof(float));
            bf1[r] = F1b->v[row];
            bf2[r] = F2b->v[row];
        }
    }
    std::vector<float> timg((size_t)p.tail_floats, 0.f);
    for (int r = 0; r < NC; r++) {
        memcpy(timg.data() + (size_t)r * p.ldf3, F3->v.data() + (size_t)r * FC, (size_t)FC * sizeof(float));
        timg[(size_t)NC * p.ldf3 + r] = F3b->v[r];
    }
    if (p.row_image) cudaFree(p.row_image);
    if (p.tail_image) cudaFree(p.tail_image);
    p.row_image = p.tail_image = nullptr;
    WRNN_CUDA(h, cudaMalloc((void **)&p.row_image, img.size() * sizeof(float)));
    WRNN_CUDA(h, cudaMemcpy(p.row_image, img.data(), img.size() * sizeof(float), cudaMemcpyHostToDevice));
    WRNN_CUDA(h, cudaMalloc((void **)&p.tail_image, timg.size() * sizeof(float)));
    WRNN_CUDA(h, cudaMemcpy(p.tail_image, timg.data(), timg.size() * sizeof(float), cudaMemcpyHostToDevice));
    WRNN_CUDA(h, cudaFuncSetAttribute(pipe_loop_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem_bytes));
    p.ok = true;
    return 0;
}

int pipe_max_folds() { return MAXB; }

// One launch for a group of a.B <= MAXB folds.
int pipe_run(wrnn_handle *h, const LoopArgs &a, cudaStream_t st) {
    const Dims &d = h->d;
    const PipePlan &p = h->pipe;
    if (a.B < 1 || a.B > MAXB) return fail(h, WRNN_ERR_ARG, "pipe loop kernel takes 1..%d folds per launch", MAXB);
    PipeParams P;
    P.d = d;
    P.B = (int)a.B;
    P.S = a.S;
    P.G = p.G; P.NR = p.NR; P.NT = p.NT; P.upc = p.upc; P.rF = p.rF;
    P.Kc = p.Kc; P.ldc = p.ldc; P.KH = p.KH; P.ldh = p.ldh; P.KX1 = p.KX1; P.ldx1 = p.ldx1; P.KX2 = p.KX2; P.ldx2 = p.ldx2;
    P.ldx = p.ldx; P.KF = p.KF; P.ldf3 = p.ldf3;
    P.off_Wc = p.off_Wc; P.off_w0 = p.off_w0; P.off_hh1 = p.off_hh1; P.off_ih2 = p.off_ih2; P.off_hh2 = p.off_hh2;
    P.off_f1 = p.off_f1; P.off_f2 = p.off_f2; P.off_bias = p.off_bias; P.img_floats = p.img_floats;
    P.tail_floats = p.tail_floats;
    P.row_image = p.row_image;
    P.tail_image = p.tail_image;
    P.a = a;
    const size_t BH = (size_t)a.B * d.H, BF = (size_t)a.B * d.FC;
    const size_t act_floats = 2 * (4 * BH + 2 * BF);
    const size_t xprev_bytes = 2 * MAXB * sizeof(unsigned long long);
    WRNN_TRY(ws_reserve(h, h->ws_act, act_floats * sizeof(float) + xprev_bytes + 256));
    WRNN_CUDA(h, cudaMemsetAsync(h->ws_act.p, 0, act_floats * sizeof(float) + xprev_bytes + 256, st));
    float *base = (float *)h->ws_act.p;
    P.x1 = base; P.h1 = P.x1 + 2 * BH; P.x2 = P.h1 + 2 * BH; P.h2 = P.x2 + 2 * BH;
    P.y1 = P.h2 + 2 * BH; P.y2 = P.y1 + 2 * BF;
    P.xprev = (unsigned long long *)(((uintptr_t)(P.y2 + 2 * BF) + 63) & ~(uintptr_t)63);
    WRNN_TRY(ws_reserve(h, h->ws_sync, 128));                        // cleared by the caller once per run
    P.abort_flag = (int *)((char *)h->ws_sync.p + 64);
    static const bool prof_env = getenv("WRNN_PROF") != nullptr;
    P.prof = nullptr;
    if (prof_env) {
        WRNN_TRY(ws_reserve(h, h->ws_prof, 32 * sizeof(long long)));
        WRNN_CUDA(h, cudaMemsetAsync(h->ws_prof.p, 0, 32 * sizeof(long long), st));
        P.prof = (long long *)h->ws_prof.p;
    }
    WRNN_CUDA(h, cudaFuncSetAttribute(pipe_loop_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem_bytes));
    void *args[] = {(void *)&P};
    WRNN_CUDA(h, cudaLaunchCooperativeKernel((const void *)pipe_loop_kernel, dim3(p.G), dim3(PT), args, p.smem_bytes, st));
    h->launches++;
    if (prof_env) {      // debugging aid: per-segment clocks of row CTA 0 and tail CTA 0 (synchronises!)
        long long pr[32];
        WRNN_CUDA(h, cudaMemcpyAsync(pr, h->ws_prof.p, sizeof pr, cudaMemcpyDeviceToHost, st));
        WRNN_CUDA(h, cudaStreamSynchronize(st));
        static const char *rn[13] = {"wait_xprev", "gates1+shadowA", "gather_x1", "job_ih2", "gates2+pub", "gather_h1",
                                     "job_hh1+pre", "gather_x2", "job_fc1+pub", "gather_y1", "job_fc2+pub", "gather_h2",
                                     "job_hh2+pre"};
        fprintf(stderr, "[wrnn prof] row CTA 0, clocks per step:");
        for (int i = 0; i < 13; i++) fprintf(stderr, " %s=%.0f", rn[i], (double)pr[i] / (double)a.S);
        fprintf(stderr, "\n[wrnn prof] tail CTA 0, clocks per step: gather_y2=%.0f fc3=%.0f sample+pub=%.0f\n",
                (double)pr[16] / (double)a.S, (double)pr[17] / (double)a.S, (double)pr[18] / (double)a.S);
    }
    return 0;
}

}  // namespace wrnn

// Per-sample loop, fp32 FMA kernel: the exact-parity path and the path for fold counts < 16.
// One persistent cooperative kernel runs all S steps of `generate`'s hot loop
// (models/fatchord_version.py:201-241).  Every CTA (one per SM) keeps its row slice of every
// layer's weights in shared memory for the whole call; per step the layers exchange activations
// through small L2-resident global buffers separated by grid barriers.
//
// Step schedule (6 grid barriers):
//   P0  sample step t-1 from its logits (every CTA redundantly; CTA 0 records it)  +  I layer
//   P1  rnn1 (W_ih.xI, W_hh.h1, fused gates, residual)      P2  rnn2 (same, input [x, a2])
//   P3  fc1+ReLU      P4  fc2+ReLU      P5  fc3 -> logits
// Dense work is cut into warp jobs (matrix, K-split): a warp owns RT rows x 8 folds of accumulators,
// lanes stride K in float4, and a shuffle reduce-scatter leaves each lane with RT*8/32 results.
#include <stdio.h>

#include "internal.h"
#include "loop_common.cuh"

namespace wrnn {

constexpr int BT = 8;            // folds per chunk (accumulator tile width)
constexpr int FMA_THREADS = 256;
constexpr int FMA_WARPS = FMA_THREADS / 32;
constexpr int MAXQ = 8;          // max K-split

struct FmaParams {
    Dims d;
    int upc, rF, r3;
    int KI, KH, KHA, KFA, KF;
    int64_t img_floats;
    int64_t off_I, off_r1i, off_r1h, off_r2i, off_r2h, off_f1, off_f2, off_f3, off_bias;
    const float *image;
    LoopArgs a;
    // global exchange buffers
    float *xI, *x1, *x2, *y1, *y2, *h1, *h2, *logits;   // h1/h2: [2][B][H]
    GridSync gs;
    int xbuf_floats, hbuf_floats;
};

// acc[r][f] += W[r][k] * X[f][k] over this warp's K share.  W4/X4 are float4 views in smem.
template <int RT>
__device__ __forceinline__ void dense_job(const float4 *W4, int ldw4, int nrows, const float4 *X4, int ldx4,
                                          int K4, int q, int nq, int lane, float *obuf) {
    float acc[RT * BT];
#pragma unroll
    for (int i = 0; i < RT * BT; i++) acc[i] = 0.f;
    for (int k4 = q * 32 + lane; k4 < K4; k4 += 32 * nq) {
        float4 x[BT];
#pragma unroll
        for (int f = 0; f < BT; f++) x[f] = X4[f * ldx4 + k4];
#pragma unroll
        for (int r = 0; r < RT; r++) {
            if (r < nrows) {
                const float4 w = W4[r * ldw4 + k4];
#pragma unroll
                for (int f = 0; f < BT; f++) {
                    float a = acc[r * BT + f];
                    a = fmaf(w.x, x[f].x, a);
                    a = fmaf(w.y, x[f].y, a);
                    a = fmaf(w.z, x[f].z, a);
                    a = fmaf(w.w, x[f].w, a);
                    acc[r * BT + f] = a;
                }
            }
        }
    }
    warp_reduce_scatter<RT * BT>(acc, lane);
    const int base = reduce_scatter_base<RT * BT>(lane);
#pragma unroll
    for (int i = 0; i < RT * BT / 32; i++) obuf[base + i] = acc[i];
}

// Run `nmat` matrices (each nrows x K4 float4, its own input buffer) split nq ways over the warps.
// Partials land in obuf[(m*nq + q)][RT*BT]; caller sums over q in order.
struct MatJob {
    const float4 *W4;
    const float4 *X4;
    int ldw4, ldx4, K4, nrows;
};
template <int RT>
__device__ __forceinline__ void run_jobs(const MatJob *jobs, int nmat, int nq, float *obuf, int warp, int lane) {
    for (int j = warp; j < nmat * nq; j += FMA_WARPS) {
        const int m = j / nq, q = j % nq;
        dense_job<RT>(jobs[m].W4, jobs[m].ldw4, jobs[m].nrows, jobs[m].X4, jobs[m].ldx4, jobs[m].K4, q, nq, lane,
                      obuf + (size_t)j * RT * BT);
    }
}
__device__ __forceinline__ float sum_q(const float *obuf, int m, int nq, int stride, int idx) {
    float s = obuf[(size_t)(m * nq) * stride + idx];
    for (int q = 1; q < nq; q++) s += obuf[(size_t)(m * nq + q) * stride + idx];
    return s;
}


// Stage one chunk of fold rows into smem: dst[f][0..ld) = [ main row (Kmain floats, written by other
// CTAs this step -> .cg loads) | tail (ntail conditioning floats, read-only) | zero padding ].
// One warp per fold row; all global loads of a row are issued before the first store so a fill
// costs one L2 round trip instead of one per element.
__device__ __forceinline__ void fill_rows(float *dst, int ld, int nb, const float *main_base, int Kmain,
                                          const float *tail_base, int64_t tail_ld, int ntail, int warp, int lane) {
    for (int f = warp; f < BT; f += FMA_WARPS) {
        float *row = dst + (size_t)f * ld;
        if (f >= nb) {
            for (int k = lane; k < ld; k += 32) row[k] = 0.f;
            continue;
        }
        const float *src = main_base + (size_t)f * Kmain;
        if ((Kmain & 3) == 0) {
            const float4 *s4 = reinterpret_cast<const float4 *>(src);
            float4 *d4 = reinterpret_cast<float4 *>(row);
            const int K4 = Kmain >> 2;
            for (int base = 0; base < K4; base += 128) {
                float4 v[4];
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const int i = base + j * 32 + lane;
                    if (i < K4) v[j] = ld_cg4(s4 + i);
                }
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const int i = base + j * 32 + lane;
                    if (i < K4) d4[i] = v[j];
                }
            }
        } else {
            for (int k = lane; k < Kmain; k += 32) row[k] = ld_cg(src + k);
        }
        const float *tl = tail_base + (size_t)f * tail_ld;
        for (int k = lane; k < ntail; k += 32) row[Kmain + k] = __ldg(tl + k);
        for (int k = Kmain + ntail + lane; k < ld; k += 32) row[k] = 0.f;
    }
}

__global__ void __launch_bounds__(FMA_THREADS, 1) fma_loop_kernel(const FmaParams P) {
    extern __shared__ __align__(16) float smem[];
    const Dims &d = P.d;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, cta = blockIdx.x;
    const int H = d.H, FC = d.FC, F = d.F, A = d.A, NC = d.NC;
    const int64_t B = P.a.B, S = P.a.S;
    float *wimg = smem;
    float *xbuf = wimg + P.img_floats;                 // [BT][Kmax]
    float *hbuf = xbuf + P.xbuf_floats;                // [BT][KH]
    float *obuf = hbuf + P.hbuf_floats;                // [2*MAXQ][12*BT]
    __shared__ MatJob jobs[2];
    __shared__ float xprev[BT];

    for (int64_t i = tid; i < P.img_floats / 4; i += FMA_THREADS)
        reinterpret_cast<float4 *>(wimg)[i] = reinterpret_cast<const float4 *>(P.image + (size_t)cta * P.img_floats)[i];
    __syncthreads();
    const float *bias = wimg + P.off_bias;
    const float *bI = bias, *b1i = bI + P.upc, *b1h = b1i + 3 * P.upc, *b2i = b1h + 3 * P.upc,
                *b2h = b2i + 3 * P.upc, *bf1 = b2h + 3 * P.upc, *bf2 = bf1 + P.rF, *bf3 = bf2 + P.rF;
    const int u0 = cta * P.upc, nu = max(0, min(P.upc, H - u0));          // GRU units / I rows owned
    const int f0 = cta * P.rF, nf1 = max(0, min(P.rF, FC - f0));          // fc1/fc2 rows owned
    const int c0 = cta * P.r3, nc3 = max(0, min(P.r3, NC - c0));          // fc3 rows owned
    const int nchunks = (int)((B + BT - 1) / BT);
    const uint2 key = make_uint2((uint32_t)P.a.seed, (uint32_t)(P.a.seed >> 32));
    const int nmix = NC / 3;
    unsigned long long gen = 0;
    auto pick_nq = [](int nmat, int K4) {
        int nq = FMA_WARPS / nmat;
        while (nq > 1 && (nq - 1) * 32 >= K4) nq--;
        return nq < 1 ? 1 : nq;
    };

    for (int64_t t = 0; t <= S; t++) {
        // ------------------------------------------------------------------ P0: sample(t-1) + I layer
        for (int ch = 0; ch < nchunks; ch++) {
            const int64_t b0 = (int64_t)ch * BT;
            const int nb = (int)min((int64_t)BT, B - b0);
            // previous sample of each fold in the chunk (one warp per fold)
            for (int f = warp; f < nb; f += FMA_WARPS) {
                const int64_t b = b0 + f;
                float xv = 0.f;
                if (t > 0) {
                    const int64_t tp = t - 1;
                    const float *lg = P.logits + (size_t)b * NC;
                    int pick = 0;
                    if (d.mode == WRNN_MODE_MOL) {
                        float myu1 = 0.5f, u2;
                        if (P.a.u1) {
                            if (lane < nmix) myu1 = P.a.u1[((size_t)tp * P.a.Btot + P.a.b0 + b) * nmix + lane];
                            u2 = P.a.u2[(size_t)tp * P.a.Btot + P.a.b0 + b];
                        } else {
                            const uint4 r = philox4x32(make_uint4((uint32_t)tp, (uint32_t)(P.a.b0 + b), (uint32_t)(lane >> 2), 1u), key);
                            const uint32_t rr[4] = {r.x, r.y, r.z, r.w};
                            myu1 = u_clip(rr[lane & 3]);
                            const uint4 r2 = philox4x32(make_uint4((uint32_t)tp, (uint32_t)(P.a.b0 + b), 0u, 2u), key);
                            u2 = u_clip(r2.x);
                        }
                        xv = mol_sample_warp(lg, nmix, myu1, u2, lane, &pick);
                    } else {
                        // logits were written by other CTAs: stage through registers with .cg loads
                        // into this warp's scratch row of obuf before the generic sampler reads them
                        float *lrow = obuf + (size_t)warp * NC;     // obuf is free during P0's sampling
                        for (int i = lane; i < NC; i += 32) lrow[i] = ld_cg(lg + i);
                        __syncwarp();
                        const float *qsrc = P.a.q ? P.a.q + ((size_t)tp * P.a.Btot + P.a.b0 + b) * NC : nullptr;
                        pick = categorical_warp(lrow, NC, qsrc, key, (uint32_t)tp, (uint32_t)(P.a.b0 + b), 3u, lane);
                        xv = raw_value(pick, NC);
                        __syncwarp();
                    }
                    if (tp >= P.a.fold_len[b]) xv = 0.f;
                    if (cta == 0 && lane == 0 && tp < P.a.fold_len[b]) {
                        P.a.samples[(size_t)b * S + tp] = xv;
                        P.a.picks[(size_t)b * S + tp] = pick;
                    }
                    if (P.a.teacher) xv = 0.f;   // replaced below
                }
                if (P.a.teacher && t < S) xv = P.a.teacher[(size_t)b * S + t];
                if (lane == 0) xprev[f] = xv;
            }
            if (t == S) continue;     // the extra iteration only samples the last step
            __syncthreads();
            // x = [x_prev, m_t, a1_t]  (fatchord_version.py:208)
            for (int f = warp; f < BT; f += FMA_WARPS) {
                float *row = xbuf + (size_t)f * P.KI;
                const float *crow = P.a.cond + ((size_t)(b0 + f) * S + t) * (F + d.R);
                float v[4];
#pragma unroll
                for (int j = 0; j < 4; j++) {              // KI <= 128 covered in one batch, else loop below
                    const int k = j * 32 + lane;
                    v[j] = (f < nb && k >= 1 && k <= F + A) ? __ldg(crow + k - 1) : 0.f;
                }
                if (lane == 0 && f < nb) v[0] = xprev[f];
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const int k = j * 32 + lane;
                    if (k < P.KI) row[k] = v[j];
                }
                for (int k = 128 + lane; k < P.KI; k += 32)
                    row[k] = (f < nb && k <= F + A) ? __ldg(crow + k - 1) : 0.f;
            }
            __syncthreads();
            if (nu > 0) {
                if (tid == 0)
                    jobs[0] = MatJob{reinterpret_cast<const float4 *>(wimg + P.off_I), reinterpret_cast<const float4 *>(xbuf),
                                     P.KI / 4, P.KI / 4, P.KI / 4, nu};
                __syncthreads();
                const int nq = pick_nq(1, P.KI / 4);
                run_jobs<4>(jobs, 1, nq, obuf, warp, lane);
                __syncthreads();
                for (int i = tid; i < nu * nb; i += FMA_THREADS) {
                    const int r = i / nb, f = i % nb;
                    P.xI[(size_t)(b0 + f) * H + u0 + r] = sum_q(obuf, 0, nq, 4 * BT, r * BT + f) + bI[r];
                }
            }
            __syncthreads();
        }
        if (t == S) break;
        grid_barrier(P.gs, gen);

        // ------------------------------------------------------------------ P1, P2: GRUs
        for (int layer = 0; layer < 2; layer++) {
            const float *xin = layer == 0 ? P.xI : P.x1;
            float *xout = layer == 0 ? P.x1 : P.x2;
            float *hbase = layer == 0 ? P.h1 : P.h2;
            const float *hcur = hbase + (size_t)(t & 1) * B * H;
            float *hnxt = hbase + (size_t)((t + 1) & 1) * B * H;
            const int Kin = layer == 0 ? P.KH : P.KHA;
            const float *Wi = wimg + (layer == 0 ? P.off_r1i : P.off_r2i), *Wh = wimg + (layer == 0 ? P.off_r1h : P.off_r2h);
            const float *bi = layer == 0 ? b1i : b2i, *bh = layer == 0 ? b1h : b2h;
            const int aoff = F + A;      // a2_t = aux channels [A, 2A)
            for (int ch = 0; ch < nchunks; ch++) {
                const int64_t b0 = (int64_t)ch * BT;
                const int nb = (int)min((int64_t)BT, B - b0);
                fill_rows(xbuf, Kin, nb, xin + (size_t)b0 * H, H,
                          P.a.cond + ((size_t)b0 * S + t) * (F + d.R) + aoff, (int64_t)S * (F + d.R),
                          layer == 1 ? A : 0, warp, lane);
                fill_rows(hbuf, P.KH, nb, hcur + (size_t)b0 * H, H, P.a.cond, 0, 0, warp, lane);
                __syncthreads();
                if (nu > 0) {
                    if (tid == 0) {
                        jobs[0] = MatJob{reinterpret_cast<const float4 *>(Wi), reinterpret_cast<const float4 *>(xbuf),
                                         Kin / 4, Kin / 4, Kin / 4, 3 * nu};
                        jobs[1] = MatJob{reinterpret_cast<const float4 *>(Wh), reinterpret_cast<const float4 *>(hbuf),
                                         P.KH / 4, P.KH / 4, P.KH / 4, 3 * nu};
                    }
                    __syncthreads();
                    const int nq = pick_nq(2, P.KH / 4);
                    run_jobs<12>(jobs, 2, nq, obuf, warp, lane);
                    __syncthreads();
                    for (int i = tid; i < nu * nb; i += FMA_THREADS) {
                        const int u = i / nb, f = i % nb;
                        // image rows are [gate][unit-in-CTA] with `upc` units per gate
                        const float ir = sum_q(obuf, 0, nq, 12 * BT, (0 * P.upc + u) * BT + f) + bi[0 * P.upc + u];
                        const float iz = sum_q(obuf, 0, nq, 12 * BT, (1 * P.upc + u) * BT + f) + bi[1 * P.upc + u];
                        const float in_ = sum_q(obuf, 0, nq, 12 * BT, (2 * P.upc + u) * BT + f) + bi[2 * P.upc + u];
                        const float hr = sum_q(obuf, 1, nq, 12 * BT, (0 * P.upc + u) * BT + f) + bh[0 * P.upc + u];
                        const float hz = sum_q(obuf, 1, nq, 12 * BT, (1 * P.upc + u) * BT + f) + bh[1 * P.upc + u];
                        const float hn = sum_q(obuf, 1, nq, 12 * BT, (2 * P.upc + u) * BT + f) + bh[2 * P.upc + u];
                        const float hp = hbuf[f * P.KH + u0 + u];
                        const float hnew = gru_update(ir, iz, in_, hr, hz, hn, hp);
                        hnxt[(size_t)(b0 + f) * H + u0 + u] = hnew;
                        xout[(size_t)(b0 + f) * H + u0 + u] = xbuf[f * Kin + u0 + u] + hnew;   // x = x + h (:212,216)
                    }
                }
                __syncthreads();
            }
            grid_barrier(P.gs, gen);
        }

        // ------------------------------------------------------------------ P3, P4: fc1, fc2 (+ReLU)
        for (int layer = 0; layer < 2; layer++) {
            const float *xin = layer == 0 ? P.x2 : P.y1;
            float *yout = layer == 0 ? P.y1 : P.y2;
            const int Kmain = layer == 0 ? H : FC, Kin = layer == 0 ? P.KHA : P.KFA;
            const float *W = wimg + (layer == 0 ? P.off_f1 : P.off_f2), *bb = layer == 0 ? bf1 : bf2;
            const int aoff = F + (2 + layer) * A;   // a3_t, a4_t
            for (int ch = 0; ch < nchunks; ch++) {
                const int64_t b0 = (int64_t)ch * BT;
                const int nb = (int)min((int64_t)BT, B - b0);
                fill_rows(xbuf, Kin, nb, xin + (size_t)b0 * Kmain, Kmain,
                          P.a.cond + ((size_t)b0 * S + t) * (F + d.R) + aoff, (int64_t)S * (F + d.R), A, warp, lane);
                __syncthreads();
                if (nf1 > 0) {
                    if (tid == 0)
                        jobs[0] = MatJob{reinterpret_cast<const float4 *>(W), reinterpret_cast<const float4 *>(xbuf),
                                         Kin / 4, Kin / 4, Kin / 4, nf1};
                    __syncthreads();
                    const int nq = pick_nq(1, Kin / 4);
                    run_jobs<4>(jobs, 1, nq, obuf, warp, lane);
                    __syncthreads();
                    for (int i = tid; i < nf1 * nb; i += FMA_THREADS) {
                        const int r = i / nb, f = i % nb;
                        yout[(size_t)(b0 + f) * FC + f0 + r] = fmaxf(sum_q(obuf, 0, nq, 4 * BT, r * BT + f) + bb[r], 0.f);
                    }
                }
                __syncthreads();
            }
            grid_barrier(P.gs, gen);
        }

        // ------------------------------------------------------------------ P5: fc3 -> logits
        for (int ch = 0; ch < nchunks; ch++) {
            const int64_t b0 = (int64_t)ch * BT;
            const int nb = (int)min((int64_t)BT, B - b0);
            if (nc3 > 0) {      // CTA-uniform
                fill_rows(xbuf, P.KF, nb, P.y2 + (size_t)b0 * FC, FC, P.a.cond, 0, 0, warp, lane);
                if (tid == 0)
                    jobs[0] = MatJob{reinterpret_cast<const float4 *>(wimg + P.off_f3), reinterpret_cast<const float4 *>(xbuf),
                                     P.KF / 4, P.KF / 4, P.KF / 4, nc3};
                __syncthreads();
                const int nq = pick_nq(1, P.KF / 4);
                run_jobs<4>(jobs, 1, nq, obuf, warp, lane);
                __syncthreads();
                for (int i = tid; i < nc3 * nb; i += FMA_THREADS) {
                    const int r = i / nb, f = i % nb;
                    const float v = sum_q(obuf, 0, nq, 4 * BT, r * BT + f) + bf3[r];
                    P.logits[(size_t)(b0 + f) * NC + c0 + r] = v;
                    if (P.a.logits_out && t < P.a.fold_len[b0 + f])
                        P.a.logits_out[((size_t)(b0 + f) * S + t) * NC + c0 + r] = v;
                }
                __syncthreads();
            }
        }
        grid_barrier(P.gs, gen);
    }
}

// ============================================================================ host: plan + launch
static inline int pad4(int k) { return (k + 3) & ~3; }

// Build the per-CTA shared-memory weight images from the checkpoint tensors
// (`get_gru_cell` aliasing at fatchord_version.py:273-279: rows of weight_ih/hh are [r; z; n]).
int fma_plan(wrnn_handle *h) {
    const Dims &d = h->d;
    FmaPlan &p = h->fma;
    const int G = h->num_sms;
    p.G = G;
    p.upc = (d.H + G - 1) / G;
    p.rF = (d.FC + G - 1) / G;
    p.r3 = (d.NC + G - 1) / G;
    if (p.upc > 4 || p.rF > 4 || p.r3 > 4)
        return fail(h, WRNN_ERR_ARG, "dims too large for the FMA loop kernel (rnn/fc/classes per SM > 4)");
    p.KI = pad4(1 + d.F + d.A);
    p.KH = pad4(d.H);
    p.KHA = pad4(d.H + d.A);
    p.KFA = pad4(d.FC + d.A);
    p.KF = pad4(d.FC);
    int64_t o = 0;
    p.off_I = o;   o += (int64_t)p.upc * p.KI;
    p.off_r1i = o; o += (int64_t)3 * p.upc * p.KH;
    p.off_r1h = o; o += (int64_t)3 * p.upc * p.KH;
    p.off_r2i = o; o += (int64_t)3 * p.upc * p.KHA;
    p.off_r2h = o; o += (int64_t)3 * p.upc * p.KH;
    p.off_f1 = o;  o += (int64_t)p.rF * p.KHA;
    p.off_f2 = o;  o += (int64_t)p.rF * p.KFA;
    p.off_f3 = o;  o += (int64_t)p.r3 * p.KF;
    p.off_bias = o; o += p.upc + 12 * p.upc + 2 * p.rF + p.r3;
    p.img_floats = (o + 3) & ~(int64_t)3;

    auto get = [&](const char *k, std::vector<int64_t> shape) -> const HostTensor * {
        auto it = h->w.find(k);
        if (it == h->w.end()) { fail(h, WRNN_ERR_STATE, "checkpoint key missing: %s", k); return nullptr; }
        if (it->second.shape != shape) { fail(h, WRNN_ERR_ARG, "checkpoint key %s has the wrong shape", k); return nullptr; }
        return &it->second;
    };
    const int H = d.H, FC = d.FC, NC = d.NC, KinI = 1 + d.F + d.A;
    const HostTensor *WI = get("I.weight", {H, KinI}), *BI = get("I.bias", {H});
    const HostTensor *W1i = get("rnn1.weight_ih_l0", {3 * H, H}), *W1h = get("rnn1.weight_hh_l0", {3 * H, H});
    const HostTensor *B1i = get("rnn1.bias_ih_l0", {3 * H}), *B1h = get("rnn1.bias_hh_l0", {3 * H});
    const HostTensor *W2i = get("rnn2.weight_ih_l0", {3 * H, H + d.A}), *W2h = get("rnn2.weight_hh_l0", {3 * H, H});
    const HostTensor *B2i = get("rnn2.bias_ih_l0", {3 * H}), *B2h = get("rnn2.bias_hh_l0", {3 * H});
    const HostTensor *F1 = get("fc1.weight", {FC, H + d.A}), *F1b = get("fc1.bias", {FC});
    const HostTensor *F2 = get("fc2.weight", {FC, FC + d.A}), *F2b = get("fc2.bias", {FC});
    const HostTensor *F3 = get("fc3.weight", {NC, FC}), *F3b = get("fc3.bias", {NC});
    if (!WI || !BI || !W1i || !W1h || !B1i || !B1h || !W2i || !W2h || !B2i || !B2h || !F1 || !F1b || !F2 || !F2b ||
        !F3 || !F3b)
        return WRNN_ERR_STATE;

    std::vector<float> img((size_t)G * p.img_floats, 0.f);
    auto copy_row = [](float *dst, const HostTensor *W, int row, int K) {
        memcpy(dst, W->v.data() + (size_t)row * K, (size_t)K * sizeof(float));
    };
    for (int c = 0; c < G; c++) {
        float *im = img.data() + (size_t)c * p.img_floats;
        float *bias = im + p.off_bias;
        for (int u = 0; u < p.upc; u++) {
            const int unit = c * p.upc + u;
            if (unit >= H) break;
            copy_row(im + p.off_I + (size_t)u * p.KI, WI, unit, KinI);
            bias[u] = BI->v[unit];
            for (int g = 0; g < 3; g++) {
                const int row = g * H + unit, lr = g * p.upc + u;
                copy_row(im + p.off_r1i + (size_t)lr * p.KH, W1i, row, H);
                copy_row(im + p.off_r1h + (size_t)lr * p.KH, W1h, row, H);
                copy_row(im + p.off_r2i + (size_t)lr * p.KHA, W2i, row, H + d.A);
                copy_row(im + p.off_r2h + (size_t)lr * p.KH, W2h, row, H);
                bias[p.upc + lr] = B1i->v[row];
                bias[p.upc + 3 * p.upc + lr] = B1h->v[row];
                bias[p.upc + 6 * p.upc + lr] = B2i->v[row];
                bias[p.upc + 9 * p.upc + lr] = B2h->v[row];
            }
        }
        for (int r = 0; r < p.rF; r++) {
            const int row = c * p.rF + r;
            if (row >= FC) break;
            copy_row(im + p.off_f1 + (size_t)r * p.KHA, F1, row, H + d.A);
            copy_row(im + p.off_f2 + (size_t)r * p.KFA, F2, row, FC + d.A);
            bias[13 * p.upc + r] = F1b->v[row];
            bias[13 * p.upc + p.rF + r] = F2b->v[row];
        }
        for (int r = 0; r < p.r3; r++) {
            const int row = c * p.r3 + r;
            if (row >= NC) break;
            copy_row(im + p.off_f3 + (size_t)r * p.KF, F3, row, FC);
            bias[13 * p.upc + 2 * p.rF + r] = F3b->v[row];
        }
    }
    if (p.image) cudaFree(p.image);
    p.image = nullptr;
    WRNN_CUDA(h, cudaMalloc((void **)&p.image, img.size() * sizeof(float)));
    WRNN_CUDA(h, cudaMemcpy(p.image, img.data(), img.size() * sizeof(float), cudaMemcpyHostToDevice));

    int kmax = p.KI;
    kmax = kmax > p.KHA ? kmax : p.KHA;
    kmax = kmax > p.KFA ? kmax : p.KFA;
    kmax = kmax > p.KF ? kmax : p.KF;
    const int64_t obuf = (int64_t)FMA_WARPS * (12 * BT > NC ? 12 * BT : NC);
    p.smem_bytes = (size_t)(p.img_floats + (int64_t)BT * kmax + (int64_t)BT * p.KH + obuf) * sizeof(float);
    if (p.smem_bytes > h->smem_optin)
        return fail(h, WRNN_ERR_ARG, "FMA loop kernel needs %zu B of shared memory, device offers %zu", p.smem_bytes,
                    h->smem_optin);
    WRNN_CUDA(h, cudaFuncSetAttribute(fma_loop_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem_bytes));
    return 0;
}

int fma_run(wrnn_handle *h, const LoopArgs &a, cudaStream_t st) {
    const Dims &d = h->d;
    const FmaPlan &p = h->fma;
    FmaParams P;
    P.d = d;
    P.upc = p.upc; P.rF = p.rF; P.r3 = p.r3;
    P.KI = p.KI; P.KH = p.KH; P.KHA = p.KHA; P.KFA = p.KFA; P.KF = p.KF;
    P.img_floats = p.img_floats;
    P.off_I = p.off_I; P.off_r1i = p.off_r1i; P.off_r1h = p.off_r1h; P.off_r2i = p.off_r2i; P.off_r2h = p.off_r2h;
    P.off_f1 = p.off_f1; P.off_f2 = p.off_f2; P.off_f3 = p.off_f3; P.off_bias = p.off_bias;
    P.image = p.image;
    P.a = a;
    int kmax = p.KI;
    kmax = kmax > p.KHA ? kmax : p.KHA;
    kmax = kmax > p.KFA ? kmax : p.KFA;
    kmax = kmax > p.KF ? kmax : p.KF;
    P.xbuf_floats = BT * kmax;
    P.hbuf_floats = BT * p.KH;
    // exchange buffers: xI, x1, x2 [B][H]; y1, y2 [B][FC]; h1, h2 [2][B][H]; logits [B][NC]
    const size_t BH = (size_t)a.B * d.H, BF = (size_t)a.B * d.FC, BN = (size_t)a.B * d.NC;
    const size_t total = 3 * BH + 2 * BF + 4 * BH + BN;
    WRNN_TRY(ws_reserve(h, h->ws_act, total * sizeof(float)));
    WRNN_CUDA(h, cudaMemsetAsync(h->ws_act.p, 0, total * sizeof(float), st));   // h1 = h2 = 0 (:194-195)
    float *base = (float *)h->ws_act.p;
    P.xI = base; P.x1 = P.xI + BH; P.x2 = P.x1 + BH;
    P.y1 = P.x2 + BH; P.y2 = P.y1 + BF;
    P.h1 = P.y2 + BF; P.h2 = P.h1 + 2 * BH;
    P.logits = P.h2 + 2 * BH;
    WRNN_TRY(ws_reserve(h, h->ws_sync, 128));
    WRNN_CUDA(h, cudaMemsetAsync(h->ws_sync.p, 0, 16, st));          // the abort flag is cleared once per run
    P.gs.counter = (unsigned long long *)h->ws_sync.p;
    P.gs.abort_flag = (int *)((char *)h->ws_sync.p + 64);
    // the attribute is per function, not per handle: another handle (other dims) may have changed it
    WRNN_CUDA(h, cudaFuncSetAttribute(fma_loop_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem_bytes));
    void *args[] = {(void *)&P};
    WRNN_CUDA(h, cudaLaunchCooperativeKernel((const void *)fma_loop_kernel, dim3(p.G), dim3(FMA_THREADS), args,
                                             p.smem_bytes, st));
    h->launches++;
    return 0;
}

}  // namespace wrnn

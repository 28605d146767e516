// Conditioning (MelResNet + UpsampleNetwork), fold gather, and the float64 output tail.
// Reference behaviour: models/fatchord_version.py:13-89 (networks), :281-340 (pad, fold),
// :243-258 + :342-405 (mu-law decode, xfade_and_unfold, fade-out), utils/dsp.py:98-103.
// All of this is HBM-bound streaming work: coalesced 128-bit accesses, no tensor cores.
#include <math.h>
#include <stdio.h>
#include <string.h>

#include "internal.h"

namespace wrnn {

// ============================================================================ packing
static int upload(wrnn_handle *h, float **dst, const std::vector<float> &v) {
    if (*dst) cudaFree(*dst);
    *dst = nullptr;
    WRNN_CUDA(h, cudaMalloc((void **)dst, v.size() * sizeof(float)));
    WRNN_CUDA(h, cudaMemcpy(*dst, v.data(), v.size() * sizeof(float), cudaMemcpyHostToDevice));
    return 0;
}

static const HostTensor *need(wrnn_handle *h, const std::string &k, std::vector<int64_t> shape) {
    auto it = h->w.find(k);
    if (it == h->w.end()) {
        fail(h, WRNN_ERR_STATE, "checkpoint key missing: %s", k.c_str());
        return nullptr;
    }
    if (it->second.shape != shape) {
        fail(h, WRNN_ERR_ARG, "checkpoint key %s has the wrong shape", k.c_str());
        return nullptr;
    }
    return &it->second;
}

int pack_cond(wrnn_handle *h) {
    const Dims &d = h->d;
    const int C = d.C, F = d.F, R = d.R, ks = d.ksize;
    const std::string p = "upsample.resnet.";
    // conv_in (C, F, ks) -> [(f*ks + j)][c]
    const HostTensor *t = need(h, p + "conv_in.weight", {C, F, ks});
    if (!t) return WRNN_ERR_STATE;
    std::vector<float> v((size_t)F * ks * C);
    for (int c = 0; c < C; c++)
        for (int f = 0; f < F; f++)
            for (int j = 0; j < ks; j++) v[((size_t)f * ks + j) * C + c] = t->v[((size_t)c * F + f) * ks + j];
    WRNN_TRY(upload(h, &h->cp.conv_in_t, v));
    // BatchNorm (eval): y = x * scale + shift, eps = 1e-5 (fatchord_version.py:18-19,36)
    const int nbn = 1 + 2 * d.nres;
    std::vector<float> sc((size_t)nbn * C), sh((size_t)nbn * C);
    auto bn = [&](int idx, const std::string &pre) -> int {
        const HostTensor *w = need(h, pre + ".weight", {C}), *b = need(h, pre + ".bias", {C});
        const HostTensor *m = need(h, pre + ".running_mean", {C}), *var = need(h, pre + ".running_var", {C});
        if (!w || !b || !m || !var) return WRNN_ERR_STATE;
        for (int c = 0; c < C; c++) {
            float s = w->v[c] / sqrtf(var->v[c] + 1e-5f);
            sc[(size_t)idx * C + c] = s;
            sh[(size_t)idx * C + c] = b->v[c] - m->v[c] * s;
        }
        return 0;
    };
    WRNN_TRY(bn(0, p + "batch_norm"));
    std::vector<float> rt((size_t)2 * d.nres * C * C);
    for (int i = 0; i < d.nres; i++) {
        const std::string q = p + "layers." + std::to_string(i);
        for (int j = 0; j < 2; j++) {
            const HostTensor *cw = need(h, q + ".conv" + std::to_string(j + 1) + ".weight", {C, C, 1});
            if (!cw) return WRNN_ERR_STATE;
            float *dst = rt.data() + (size_t)(2 * i + j) * C * C;
            for (int o = 0; o < C; o++)
                for (int k = 0; k < C; k++) dst[(size_t)k * C + o] = cw->v[(size_t)o * C + k];
            WRNN_TRY(bn(1 + 2 * i + j, q + ".batch_norm" + std::to_string(j + 1)));
        }
    }
    WRNN_TRY(upload(h, &h->cp.bn_scale, sc));
    WRNN_TRY(upload(h, &h->cp.bn_shift, sh));
    if (d.nres > 0) WRNN_TRY(upload(h, &h->cp.res_t, rt));
    const HostTensor *co = need(h, p + "conv_out.weight", {R, C, 1}), *cb = need(h, p + "conv_out.bias", {R});
    if (!co || !cb) return WRNN_ERR_STATE;
    std::vector<float> ot((size_t)C * R);
    for (int r = 0; r < R; r++)
        for (int k = 0; k < C; k++) ot[(size_t)k * R + r] = co->v[(size_t)r * C + k];
    WRNN_TRY(upload(h, &h->cp.conv_out_t, ot));
    WRNN_TRY(upload(h, &h->cp.conv_out_b, cb->v));
    for (int j = 0; j < d.nup; j++) {
        const HostTensor *fw = need(h, "upsample.up_layers." + std::to_string(2 * j + 1) + ".weight",
                                    {1, 1, 1, 2 * d.up[j] + 1});
        if (!fw) return WRNN_ERR_STATE;
        WRNN_TRY(upload(h, &h->cp.fir[j], fw->v));
    }
    return 0;
}

// ============================================================================ MelResNet
// One CTA owns RES_TT consecutive frames and carries them through every layer: after conv_in all
// layers are 1x1 in time, so tiles are independent.  Thread c owns output channel c; weights are
// pre-transposed [k][c] so a warp reads 128 contiguous bytes per k; activations are smem broadcasts.
constexpr int RES_TT = 8;

__device__ __forceinline__ void res_matvec(const float *__restrict__ Wt, int K, int ldw, int c,
                                           const float *act, float (&acc)[RES_TT]) {
#pragma unroll
    for (int t = 0; t < RES_TT; t++) acc[t] = 0.f;
    for (int k = 0; k < K; k++) {
        const float w = __ldg(Wt + (size_t)k * ldw + c);
        const float *a = act + k * RES_TT;
#pragma unroll
        for (int t = 0; t < RES_TT; t++) acc[t] = fmaf(w, a[t], acc[t]);
    }
}

__global__ void __launch_bounds__(128) resnet_kernel(const float *__restrict__ mel_padded, int Tp, int T,
                                                      Dims d, CondPack p, float *__restrict__ res_out) {
    extern __shared__ float sm[];
    const int C = d.C, F = d.F, ks = d.ksize, R = d.R;
    float *in = sm;                              // [F*ks][RES_TT]   im2col of the mel tile
    float *a0 = in + (size_t)F * ks * RES_TT;    // [C][RES_TT]
    float *a1 = a0 + (size_t)C * RES_TT;         // [C][RES_TT]
    const int t0 = blockIdx.x * RES_TT;
    for (int i = threadIdx.x; i < F * ks * RES_TT; i += blockDim.x) {
        const int t = i % RES_TT, fj = i / RES_TT, f = fj / ks, j = fj % ks;
        const int tt = t0 + t + j;               // valid conv over the padded mel
        in[i] = (t0 + t < T && tt < Tp) ? mel_padded[(size_t)f * Tp + tt] : 0.f;
    }
    __syncthreads();
    float acc[RES_TT];
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        res_matvec(p.conv_in_t, F * ks, C, c, in, acc);
        const float s = p.bn_scale[c], b = p.bn_shift[c];
#pragma unroll
        for (int t = 0; t < RES_TT; t++) a0[c * RES_TT + t] = fmaxf(fmaf(acc[t], s, b), 0.f);
    }
    __syncthreads();
    for (int i = 0; i < d.nres; i++) {
        const float *W1 = p.res_t + (size_t)(2 * i) * C * C, *W2 = W1 + (size_t)C * C;
        const float *s1 = p.bn_scale + (size_t)(1 + 2 * i) * C, *b1 = p.bn_shift + (size_t)(1 + 2 * i) * C;
        const float *s2 = s1 + C, *b2 = b1 + C;
        for (int c = threadIdx.x; c < C; c += blockDim.x) {
            res_matvec(W1, C, C, c, a0, acc);
#pragma unroll
            for (int t = 0; t < RES_TT; t++) a1[c * RES_TT + t] = fmaxf(fmaf(acc[t], s1[c], b1[c]), 0.f);
        }
        __syncthreads();
        float nxt[RES_TT];   // C <= blockDim handled in registers; larger C loops (values written after sync)
        for (int c0 = 0; c0 < C; c0 += blockDim.x) {
            const int c = c0 + threadIdx.x;
            if (c < C) {
                res_matvec(W2, C, C, c, a1, acc);
#pragma unroll
                for (int t = 0; t < RES_TT; t++) nxt[t] = fmaf(acc[t], s2[c], b2[c]) + a0[c * RES_TT + t];
#pragma unroll
                for (int t = 0; t < RES_TT; t++) a0[c * RES_TT + t] = nxt[t];   // a0[c] is only read by its owner here
            }
        }
        __syncthreads();
    }
    for (int r = threadIdx.x; r < R; r += blockDim.x) {
        res_matvec(p.conv_out_t, C, R, r, a0, acc);
        const float b = p.conv_out_b[r];
#pragma unroll
        for (int t = 0; t < RES_TT; t++)
            if (t0 + t < T) res_out[(size_t)(t0 + t) * R + r] = acc[t] + b;
    }
}

// ============================================================================ mel upsampling
// One stage of UpsampleNetwork's mel branch (fatchord_version.py:73-80,86-87): nearest-repeat by s
// along time, then a (2s+1)-tap FIR with ZERO padding of s at both ends of the repeated signal.
// in [F][n_in] -> out [F][n_in*s], channel-major.
__global__ void fir_stage_kernel(const float *__restrict__ in, int64_t n_in, int s,
                                 const float *__restrict__ w, float *__restrict__ out) {
    const int f = blockIdx.y;
    const int64_t n_out = n_in * s;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_out) return;
    const float *row = in + (size_t)f * n_in;
    float acc = 0.f;
    for (int j = 0; j <= 2 * s; j++) {
        const int64_t pidx = i + j - s;
        if (pidx >= 0 && pidx < n_out) acc = fmaf(w[j], row[pidx / s], acc);
    }
    out[(size_t)f * n_out + i] = acc;
}

// Last stage: same FIR, but crop `indent` at both ends and write time-major [L][F]
// (the `.transpose(1, 2)` at fatchord_version.py:88-89) through a smem tile so the global stores
// are full contiguous rows.
constexpr int FIR_TI = 128;
__global__ void __launch_bounds__(256) fir_last_kernel(const float *__restrict__ in, int64_t n_in, int s,
                                                        const float *__restrict__ w, int F, int64_t indent,
                                                        int64_t L, float *__restrict__ out) {
    extern __shared__ float tile[];              // [FIR_TI][F + 1]
    const int64_t n_out = n_in * s;
    const int64_t l0 = (int64_t)blockIdx.x * FIR_TI;
    const int ldt = F + 1;
    for (int e = threadIdx.x; e < FIR_TI * F; e += blockDim.x) {
        const int f = e / FIR_TI, tl = e % FIR_TI;   // consecutive threads walk time: reads hit one row
        const int64_t l = l0 + tl;
        float acc = 0.f;
        if (l < L) {
            const int64_t i = l + indent;
            const float *row = in + (size_t)f * n_in;
            for (int j = 0; j <= 2 * s; j++) {
                const int64_t pidx = i + j - s;
                if (pidx >= 0 && pidx < n_out) acc = fmaf(w[j], row[pidx / s], acc);
            }
        }
        tile[tl * ldt + f] = acc;
    }
    __syncthreads();
    for (int e = threadIdx.x; e < FIR_TI * F; e += blockDim.x) {
        const int tl = e / F, f = e % F;
        if (l0 + tl < L) out[(size_t)(l0 + tl) * F + f] = tile[tl * ldt + f];
    }
}

__global__ void pad_mel_kernel(const float *__restrict__ mel, int T, int pad, int F, float *__restrict__ out) {
    const int Tp = T + 2 * pad;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (int64_t)F * Tp) return;
    const int f = i / Tp, t = i % Tp - pad;
    out[i] = (t >= 0 && t < T) ? mel[(size_t)f * T + t] : 0.f;
}

__global__ void repeat_aux_kernel(const float *__restrict__ res_out, int64_t L, int hop, int R,
                                  float *__restrict__ aux) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= L * R) return;
    const int64_t l = i / R;
    aux[i] = res_out[(size_t)(l / hop) * R + (i % R)];
}

int launch_pad_mel(wrnn_handle *h, const float *mel, int T, float *mel_padded, cudaStream_t st) {
    const int64_t n = (int64_t)h->d.F * (T + 2 * h->d.pad);
    pad_mel_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(mel, T, h->d.pad, h->d.F, mel_padded);
    h->launches++;
    WRNN_CUDA(h, cudaGetLastError());
    return 0;
}

int launch_repeat_aux(wrnn_handle *h, const float *res_out, int T, float *aux, cudaStream_t st) {
    const int64_t L = (int64_t)T * h->d.total_scale, n = L * h->d.R;
    repeat_aux_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(res_out, L, h->d.total_scale, h->d.R, aux);
    h->launches++;
    WRNN_CUDA(h, cudaGetLastError());
    return 0;
}

int run_conditioning(wrnn_handle *h, const float *mel_padded, int Tp, float *up_mels, float *res_out,
                     cudaStream_t st) {
    const Dims &d = h->d;
    const int T = Tp - 2 * d.pad;
    if (T < 1) return fail(h, WRNN_ERR_ARG, "mel too short: T=%d", T);
    // MelResNet
    const size_t smem = ((size_t)d.F * d.ksize + 2 * (size_t)d.C) * RES_TT * sizeof(float);
    if (smem > 48 * 1024)
        WRNN_CUDA(h, cudaFuncSetAttribute(resnet_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    resnet_kernel<<<(T + RES_TT - 1) / RES_TT, 128, smem, st>>>(mel_padded, Tp, T, d, h->cp, res_out);
    h->launches++;
    WRNN_CUDA(h, cudaGetLastError());
    // mel branch: nup stages; intermediates live in two scratch buffers
    int64_t n_in = Tp;
    const float *cur = mel_padded;
    size_t max_mid = 0;
    {
        int64_t n = Tp;
        for (int j = 0; j + 1 < d.nup; j++) { n *= d.up[j]; max_mid = (size_t)n > max_mid ? (size_t)n : max_mid; }
    }
    if (d.nup > 1) {
        WRNN_TRY(ws_reserve(h, h->ws_tmp0, max_mid * d.F * sizeof(float)));
        WRNN_TRY(ws_reserve(h, h->ws_tmp1, max_mid * d.F * sizeof(float)));
    }
    for (int j = 0; j < d.nup; j++) {
        const int s = d.up[j];
        const int64_t n_out = n_in * s;
        if (j + 1 < d.nup) {
            float *dst = (float *)((j & 1) ? h->ws_tmp1.p : h->ws_tmp0.p);
            dim3 grid((unsigned)((n_out + 255) / 256), d.F);
            fir_stage_kernel<<<grid, 256, 0, st>>>(cur, n_in, s, h->cp.fir[j], dst);
            cur = dst;
        } else {
            const int64_t L = (int64_t)T * d.total_scale;
            const size_t sm = (size_t)FIR_TI * (d.F + 1) * sizeof(float);
            if (sm > 48 * 1024)
                WRNN_CUDA(h, cudaFuncSetAttribute(fir_last_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
            fir_last_kernel<<<(unsigned)((L + FIR_TI - 1) / FIR_TI), 256, sm, st>>>(cur, n_in, s, h->cp.fir[j], d.F,
                                                                                 d.indent, L, up_mels);
        }
        h->launches++;
        WRNN_CUDA(h, cudaGetLastError());
        n_in = n_out;
    }
    return 0;
}

// ============================================================================ fold (gather)
// Generic fold_with_overlap (fatchord_version.py:293-340): out[b][s][:] = x[b*(t+o)+s][:] or 0.
__global__ void fold_kernel(const float *__restrict__ x, int64_t L, int F, int64_t stride, int64_t len,
                            int64_t folds, float *__restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t n = folds * len * F;
    if (i >= n) return;
    const int64_t f = i % F, s = (i / F) % len, b = i / (F * len);
    const int64_t src = b * stride + s;
    out[i] = src < L ? x[(size_t)src * F + f] : 0.f;
}

int launch_fold(wrnn_handle *h, const float *x, int64_t L, int F, int64_t target, int64_t overlap,
                int64_t folds, float *out, cudaStream_t st) {
    const int64_t len = target + 2 * overlap, n = folds * len * F;
    fold_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(x, L, F, target + overlap, len, folds, out);
    h->launches++;
    WRNN_CUDA(h, cudaGetLastError());
    return 0;
}

// The fold the loop consumes: mel features and the (frame-rate) aux features of fold b, step s,
// gathered into one [B][S][F+R] row so the loop reads 832 contiguous bytes per fold-step.  The
// x275 nearest-repeat of aux (Stretch2d, :83-85) is done here by index, never materialised.
template <int VEC>
__global__ void fold_cond_kernel(const float *__restrict__ up_mels, const float *__restrict__ res_out,
                                 const FoldMeta *__restrict__ meta, int64_t B, int64_t S, int F, int R, int hop,
                                 float *__restrict__ cond) {
    const int W = (F + R) / VEC;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B * S * W) return;
    const int e = (int)(i % W) * VEC;
    const int64_t s = (i / W) % S, b = i / ((int64_t)W * S);
    const FoldMeta m = meta[b];
    const int64_t l = m.start + s;
    float v[VEC];
#pragma unroll
    for (int k = 0; k < VEC; k++) v[k] = 0.f;
    if (l < m.L) {
        const float *src = e < F ? up_mels + (size_t)(m.up_off + l) * F + e
                                 : res_out + (size_t)(m.res_off + l / hop) * R + (e - F);
        if (VEC == 4) {
            const float4 q = *reinterpret_cast<const float4 *>(src);
            v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
        } else {
            v[0] = src[0];
        }
    }
    float *dst = cond + (size_t)i * VEC;
    if (VEC == 4) *reinterpret_cast<float4 *>(dst) = make_float4(v[0], v[1], v[2], v[3]);
    else dst[0] = v[0];
}

int launch_fold_cond(wrnn_handle *h, const float *up_mels, const float *res_out, const FoldMeta *meta,
                     int64_t B, int64_t S, float *cond, cudaStream_t st) {
    const Dims &d = h->d;
    if (d.F % 4 == 0 && d.R % 4 == 0) {
        const int64_t n = B * S * ((d.F + d.R) / 4);
        fold_cond_kernel<4><<<(unsigned)((n + 255) / 256), 256, 0, st>>>(up_mels, res_out, meta, B, S, d.F, d.R,
                                                                      d.total_scale, cond);
    } else {
        const int64_t n = B * S * (d.F + d.R);
        fold_cond_kernel<1><<<(unsigned)((n + 255) / 256), 256, 0, st>>>(up_mels, res_out, meta, B, S, d.F, d.R,
                                                                      d.total_scale, cond);
    }
    h->launches++;
    WRNN_CUDA(h, cudaGetLastError());
    return 0;
}

// ============================================================================ float64 output tail
// Everything after the loop in `generate` (fatchord_version.py:243-258): cast to float64, optional
// mu-law decode (a C-entry table, utils/dsp.py:98-103), xfade_and_unfold (:342-405), truncate to
// wave_len, linear fade-out over the last 20*hop samples.  Each output index receives at most two
// addends, so the result is order-independent; __dmul_rn/__dadd_rn forbid FMA contraction so every
// product and sum rounds exactly like numpy's.
__device__ __forceinline__ double xfade_gain_mul(double v, int64_t pos, int64_t len, int64_t ov,
                                                 const double *fade_in, const double *fade_out) {
    if (pos < ov) v = __dmul_rn(v, fade_in[pos]);              // y[:, :overlap] *= fade_in
    if (pos >= len - ov) v = __dmul_rn(v, fade_out[pos - (len - ov)]);   // y[:, -overlap:] *= fade_out
    return v;
}

__global__ void tail_kernel(const float *__restrict__ samples, const int32_t *__restrict__ picks,
                            const TailMeta *__restrict__ meta, int64_t S, int batched, int64_t target,
                            int64_t overlap, const double *__restrict__ fade_in,
                            const double *__restrict__ fade_out, const double *__restrict__ fo, int64_t fo_len,
                            const double *__restrict__ lut, double *__restrict__ out) {
    const TailMeta m = meta[blockIdx.y];
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m.wave_len) return;
    auto val = [&](int64_t b, int64_t pos) -> double {
        const size_t idx = (size_t)(m.fold0 + b) * S + pos;
        return lut ? lut[picks[idx]] : (double)samples[idx];
    };
    double v;
    if (batched) {
        const int64_t stride = target + overlap, len = target + 2 * overlap;
        const int64_t b = i / stride, pos = i - b * stride;
        v = 0.0;
        if (b >= 1 && pos < overlap)        // tail of the previous fold, added first like the numpy loop
            v = __dadd_rn(v, xfade_gain_mul(val(b - 1, pos + stride), pos + stride, len, overlap, fade_in, fade_out));
        if (b < m.nfolds) v = __dadd_rn(v, xfade_gain_mul(val(b, pos), pos, len, overlap, fade_in, fade_out));
    } else {
        v = val(0, i);
    }
    const int64_t k = i - (m.wave_len - fo_len);
    if (k >= 0) v = __dmul_rn(v, fo[k]);
    out[m.out_off + i] = v;
}

int launch_tail(wrnn_handle *h, const float *samples, const int32_t *picks, const TailMeta *meta, int n_utt,
                int64_t max_wave, int64_t S, int batched, int64_t target, int64_t overlap, const double *fade_in,
                const double *fade_out, const double *fo, int64_t fo_len, const double *lut, double *out,
                cudaStream_t st) {
    dim3 grid((unsigned)((max_wave + 255) / 256), n_utt);
    tail_kernel<<<grid, 256, 0, st>>>(samples, picks, meta, S, batched, target, overlap, fade_in, fade_out, fo,
                                      fo_len, lut, out);
    h->launches++;
    WRNN_CUDA(h, cudaGetLastError());
    return 0;
}

// Stand-alone xfade_and_unfold on caller-provided float64 folds (parity entry point).
__global__ void unfold_kernel(const double *__restrict__ y, int64_t folds, int64_t target, int64_t overlap,
                              const double *__restrict__ fade_in, const double *__restrict__ fade_out,
                              double *__restrict__ out) {
    const int64_t stride = target + overlap, len = target + 2 * overlap, total = folds * stride + overlap;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const int64_t b = i / stride, pos = i - b * stride;
    double v = 0.0;
    if (b >= 1 && pos < overlap)
        v = __dadd_rn(v, xfade_gain_mul(y[(size_t)(b - 1) * len + pos + stride], pos + stride, len, overlap, fade_in, fade_out));
    if (b < folds) v = __dadd_rn(v, xfade_gain_mul(y[(size_t)b * len + pos], pos, len, overlap, fade_in, fade_out));
    out[i] = v;
}

int launch_unfold(wrnn_handle *h, const double *y, int64_t folds, int64_t target, int64_t overlap,
                  const double *fade_in, const double *fade_out, double *out, cudaStream_t st) {
    const int64_t total = folds * (target + overlap) + overlap;
    unfold_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(y, folds, target, overlap, fade_in, fade_out, out);
    h->launches++;
    WRNN_CUDA(h, cudaGetLastError());
    return 0;
}

// decode_mu_law(y, mu, from_labels=False) for y = 2k/(mu-1)-1: recover k exactly, look up the table.
__global__ void mulaw_kernel(const float *__restrict__ y, int64_t n, int mu, const double *__restrict__ lut,
                             double *__restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int k = __float2int_rn((y[i] + 1.f) * 0.5f * (float)(mu - 1));
    k = min(max(k, 0), mu - 1);
    out[i] = lut[k];
}

int launch_mulaw(wrnn_handle *h, const float *y, int64_t n, int mu, const double *lut, double *out,
                 cudaStream_t st) {
    mulaw_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(y, n, mu, lut, out);
    h->launches++;
    WRNN_CUDA(h, cudaGetLastError());
    return 0;
}

}  // namespace wrnn

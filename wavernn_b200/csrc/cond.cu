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
    // ---- the mel branch as one polyphase filter -------------------------------------------------
    // UpsampleNetwork.forward (fatchord_version.py:86-88) is, per stage, nearest-repeat by s and a (2s+1)-tap FIR with
    // zero padding.  Repeat + FIR is linear and shift-invariant in whole frames, so the composition of all stages maps
    // padded mel frames to upsampled samples as  up[i] = sum_j g(i - j * hop) * mel[j].  g is the response to a unit
    // impulse, formed here once in double by running the stages on an impulse; it is non-zero for a few frames around
    // i / hop only, which gives a [hop][taps] table.  Zero padding per stage equals extending the padded mel with zeros:
    // a stage reaches one input element to each side, and outside the signal both read zero.
    {
        const int hop = d.total_scale, Tn = 2 * d.nup + 5, c = Tn / 2;
        std::vector<double> x(Tn, 0.0);
        x[c] = 1.0;
        for (int j = 0; j < d.nup; j++) {
            const int sc_ = d.up[j];
            const HostTensor *fw = need(h, "upsample.up_layers." + std::to_string(2 * j + 1) + ".weight", {1, 1, 1, 2 * sc_ + 1});
            if (!fw) return WRNN_ERR_STATE;
            const int64_t n_out = (int64_t)x.size() * sc_;
            std::vector<double> y((size_t)n_out, 0.0);
            for (int64_t i = 0; i < n_out; i++) {
                double acc = 0.0;
                for (int t2 = 0; t2 <= 2 * sc_; t2++) {
                    const int64_t pidx = i + t2 - sc_;
                    if (pidx >= 0 && pidx < n_out) acc += (double)fw->v[t2] * x[(size_t)(pidx / sc_)];
                }
                y[(size_t)i] = acc;
            }
            x.swap(y);
        }
        // x[i] = g(i - c * hop); tap k of phase p is g(p - k * hop) = x[(c - k) * hop + p]
        int kmin = 0, kmax = 0;
        for (int k = -(Tn - 1 - c); k <= c; k++)
            for (int ph = 0; ph < hop; ph++)
                if (x[(size_t)((c - k) * hop + ph)] != 0.0) { kmin = k < kmin ? k : kmin; kmax = k > kmax ? k : kmax; }
        if (kmin <= -(Tn - 1 - c) || kmax >= c) return fail(h, WRNN_ERR_ARG, "upsample network reaches further than expected");
        const int taps = kmax - kmin + 1;
        std::vector<float> tab((size_t)hop * taps);
        for (int ph = 0; ph < hop; ph++)
            for (int k = kmin; k <= kmax; k++) tab[(size_t)ph * taps + (k - kmin)] = (float)x[(size_t)((c - k) * hop + ph)];
        WRNN_TRY(upload(h, &h->cp.poly, tab));
        h->cp.poly_taps = taps;
        h->cp.poly_kmin = kmin;
    }
    return 0;
}

// ============================================================================ pad + transpose
// pad_tensor (fatchord_version.py:281-291) for every utterance of the call, written frame-major: melT[frame][f], so the
// kernels below read whole frames with 16-byte loads.  Raw mels are (F, Tin) row-major; `pad_` zero frames each side.
__global__ void pad_mel_batch_kernel(const float *__restrict__ mel_raw, const UttMeta *__restrict__ utts, int n_utt, int F,
                                     int64_t total, float *__restrict__ melT) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total * F) return;
    const int64_t fr = i / F;
    const int f = (int)(i % F);
    int lo = 0, hi = n_utt - 1;                    // utterance of this frame: pm_off is ascending
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (utts[mid].pm_off <= fr) lo = mid;
        else hi = mid - 1;
    }
    const UttMeta m = utts[lo];
    const int t = (int)(fr - m.pm_off) - m.pad_;
    melT[i] = (t >= 0 && t < m.Tin) ? mel_raw[m.raw_off + (size_t)f * m.Tin + t] : 0.f;
}

int launch_pad_mel_batch(wrnn_handle *h, const float *mel_raw, const UttMeta *utts, int n_utt, int64_t total_frames,
                         float *melT, cudaStream_t st) {
    const int64_t n = total_frames * h->d.F;
    pad_mel_batch_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(mel_raw, utts, n_utt, h->d.F, total_frames, melT);
    h->launches++;
    WRNN_CUDA(h, cudaGetLastError());
    return 0;
}

// ============================================================================ MelResNet
// MelResNet (fatchord_version.py:31-48) for every utterance of the call in one launch.  One CTA carries RES_TT
// consecutive frames of one utterance through every layer (after conv_in all layers are 1x1 in time, so tiles are
// independent).  Each layer is a [Cout x K] . [K x RES_TT] product: the weights stream from L2 through a double-buffered
// 2 x 64 KB shared-memory stage (cp.async, next chunk in flight while this one is used: the first version read them with
// dependent __ldg loads and was latency-bound at 0.8 TFLOP/s), activations are shared-memory broadcasts, thread
// (c, half) owns channel c for four frames.  BatchNorm is folded to scale/shift (eval mode, eps 1e-5).
constexpr int RES_TT = 8, RES_THREADS = 256, RES_WBUF = 16384, RES_MAXCI = 4;
int resnet_tile_frames() { return RES_TT; }

struct ResPass { const float *W; int K, Cout; };
__device__ __forceinline__ ResPass res_pass(const Dims &d, const CondPack &p, int i) {
    if (i == 0) return ResPass{p.conv_in_t, d.F * d.ksize, d.C};
    if (i <= 2 * d.nres) return ResPass{p.res_t + (size_t)(i - 1) * d.C * d.C, d.C, d.C};
    return ResPass{p.conv_out_t, d.C, d.R};
}
__device__ __forceinline__ void res_stage(float *buf, const ResPass &ps, int ch, int tid) {
    const int KR = RES_WBUF / ps.Cout, k0 = ch * KR, rows = min(KR, ps.K - k0);
    const float4 *src = reinterpret_cast<const float4 *>(ps.W + (size_t)k0 * ps.Cout);
    const uint32_t dst = (uint32_t)__cvta_generic_to_shared(buf);
    for (int i = tid; i < rows * ps.Cout / 4; i += RES_THREADS)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + 16u * i), "l"(src + i) : "memory");
    asm volatile("cp.async.commit_group;" ::: "memory");
}

__global__ void __launch_bounds__(RES_THREADS) resnet_batch_kernel(const float *__restrict__ melT, const UttMeta *__restrict__ utts,
                                                                    const ResTile *__restrict__ tiles, Dims d, CondPack p,
                                                                    float *__restrict__ res_out) {
    extern __shared__ __align__(16) float sm[];
    const int C = d.C, F = d.F, ks = d.ksize, R = d.R, tid = threadIdx.x;
    float *wbuf = sm;                                    // [2][RES_WBUF]
    float *in = wbuf + 2 * RES_WBUF;                     // [F*ks][RES_TT]   im2col of the mel tile
    float *a0 = in + (size_t)F * ks * RES_TT;            // [C][RES_TT]
    float *a1 = a0 + (size_t)C * RES_TT;                 // [C][RES_TT]
    const ResTile tl = tiles[blockIdx.x];
    const UttMeta um = utts[tl.utt];
    const int t0 = tl.t0, T = um.T, Tp = um.Tp;
    const int npass = 2 * d.nres + 2;
    ResPass ps = res_pass(d, p, 0);
    res_stage(wbuf, ps, 0, tid);                         // first weights in flight while the tile is gathered
    for (int i = tid; i < F * ks * RES_TT; i += RES_THREADS) {
        const int t = i % RES_TT, fj = i / RES_TT, f = fj / ks, j = fj % ks;
        const int tt = t0 + t + j;                       // valid conv over the padded mel
        in[i] = (t0 + t < T && tt < Tp) ? melT[(size_t)(um.pm_off + tt) * F + f] : 0.f;
    }
    const int cl = tid & 127, fh = tid >> 7;
    float acc[RES_MAXCI][4];
    int g = 0;
    for (int pi = 0; pi < npass; pi++) {
        const int KR = RES_WBUF / ps.Cout, nch = (ps.K + KR - 1) / KR, nci = (ps.Cout + 127) / 128;
        const float *act = pi == 0 ? in : ((pi & 1) ? a0 : a1);          // odd passes (conv1, conv_out) read a0, even read a1
#pragma unroll
        for (int ci = 0; ci < RES_MAXCI; ci++)
#pragma unroll
            for (int t = 0; t < 4; t++) acc[ci][t] = 0.f;
        for (int ch = 0; ch < nch; ch++, g++) {
            // next chunk (of this pass or the first of the next) into the other buffer
            const bool more = ch + 1 < nch || pi + 1 < npass;
            if (more) {
                const ResPass nx = ch + 1 < nch ? ps : res_pass(d, p, pi + 1);
                res_stage(wbuf + ((g + 1) & 1) * RES_WBUF, nx, ch + 1 < nch ? ch + 1 : 0, tid);
                asm volatile("cp.async.wait_group 1;" ::: "memory");
            } else {
                asm volatile("cp.async.wait_group 0;" ::: "memory");
            }
            __syncthreads();
            const float *wb = wbuf + (g & 1) * RES_WBUF;
            const int k0 = ch * KR, rows = min(KR, ps.K - k0);
            for (int k = 0; k < rows; k++) {
                const float4 a = *reinterpret_cast<const float4 *>(act + (size_t)(k0 + k) * RES_TT + 4 * fh);
#pragma unroll
                for (int ci = 0; ci < RES_MAXCI; ci++) {
                    const int c = cl + 128 * ci;
                    if (ci < nci && c < ps.Cout) {
                        const float wv = wb[k * ps.Cout + c];
                        acc[ci][0] = fmaf(wv, a.x, acc[ci][0]);
                        acc[ci][1] = fmaf(wv, a.y, acc[ci][1]);
                        acc[ci][2] = fmaf(wv, a.z, acc[ci][2]);
                        acc[ci][3] = fmaf(wv, a.w, acc[ci][3]);
                    }
                }
            }
            __syncthreads();
        }
        // ---- epilogue of the pass
#pragma unroll
        for (int ci = 0; ci < RES_MAXCI; ci++) {
            const int c = cl + 128 * ci;
            if (ci >= nci || c >= ps.Cout) continue;
            if (pi == npass - 1) {                       // conv_out (+ bias) -> res_out
                const float b = p.conv_out_b[c];
#pragma unroll
                for (int t = 0; t < 4; t++)
                    if (t0 + 4 * fh + t < T) res_out[(size_t)(um.res_off + t0 + 4 * fh + t) * R + c] = acc[ci][t] + b;
            } else if (pi == 0 || (pi & 1)) {            // conv_in / conv1: BN + ReLU -> a0 / a1
                const float s = p.bn_scale[(size_t)pi * C + c], b = p.bn_shift[(size_t)pi * C + c];
                float *dst = pi == 0 ? a0 : a1;
#pragma unroll
                for (int t = 0; t < 4; t++) dst[c * RES_TT + 4 * fh + t] = fmaxf(fmaf(acc[ci][t], s, b), 0.f);
            } else {                                     // conv2: BN + residual -> a0 (each element read and written by its owner)
                const float s = p.bn_scale[(size_t)pi * C + c], b = p.bn_shift[(size_t)pi * C + c];
#pragma unroll
                for (int t = 0; t < 4; t++) a0[c * RES_TT + 4 * fh + t] = fmaf(acc[ci][t], s, b) + a0[c * RES_TT + 4 * fh + t];
            }
        }
        if (pi + 1 < npass) ps = res_pass(d, p, pi + 1);
        // the next pass's first __syncthreads (after its cp.async wait) orders these writes before its reads
    }
}

int launch_resnet_batch(wrnn_handle *h, const float *melT, const UttMeta *utts, const ResTile *tiles, int n_tiles,
                        float *res_out, cudaStream_t st) {
    const Dims &d = h->d;
    if (d.C % 4 || d.R % 4 || d.C > 128 * RES_MAXCI || d.R > 128 * RES_MAXCI)
        return fail(h, WRNN_ERR_ARG, "compute_dims / res_out_dims must be multiples of 4 and <= %d", 128 * RES_MAXCI);
    const size_t smem = (2 * (size_t)RES_WBUF + ((size_t)d.F * d.ksize + 2 * (size_t)d.C) * RES_TT) * sizeof(float);
    if (smem > h->smem_optin) return fail(h, WRNN_ERR_ARG, "MelResNet tile needs %zu B of shared memory", smem);
    WRNN_CUDA(h, cudaFuncSetAttribute(resnet_batch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    resnet_batch_kernel<<<n_tiles, RES_THREADS, smem, st>>>(melT, utts, tiles, d, h->cp, res_out);
    h->launches++;
    WRNN_CUDA(h, cudaGetLastError());
    return 0;
}

// ============================================================================ mel upsampling (polyphase) + repeat
// Four consecutive features of upsampled sample i (stage coordinates: i = l + indent) of one utterance:
//   up[i][f] = sum_k poly[i % hop][k] * melT[i / hop + kmin + k][f]      (zero outside the padded mel)
// which is UpsampleNetwork's mel branch (fatchord_version.py:86-88), all stages at once (see pack_cond).
__device__ __forceinline__ float4 upsample4(const float *__restrict__ melT, int64_t pm_off, int Tp, int F, const float *__restrict__ poly,
                                            int taps, int kmin, int hop, int64_t i, int f) {
    const int fr = (int)(i / hop), ph = (int)(i - (int64_t)fr * hop);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int k = 0; k < taps; k++) {
        const int j = fr + kmin + k;
        if (j < 0 || j >= Tp) continue;
        const float wv = __ldg(poly + ph * taps + k);
        const float4 m = __ldg(reinterpret_cast<const float4 *>(melT + (size_t)(pm_off + j) * F + f));
        acc.x = fmaf(wv, m.x, acc.x);
        acc.y = fmaf(wv, m.y, acc.y);
        acc.z = fmaf(wv, m.z, acc.z);
        acc.w = fmaf(wv, m.w, acc.w);
    }
    return acc;
}

// component entry (wrnn_upsample): the upsampled mel of one utterance, time-major [L][F]
__global__ void upsample_flat_kernel(const float *__restrict__ melT, int Tp, int F, const float *__restrict__ poly, int taps, int kmin,
                                     int hop, int64_t indent, int64_t L, float *__restrict__ out) {
    const int W = F / 4;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= L * W) return;
    const int64_t l = i / W;
    const int f = (int)(i % W) * 4;
    *reinterpret_cast<float4 *>(out + (size_t)l * F + f) = upsample4(melT, 0, Tp, F, poly, taps, kmin, hop, l + indent, f);
}

int launch_upsample_flat(wrnn_handle *h, const float *melT, int Tp, int64_t L, float *up_mels, cudaStream_t st) {
    const Dims &d = h->d;
    if (d.F % 4) return fail(h, WRNN_ERR_ARG, "feat_dims must be a multiple of 4");
    const int64_t n = L * (d.F / 4);
    upsample_flat_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(melT, Tp, d.F, h->cp.poly, h->cp.poly_taps, h->cp.poly_kmin,
                                                                   d.total_scale, d.indent, L, up_mels);
    h->launches++;
    WRNN_CUDA(h, cudaGetLastError());
    return 0;
}

__global__ void repeat_aux_kernel(const float *__restrict__ res_out, int64_t L, int hop, int R,
                                  float *__restrict__ aux) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= L * R) return;
    const int64_t l = i / R;
    aux[i] = res_out[(size_t)(l / hop) * R + (i % R)];
}

int launch_repeat_aux(wrnn_handle *h, const float *res_out, int T, float *aux, cudaStream_t st) {
    const int64_t L = (int64_t)T * h->d.total_scale, n = L * h->d.R;
    repeat_aux_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(res_out, L, h->d.total_scale, h->d.R, aux);
    h->launches++;
    WRNN_CUDA(h, cudaGetLastError());
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

// The fold the loop consumes: mel features and the (frame-rate) aux features of fold b, step s, as one [B][S][F+R] row,
// so the loop reads 832 contiguous bytes per fold-step.  Nothing upsampled is materialised in between: the mel part is
// formed here from the padded frames by the polyphase table (UpsampleNetwork's three repeat + FIR stages and the crop,
// fatchord_version.py:86-88), the x275 nearest-repeat of aux (Stretch2d, :83-85) is an index.  One launch for all
// folds of all utterances of the call; HBM-bound on the row writes (float4 stores, a warp covers whole rows).
__global__ void fold_cond_kernel(const float *__restrict__ melT, const float *__restrict__ res_out,
                                 const FoldMeta *__restrict__ meta, int S, int F, int R, int hop, int indent,
                                 const float *__restrict__ poly, int taps, int kmin, float *__restrict__ cond) {
    // grid = (ceil(S * W / 256), B): the fold is the block row, everything inside a fold is 32-bit arithmetic
    const int W = (F + R) / 4;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= S * W) return;
    const int s = i / W, e = (i - s * W) * 4;
    const int64_t b = blockIdx.y;
    const FoldMeta m = meta[b];
    const int64_t l = m.start + s;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (l < m.L) {
        const int li = (int)l;                     // launch_fold_cond checks that L + indent fits 31 bits
        if (e < F) {
            const int ii = li + indent, fr = ii / hop, ph = ii - fr * hop;
            for (int k = 0; k < taps; k++) {
                const int j = fr + kmin + k;
                if (j < 0 || j >= m.Tp) continue;
                const float wv = __ldg(poly + ph * taps + k);
                const float4 q = __ldg(reinterpret_cast<const float4 *>(melT + (size_t)(m.pm_off + j) * F + e));
                v.x = fmaf(wv, q.x, v.x);
                v.y = fmaf(wv, q.y, v.y);
                v.z = fmaf(wv, q.z, v.z);
                v.w = fmaf(wv, q.w, v.w);
            }
        } else {
            v = __ldg(reinterpret_cast<const float4 *>(res_out + (size_t)(m.res_off + li / hop) * R + (e - F)));
        }
    }
    __stcs(reinterpret_cast<float4 *>(cond + ((size_t)b * S * W + i) * 4), v);     // streamed: the loop reads each row once, later
}

int launch_fold_cond(wrnn_handle *h, const float *melT, const float *res_out, const FoldMeta *meta,
                     int64_t B, int64_t S, float *cond, cudaStream_t st) {
    const Dims &d = h->d;
    if (d.F % 4 || d.R % 4) return fail(h, WRNN_ERR_ARG, "feat_dims and res_out_dims must be multiples of 4");
    const int W = (d.F + d.R) / 4;
    int64_t Lmax = 0;
    for (auto &ut : h->utts) Lmax = ut.L > Lmax ? ut.L : Lmax;
    if (Lmax + d.indent + S >= 0x7fffffff || S * W >= 0x7fffffff || B > 65535 * 32768ll)
        return fail(h, WRNN_ERR_ARG, "utterance too long for the conditioning kernel");
    for (int64_t b0 = 0; b0 < B; b0 += 65535) {           // gridDim.y limit
        const int64_t nb = B - b0 < 65535 ? B - b0 : 65535;
        dim3 grid((unsigned)((S * W + 255) / 256), (unsigned)nb);
        fold_cond_kernel<<<grid, 256, 0, st>>>(melT, res_out, meta + b0, (int)S, d.F, d.R, d.total_scale, (int)d.indent, h->cp.poly,
                                               h->cp.poly_taps, h->cp.poly_kmin, cond + (size_t)b0 * S * (d.F + d.R));
        h->launches++;
    }
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
                            const double *__restrict__ lut, double *__restrict__ out, int64_t i0, int64_t i1) {
    const TailMeta m = meta[blockIdx.y];
    const int64_t i = i0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;       // [i0, i1): all of it, or one streamed chunk
    if (i >= m.wave_len || i >= i1) return;
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
                cudaStream_t st, int64_t first) {
    // samples [first, max_wave) of every utterance (first > 0: a streamed chunk of one utterance)
    dim3 grid((unsigned)((max_wave - first + 255) / 256), n_utt);
    tail_kernel<<<grid, 256, 0, st>>>(samples, picks, meta, S, batched, target, overlap, fade_in, fade_out, fo,
                                      fo_len, lut, out, first, max_wave);
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

// C ABI of libwavernn_b200.so (include/wavernn_b200.h): handle lifetime, checkpoint ingest,
// orchestration of conditioning -> fold -> per-sample loop -> float64 tail.  No CPU fallback:
// every compute entry point needs a CUDA device and fails with WRNN_ERR_CUDA otherwise.
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>

#include "internal.h"

namespace wrnn {

static thread_local std::string g_create_err;

int fail(wrnn_handle *h, int code, const char *fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (h) h->err = buf;
    else g_create_err = buf;
    if (code == WRNN_ERR_CUDA) cudaGetLastError();   // clear the sticky (non-fatal) error state
    return code;
}

int ws_reserve(wrnn_handle *h, Workspace &w, size_t bytes) {
    if (bytes <= w.bytes && w.p) return 0;
    if (w.p) cudaFree(w.p);
    w.p = nullptr;
    w.bytes = 0;
    const size_t want = bytes + bytes / 8 + 256;
    cudaError_t e = cudaMalloc(&w.p, want);
    if (e != cudaSuccess) {
        w.p = nullptr;
        return fail(h, WRNN_ERR_NOMEM, "cudaMalloc(%zu) failed: %s", want, cudaGetErrorString(e));
    }
    w.bytes = want;
    return 0;
}

// numpy.linspace(start, stop, n) semantics (float64): y[i] = i*step + start, y[n-1] = stop.
void host_linspace(double start, double stop, int64_t n, double *out) {
    if (n <= 0) return;
    if (n == 1) { out[0] = start; return; }
    const double step = (stop - start) / (double)(n - 1);
    for (int64_t i = 0; i < n; i++) {
        volatile double prod = (double)i * step;   // two roundings, never an FMA
        out[i] = prod + start;
    }
    out[n - 1] = stop;
}

static int64_t num_folds(int64_t total_len, int64_t target, int64_t overlap) {
    // fatchord_version.py:322-328
    int64_t n = (total_len - overlap) / (target + overlap);
    const int64_t extended = n * (overlap + target) + overlap;
    if (total_len - extended != 0) n += 1;
    return n;
}

int loop_kernel_choice(const wrnn_handle *h, int64_t folds) {
    const int f = h->force_kernel;
    if (f == 4 && mma_supported(h)) return 4;
    if (f == 0 && mma_supported(h) && folds >= mma_min_folds()) return 4;
    if (f == 1) return 1;
    if (f == 2) return pipe_supported(h) ? 2 : 1;
    if (f == 3 && res_supported(h)) return 3;
    if (res_supported(h)) return 3;
    if (pipe_supported(h)) return 2;
    return 1;
}

int ensure_packed_public(wrnn_handle *h) {
    if (h->packed) return 0;
    if (h->w.empty()) return fail(h, WRNN_ERR_STATE, "weights not loaded: call wrnn_load_weights first");
    WRNN_DEVICE(h);
    WRNN_TRY(pack_cond(h));
    WRNN_TRY(fma_plan(h));
    WRNN_TRY(pipe_plan(h));
    WRNN_TRY(res_plan(h));
    WRNN_TRY(mma_plan(h));
    h->packed = true;
    return 0;
}

}  // namespace wrnn

using namespace wrnn;

extern "C" {

int wrnn_create(const wrnn_config *cfg, int device, wrnn_handle **out) {
    if (!cfg || !out) return fail(nullptr, WRNN_ERR_ARG, "null argument");
    *out = nullptr;
    if (cfg->mode != WRNN_MODE_RAW && cfg->mode != WRNN_MODE_MOL)
        return fail(nullptr, WRNN_ERR_ARG, "Unknown model mode value - %d", cfg->mode);
    if (cfg->n_upsample < 1 || cfg->n_upsample > 8 || cfg->rnn_dims < 1 || cfg->fc_dims < 1 || cfg->feat_dims < 1 ||
        cfg->compute_dims < 1 || cfg->res_out_dims < 4 || cfg->res_out_dims % 4 || cfg->pad < 0 || cfg->res_blocks < 0)
        return fail(nullptr, WRNN_ERR_ARG, "bad model dimensions");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(nullptr, WRNN_ERR_CUDA, "no CUDA device (%s): this library has no CPU fallback",
                    cudaGetErrorString(e));
    if (device < 0 || device >= ndev) return fail(nullptr, WRNN_ERR_ARG, "device %d out of range (%d)", device, ndev);
    wrnn_handle *h = new wrnn_handle();
    h->cfg = *cfg;
    h->device = device;
    Dims &d = h->d;
    d.H = cfg->rnn_dims; d.FC = cfg->fc_dims; d.F = cfg->feat_dims; d.R = cfg->res_out_dims; d.A = d.R / 4;
    d.C = cfg->compute_dims; d.pad = cfg->pad; d.hop = cfg->hop_length; d.nres = cfg->res_blocks;
    d.nup = cfg->n_upsample; d.mode = cfg->mode; d.bits = cfg->bits; d.ksize = 2 * cfg->pad + 1;
    d.total_scale = 1;
    for (int i = 0; i < 8; i++) d.up[i] = i < d.nup ? cfg->upsample_factors[i] : 1;
    for (int i = 0; i < d.nup; i++) d.total_scale *= d.up[i];
    d.indent = d.pad * d.total_scale;
    d.NC = cfg->mode == WRNN_MODE_RAW ? (1 << cfg->bits) : 30;
    if (d.total_scale != d.hop) {
        delete h;
        return fail(nullptr, WRNN_ERR_ARG, "upsample_factors must multiply to hop_length");
    }
    cudaDeviceProp prop;
    DeviceGuard guard(device);
    if (guard.err != cudaSuccess || cudaGetDeviceProperties(&prop, device) != cudaSuccess) {
        delete h;
        return fail(nullptr, WRNN_ERR_CUDA, "cannot open device %d", device);
    }
    h->num_sms = prop.multiProcessorCount;
    h->smem_optin = prop.sharedMemPerBlockOptin;
    for (int i = 0; i < 4; i++) cudaEventCreate(&h->ev[i]);
    *out = h;
    return 0;
}

void wrnn_destroy(wrnn_handle *h) {
    if (!h) return;
    DeviceGuard guard(h->device);
    Workspace *all[] = {&h->ws_mel, &h->ws_up, &h->ws_res, &h->ws_cond, &h->ws_samples, &h->ws_picks, &h->ws_out,
                        &h->ws_noise, &h->ws_meta, &h->ws_tables, &h->ws_sync, &h->ws_mma_sync, &h->ws_act, &h->ws_tmp0, &h->ws_tmp1,
                        &h->ws_teacher, &h->ws_logits, &h->ws_prof, &h->mel_tab};
    for (Workspace *w : all)
        if (w->p) cudaFree(w->p);
    float *ptrs[] = {h->cp.poly, h->cp.conv_in_t, h->cp.bn_scale, h->cp.bn_shift, h->cp.res_t, h->cp.conv_out_t, h->cp.conv_out_b,
                     h->fma.image, h->pipe.row_image, h->pipe.tail_image, h->res.img, h->res.tail_img,
                     (float *)h->mma.Aimg, h->mma.bias, h->mma.wf3};
    for (float *p : ptrs)
        if (p) cudaFree(p);
    for (int i = 0; i < 8; i++)
        if (h->cp.fir[i]) cudaFree(h->cp.fir[i]);
    for (int i = 0; i < 4; i++)
        if (h->ev[i]) cudaEventDestroy(h->ev[i]);
    if (h->h_pinned_out) cudaFreeHost(h->h_pinned_out);
    if (h->h_pinned_in) cudaFreeHost(h->h_pinned_in);
    delete h;
}

const char *wrnn_last_error(const wrnn_handle *h) { return h ? h->err.c_str() : g_create_err.c_str(); }

int wrnn_load_weights(wrnn_handle *h, const wrnn_tensor *tensors, int32_t n) {
    if (!h || (!tensors && n > 0)) return fail(h, WRNN_ERR_ARG, "null argument");
    for (int i = 0; i < n; i++) {
        const wrnn_tensor &t = tensors[i];
        if (!t.name || !t.data || t.ndim < 0 || t.ndim > 4) return fail(h, WRNN_ERR_ARG, "bad tensor %d", i);
        if (t.dtype != 0) continue;       // int64 counters (step, num_batches_tracked) are not used by the path
        HostTensor ht;
        int64_t numel = 1;
        for (int k = 0; k < t.ndim; k++) { ht.shape.push_back(t.shape[k]); numel *= t.shape[k]; }
        ht.v.assign((const float *)t.data, (const float *)t.data + numel);
        h->w[t.name] = std::move(ht);
    }
    h->packed = false;     // repack on next use (`load()` may be called again, SURVEY §8b ownership)
    h->sparse = false;
    return 0;
}

int wrnn_mulaw_lut(int32_t mu, double *lut) {
    if (mu < 2 || !lut) return WRNN_ERR_ARG;
    // decode_mu_law(y, mu, from_labels=False) at y = float32(2k/(mu-1) - 1)  (utils/dsp.py:98-103)
    const double m = (double)(mu - 1);
    for (int k = 0; k < mu; k++) {
        volatile float q = (2.f * (float)k) / ((float)mu - 1.f);
        const float yf = q - 1.f;
        const double y = (double)yf;
        const double sgn = y > 0 ? 1.0 : (y < 0 ? -1.0 : 0.0);
        volatile double a = sgn / m;
        volatile double p = pow(1.0 + m, fabs(y)) - 1.0;
        lut[k] = a * p;
    }
    return 0;
}

int wrnn_set_mulaw_lut(wrnn_handle *h, int32_t mu, const double *lut) {
    if (!h || !lut || mu < 2) return fail(h, WRNN_ERR_ARG, "bad argument");
    h->lut_override.assign(lut, lut + mu);
    return 0;
}

int wrnn_xfade_tables(int64_t overlap, double *fade_in, double *fade_out) {
    if (overlap < 0 || !fade_in || !fade_out) return WRNN_ERR_ARG;
    // fatchord_version.py:379-391
    const int64_t silence = overlap / 2, fade = overlap - silence;
    std::vector<double> t((size_t)std::max<int64_t>(fade, 1));
    host_linspace(-1.0, 1.0, fade, t.data());
    for (int64_t i = 0; i < silence; i++) { fade_in[i] = 0.0; fade_out[i] = 1.0; }
    for (int64_t i = 0; i < fade; i++) {
        volatile double a = 1 + t[i], b = 1 - t[i];
        volatile double ha = 0.5 * a, hb = 0.5 * b;
        fade_in[silence + i] = sqrt(ha);
        fade_out[silence + i] = sqrt(hb);
    }
    return 0;
}

int64_t wrnn_num_folds(int64_t total_len, int64_t target, int64_t overlap) {
    if (target + overlap <= 0) return -1;
    return num_folds(total_len, target, overlap);
}

const char *wrnn_loop_kernel_name(wrnn_handle *h, int64_t folds) {
    if (!h) return "none";
    if (!h->packed && !h->w.empty()) ensure_packed_public(h);
    const int k = loop_kernel_choice(h, folds);
    if (k == 4) return "tcgen05-bf16x3";
    return k == 3 ? (h->res.sparse ? "fma-resident-sparse16x1" : "fma-resident") : (k == 2 ? "fma-pipelined" : "fma");
}

int wrnn_set_loop_kernel(wrnn_handle *h, int32_t which) {
    if (!h) return WRNN_ERR_ARG;
    h->force_kernel = which;
    return 0;
}

int wrnn_debug_mma_state(wrnn_handle *h, int32_t which, int64_t folds, float *out) {
    if (!h || !out || which < 0 || which > 2 || folds < 1) return fail(h, WRNN_ERR_ARG, "bad argument");
    if (!h->mma_dbg[which]) return fail(h, WRNN_ERR_STATE, "no tensor-core run yet");
    WRNN_DEVICE(h);
    WRNN_CUDA(h, cudaDeviceSynchronize());
    WRNN_CUDA(h, cudaMemcpy(out, h->mma_dbg[which], (size_t)folds * 512 * sizeof(float), cudaMemcpyDeviceToHost));
    return 0;
}

int wrnn_exchange_calibration(wrnn_handle *h, float *l2_round_trip_clk, float *delay_scale) {
    if (!h) return WRNN_ERR_ARG;
    if (l2_round_trip_clk) *l2_round_trip_clk = h->xchg_l2_clk;
    if (delay_scale) *delay_scale = h->xchg_scale;
    return 0;
}

int wrnn_set_mma_fold_tile(wrnn_handle *h, int32_t nt) {
    if (!h) return WRNN_ERR_ARG;
    if (nt != 0 && (nt < 16 || nt > 128 || nt % 16)) return fail(h, WRNN_ERR_ARG, "fold tile must be 0 (auto) or a multiple of 16 in [16, 128]");
    h->mma_force_nt = nt;
    return 0;
}

int wrnn_prune_16x1(wrnn_handle *h, float keep_fraction, float *masked_out) {
    // SURVEY.md §8a row S (the reference has no block-sparse code; nearest: per-gate magnitude pruning in
    // notebooks/Pruning - Scratchpad.ipynb:92-108).  Per matrix and per gate (H-row split), blocks of 16
    // consecutive output rows x 1 input column; keep round(keep_fraction * nblocks) blocks of largest L1 norm
    // (ties: lower block index wins), zero the rest.
    if (!h) return WRNN_ERR_ARG;
    if (!(keep_fraction > 0.f) || keep_fraction > 1.f) return fail(h, WRNN_ERR_ARG, "keep_fraction must be in (0, 1]");
    const int H = h->d.H;
    if (H % 16) return fail(h, WRNN_ERR_ARG, "rnn_dims must be a multiple of 16 for 16x1 block pruning");
    const char *keys[4] = {"rnn1.weight_ih_l0", "rnn1.weight_hh_l0", "rnn2.weight_ih_l0", "rnn2.weight_hh_l0"};
    size_t out_off = 0;
    for (int m = 0; m < 4; m++) {
        auto it = h->w.find(keys[m]);
        if (it == h->w.end() || it->second.shape.size() != 2 || it->second.shape[0] != 3 * H)
            return fail(h, WRNN_ERR_STATE, "checkpoint key missing or misshapen: %s", keys[m]);
        HostTensor &W = it->second;
        const int64_t K = W.shape[1], nbr = H / 16, nblocks = nbr * K;
        const int64_t keep = std::min<int64_t>(nblocks, std::max<int64_t>(1, (int64_t)floor((double)keep_fraction * (double)nblocks + 0.5)));
        std::vector<std::pair<double, int64_t>> norm((size_t)nblocks);
        for (int g = 0; g < 3; g++) {
            float *base = W.v.data() + (size_t)g * H * K;
            for (int64_t br = 0; br < nbr; br++)
                for (int64_t c = 0; c < K; c++) {
                    double s = 0.0;
                    for (int r = 0; r < 16; r++) s += fabs((double)base[(size_t)(br * 16 + r) * K + c]);
                    norm[(size_t)(br * K + c)] = {s, br * K + c};
                }
            std::vector<std::pair<double, int64_t>> sorted = norm;
            std::sort(sorted.begin(), sorted.end(), [](const std::pair<double, int64_t> &a, const std::pair<double, int64_t> &b) {
                return a.first > b.first || (a.first == b.first && a.second < b.second);
            });
            std::vector<char> keepmask((size_t)nblocks, 0);
            for (int64_t i = 0; i < keep; i++) keepmask[(size_t)sorted[(size_t)i].second] = 1;
            for (int64_t br = 0; br < nbr; br++)
                for (int64_t c = 0; c < K; c++)
                    if (!keepmask[(size_t)(br * K + c)])
                        for (int r = 0; r < 16; r++) base[(size_t)(br * 16 + r) * K + c] = 0.f;
        }
        if (masked_out) memcpy(masked_out + out_off, W.v.data(), W.v.size() * sizeof(float));
        out_off += W.v.size();
    }
    h->packed = false;
    h->sparse = true;
    return 0;
}

}  // extern "C"

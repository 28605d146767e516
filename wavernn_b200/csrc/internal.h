// Internal declarations shared by the .cu files of libwavernn_b200.so (not part of the ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <map>
#include <string>
#include <vector>

#include "../../include/wavernn_b200.h"

namespace wrnn {

struct Dims {
    int H, FC, F, A, R, C, NC, pad, hop, nres, nup, up[8], total_scale, indent, mode, bits, ksize;
};

struct HostTensor {            // checkpoint tensor kept on the host; packers upload device images
    std::vector<float> v;
    std::vector<int64_t> shape;
};

// ---- packed conditioning weights (cond.cu) -------------------------------------------------
struct CondPack {
    // all [k][c] (input index major, output channel fastest) so thread c reads coalesced
    float *conv_in_t = nullptr;        // [(F*ksize)][C]
    float *bn_scale = nullptr;         // [1 + 2*nres][C]  gamma / sqrt(var + eps)
    float *bn_shift = nullptr;         // [1 + 2*nres][C]  beta - mean * scale
    float *res_t = nullptr;            // [2*nres][C][C]
    float *conv_out_t = nullptr;       // [C][R]
    float *conv_out_b = nullptr;       // [R]
    float *fir[8] = {nullptr};         // [2*s+1] per stage
    // the whole mel branch of UpsampleNetwork as ONE polyphase filter over frames (cond.cu, pack_cond):
    // up[i] = sum_k poly[i % hop][k] * mel_padded[i / hop + poly_kmin + k]
    float *poly = nullptr;             // [hop][poly_taps]
    int poly_taps = 0, poly_kmin = 0;
};

// ---- per-sample loop, FMA kernel (loop_fma.cu) ----------------------------------------------
struct FmaPlan {
    int G = 0;                 // grid (CTAs), one per SM
    int upc = 0, rF = 0, r3 = 0;   // GRU units / fc rows / fc3 rows per CTA
    int KI = 0, KH = 0, KHA = 0, KFA = 0, KF = 0;   // padded (multiple of 4) reduction lengths
    int64_t img_floats = 0;    // per-CTA weight image size
    int64_t off_I, off_r1i, off_r1h, off_r2i, off_r2h, off_f1, off_f2, off_f3, off_bias;
    float *image = nullptr;    // device [G][img_floats]
    size_t smem_bytes = 0;
};

// ---- per-sample loop, phase-pipelined kernel (loop_pipe.cu) -----------------------------------
struct PipePlan {
    bool ok = false;           // dims supported and images built
    int G = 0, NR = 0, NT = 0, upc = 0, rF = 0;
    int Kc = 0, ldc = 0, KH = 0, ldh = 0, KX1 = 0, ldx1 = 0, KX2 = 0, ldx2 = 0, ldx = 0, KF = 0, ldf3 = 0;
    int off_Wc = 0, off_w0 = 0, off_hh1 = 0, off_ih2 = 0, off_hh2 = 0, off_f1 = 0, off_f2 = 0, off_bias = 0;
    int img_floats = 0, tail_floats = 0;
    float *row_image = nullptr;    // device [NR][img_floats]
    float *tail_image = nullptr;   // device fc3 rows + bias
    size_t smem_bytes = 0;
};

// ---- per-sample loop, register-resident warp-specialised kernel (loop_res.cu) ----------------
struct ResPlan {
    bool ok = false;
    bool sparse = false;           // every (CTA, gate) of the three GRU matrices keeps <= 80 columns: 16x1 block-sparse tables in use
    int G = 0;
    float *img = nullptr;          // device [128][IMG_FLOATS]
    float *tail_img = nullptr;     // device fc3 rows + bias
    size_t smem_bytes = 0;
};

// ---- per-sample loop, tensor-core kernel for large fold batches (loop_mma.cu) ------------------
struct MmaPlan {
    bool ok = false;
    uint8_t *Aimg = nullptr;       // device: bf16 hi/lo weight images of the four MMA phases
    int64_t offA[10] = {0};              // one image per job type (loop_mma.cu: J_R1A .. J_F2C)
    int64_t offP3 = 0;                   // fc3 columns as an MMA operand, per fc row tile
    float *bias = nullptr;         // device: composed biases
    float *wf3 = nullptr;          // device: fc3 rows (fp32) + bias
    size_t img_bytes = 0, smem_bytes = 0;
};

struct LoopArgs {
    const float *cond;         // [B][S][F+R]; null for the tensor-core kernel, which forms its own operand from:
    const float *melT = nullptr, *res_out = nullptr;     // padded frames [.][F] and MelResNet rows [.][R] of the call
    const void *fold_meta = nullptr;                     // FoldMeta[B] (device) of the folds of this launch
    int64_t B, S;
    int64_t Btot, b0;          // fold count of the whole call and this launch's first fold (noise / Philox addressing)
    const int32_t *fold_len;   // [B] steps valid per fold (== S when batched)
    const float *u1, *u2, *q;  // injected noise or null
    const float *teacher;      // [B][S] or null
    float *logits_out;         // [B][S][NC] or null
    float *samples;            // [B][S]
    int32_t *picks;            // [B][S] (RAW class index / MoL mixture index)
    uint64_t seed;
};

struct Workspace {             // grows on demand, never shrinks
    void *p = nullptr;
    size_t bytes = 0;
};

}  // namespace wrnn

struct wrnn_handle {
    wrnn_config cfg;
    wrnn::Dims d;
    int device = 0;
    int num_sms = 0;
    size_t smem_optin = 0;
    std::string err;
    std::map<std::string, wrnn::HostTensor> w;     // raw checkpoint tensors (host copies)
    bool packed = false;
    wrnn::CondPack cp;
    wrnn::FmaPlan fma;
    wrnn::PipePlan pipe;
    wrnn::ResPlan res;
    wrnn::MmaPlan mma;
    float *mma_dbg[3] = {nullptr, nullptr, nullptr};   // h1, h2, fc2 output [folds][512] fp32 of the last tensor-core run
    float xchg_l2_clk = 0.f, xchg_scale = 1.f;   // loop_res.cu calibrate_exchange: measured L2 round trip (SM clocks), delay scale
    int mma_force_nt = 0;              // debug / tests: fold-tile width of the tensor-core kernel (0 = auto)
    int force_kernel = 0;
    std::vector<double> lut_override;   // host-supplied mu-law table (wrnn_set_mulaw_lut)
    bool sparse = false;

    // staged call
    struct Utt { int T, Tin, Tp; int64_t L, wave_len, fold0, nfolds, mel_off, up_off, res_off, out_off, pm_off; };
    int64_t total_frames = 0;          // padded frames of all staged utterances
    int64_t range_lo = 0, range_hi = 0;    // wrnn_set_fold_range: [lo, hi) of the next single-utterance batched stage, 0/0 = all
    int64_t fold_lo = 0, Btot = 0;         // staged: first global fold of this call, fold count of the whole utterance(s)
    std::vector<Utt> utts;
    bool staged = false;
    int batched = 0, mu_law = 0;
    int64_t target = 0, overlap = 0, B = 0, S = 0, total_out = 0;
    bool have_noise = false;

    wrnn::Workspace ws_mel, ws_up, ws_res, ws_cond, ws_samples, ws_picks, ws_out, ws_noise,
        ws_meta, ws_tables, ws_sync, ws_mma_sync, ws_act, ws_tmp0, ws_tmp1, ws_teacher, ws_logits, ws_prof, mel_tab;
    wrnn_mel_params mel_tab_params = {0, 0, 0, 0, 0, 0.f, 0.f};     // what mel_tab (window, roots of unity, mel basis) was built for
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    int64_t launches = 0;
    double *h_pinned_out = nullptr;
    size_t h_pinned_out_bytes = 0;
    float *h_pinned_in = nullptr;
    size_t h_pinned_in_bytes = 0;
};

namespace wrnn {

// error plumbing: functions return 0 / negative status and fill h->err
int fail(wrnn_handle *h, int code, const char *fmt, ...);
#define WRNN_CUDA(h, expr)                                                                    \
    do {                                                                                      \
        cudaError_t e__ = (expr);                                                             \
        if (e__ != cudaSuccess)                                                               \
            return wrnn::fail((h), WRNN_ERR_CUDA, "%s failed: %s (%s:%d)", #expr,            \
                              cudaGetErrorString(e__), __FILE__, __LINE__);                   \
    } while (0)
#define WRNN_TRY(expr)                                                                        \
    do {                                                                                      \
        int r__ = (expr);                                                                     \
        if (r__ != 0) return r__;                                                             \
    } while (0)

// The current CUDA device is thread-global state shared with the host process (PyTorch): every entry point selects the
// handle's device for the duration of the call and puts the caller's device back on every exit path.
struct DeviceGuard {
    int prev = -1;
    cudaError_t err = cudaSuccess;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
        if (prev != dev) err = cudaSetDevice(dev);
        else prev = -1;                 // nothing to restore
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
    DeviceGuard(const DeviceGuard &) = delete;
    DeviceGuard &operator=(const DeviceGuard &) = delete;
};
#define WRNN_DEVICE(h)                                                                          \
    wrnn::DeviceGuard dev_guard__((h)->device);                                                 \
    if (dev_guard__.err != cudaSuccess)                                                         \
        return wrnn::fail((h), WRNN_ERR_CUDA, "cudaSetDevice(%d) failed: %s", (h)->device, cudaGetErrorString(dev_guard__.err))

int ws_reserve(wrnn_handle *h, Workspace &w, size_t bytes);
int ensure_packed_public(wrnn_handle *h);   // pack weights for the kernels if not yet done

// cond.cu
int pack_cond(wrnn_handle *h);
// mel_padded: device [F][Tp] (Tp = T + 2*pad); up_mels: device [L][F]; res_out: device [T][R]
// Conditioning of ALL utterances of a call in three launches (pad + transpose, MelResNet, conditioning rows).
struct UttMeta { int64_t raw_off, pm_off, res_off; int T, Tp, Tin, pad_; };   // raw mel floats / padded-frame / res-row offsets
struct ResTile { int utt, t0; };                                              // one MelResNet CTA: RES_TT frames of one utterance
int resnet_tile_frames();
// mel_raw: utterance u at raw_off, (F, Tin) row-major; melT: [sum Tp][F] frame-major, zero frames where the pad is
int launch_pad_mel_batch(wrnn_handle *h, const float *mel_raw, const UttMeta *utts, int n_utt, int64_t total_frames,
                         float *melT, cudaStream_t st);
int launch_resnet_batch(wrnn_handle *h, const float *melT, const UttMeta *utts, const ResTile *tiles, int n_tiles,
                        float *res_out, cudaStream_t st);
// component entry: up_mels [L][F] of one utterance (melT at frame offset 0, Tp frames)
int launch_upsample_flat(wrnn_handle *h, const float *melT, int Tp, int64_t L, float *up_mels, cudaStream_t st);
int launch_repeat_aux(wrnn_handle *h, const float *res_out, int T, float *aux, cudaStream_t st);
int launch_fold(wrnn_handle *h, const float *x, int64_t L, int F, int64_t target, int64_t overlap,
                int64_t folds, float *out, cudaStream_t st);
struct FoldMeta { int64_t pm_off, res_off, start, L; int Tp, pad_; };   // per fold: padded-frame / res-row offsets of its utterance
int launch_fold_cond(wrnn_handle *h, const float *melT, const float *res_out, const FoldMeta *meta,
                     int64_t B, int64_t S, float *cond, cudaStream_t st);
struct TailMeta { int64_t fold0, nfolds, wave_len, out_off; };   // per utterance
int launch_tail(wrnn_handle *h, const float *samples, const int32_t *picks, const TailMeta *meta,
                int n_utt, int64_t max_wave, int64_t S, int batched, int64_t target, int64_t overlap,
                const double *fade_in, const double *fade_out, const double *fo, int64_t fo_len,
                const double *lut, double *out, cudaStream_t st, int64_t first = 0);
int launch_unfold(wrnn_handle *h, const double *y, int64_t folds, int64_t target, int64_t overlap,
                  const double *fade_in, const double *fade_out, double *out, cudaStream_t st);
int launch_mulaw(wrnn_handle *h, const float *y, int64_t n, int mu, const double *lut, double *out,
                 cudaStream_t st);

// loop_fma.cu
int fma_plan(wrnn_handle *h);
int fma_run(wrnn_handle *h, const LoopArgs &a, cudaStream_t st);

// loop_pipe.cu
int pipe_plan(wrnn_handle *h);
bool pipe_supported(const wrnn_handle *h);
int pipe_max_folds();
int pipe_run(wrnn_handle *h, const LoopArgs &a, cudaStream_t st);

// loop_res.cu
int res_plan(wrnn_handle *h);
bool res_supported(const wrnn_handle *h);
int res_max_folds();
int res_run(wrnn_handle *h, const LoopArgs &a, cudaStream_t st);
// loop_mma.cu
int mma_plan(wrnn_handle *h);
bool mma_supported(const wrnn_handle *h);
int mma_min_folds();               // fold count from which the tensor-core kernel beats groups of the resident kernel
int mma_run(wrnn_handle *h, const LoopArgs &a, cudaStream_t st);
int64_t mma_max_folds(const wrnn_handle *h);   // folds one launch takes (one GRU job per CTA): larger calls run in groups
// which loop kernel a run will use: 1 = fma (grid barriers, any dims, any fold count), 2 = pipelined
// (smem weights, tagged exchange), 3 = resident (register weights, default dims), 4 = tensor-core (tcgen05, bf16 hi/lo
// operands, default dims, MoL; chosen from mma_min_folds() folds up)
int loop_kernel_choice(const wrnn_handle *h, int64_t folds);

// host tables (api.cu)
void host_linspace(double start, double stop, int64_t n, double *out);

}  // namespace wrnn

/* wavernn_b200.h — C ABI of the B200-native WaveRNN generation path.
 *
 * The reference (qingyundou/WaveRNN) is pure Python and has no FFI; the boundary it exposes for
 * this path is the method `WaveRNN.generate(mels, save_path, batched, target, overlap, mu_law)`
 * (models/fatchord_version.py:169) and `deepmind_version.WaveRNN.generate(seq_len)`
 * (models/deepmind_version.py:75).  This header is what a Python-side binding of that method
 * calls (wavernn_b200/binding.py; INTEGRATION.md shows the stub a reference maintainer adds).
 *
 * Conventions: plain pointers and sizes, no torch types.  Every function returns 0 on success
 * or a negative wrnn_status; the message is available from wrnn_last_error().  No C++ exception
 * crosses the ABI.  Unless a parameter says "device", pointers are HOST pointers owned by the
 * caller; the library owns packed weights and workspaces.  One handle per GPU; a handle is not
 * thread-safe; distinct handles may be driven from distinct host threads (the multi-GPU model:
 * no collective on the path).  There is NO CPU fallback: without a CUDA device every compute
 * entry point fails with WRNN_ERR_CUDA.
 */
#ifndef WAVERNN_B200_H
#define WAVERNN_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    WRNN_OK = 0,
    WRNN_ERR_ARG = -1,      /* bad argument / shape (Python side raises ValueError)          */
    WRNN_ERR_STATE = -2,    /* weights not loaded, wrong mode ... (RuntimeError)               */
    WRNN_ERR_CUDA = -3,     /* CUDA runtime error, no device, kernel fault (RuntimeError)      */
    WRNN_ERR_NOMEM = -4
} wrnn_status;

enum { WRNN_MODE_RAW = 0, WRNN_MODE_MOL = 1 };

/* Constructor arguments of the reference model (fatchord_version.py:93-95). */
typedef struct {
    int32_t rnn_dims, fc_dims, bits, pad;
    int32_t n_upsample;             /* number of entries used in upsample_factors (<= 8)   */
    int32_t upsample_factors[8];
    int32_t feat_dims, compute_dims, res_out_dims, res_blocks;
    int32_t hop_length, sample_rate;
    int32_t mode;                   /* WRNN_MODE_RAW / WRNN_MODE_MOL                        */
} wrnn_config;

/* One named checkpoint tensor, state_dict key and layout unchanged (fatchord_version.py:114-125,
 * loaded like `load()` at :414-417).  dtype: 0 = float32, 1 = int64.  Borrowed, copied inside. */
typedef struct {
    const char *name;
    const void *data;
    int32_t dtype;
    int32_t ndim;
    int64_t shape[4];
} wrnn_tensor;

/* Pre-drawn noise for parity runs, in the reference's draw order (utils/distribution.py:106,118;
 * Categorical.sample == argmax(p / Exp(1))).  steps x folds is the loop extent of the call.
 *   MOL: u1 [steps][folds][10], u2 [steps][folds]      (uniform in [1e-5, 1-1e-5])
 *   RAW: q  [steps][folds][n_classes]                    (Exp(1))
 * NULL noise => in-kernel Philox keyed (seed, fold, step). */
typedef struct {
    const float *u1;
    const float *u2;
    const float *q;
    int64_t steps;
    int64_t folds;
} wrnn_noise;

typedef struct wrnn_handle wrnn_handle;

/* ------------------------------------------------------------------ lifetime */
int wrnn_create(const wrnn_config *cfg, int device, wrnn_handle **out);
void wrnn_destroy(wrnn_handle *h);
const char *wrnn_last_error(const wrnn_handle *h);          /* h may be NULL: create() errors */
/* Replaces `model.load()` / the aliasing in get_gru_cell (fatchord_version.py:273-279,414-417):
 * copies and packs the checkpoint for the kernels.  Unknown keys are ignored (strict=False);
 * every key the path needs must be present by the time generate is called. */
int wrnn_load_weights(wrnn_handle *h, const wrnn_tensor *tensors, int32_t n);
/* 16x1 block pruning of the four GRU matrices: per gate, keep the
 * `keep_fraction` of 16-row x 1-column blocks with largest L1 norm, zero the rest, and switch the
 * loop kernel to the block-sparse mat-vec.  `masked_out` (optional, host) receives the four masked
 * dense matrices concatenated in the order rnn1.ih, rnn1.hh, rnn2.ih, rnn2.hh. */
int wrnn_prune_16x1(wrnn_handle *h, float keep_fraction, float *masked_out);

/* Chunked delivery of one batched utterance (the reference returns only when the last sample
 * is drawn, fatchord_version.py:241-264).  Folds run in groups of `chunk_folds` (0 = 20, the resident kernel's group);
 * after each group the samples that no later fold touches, [done so far, folds_done * (target + overlap)), get their
 * cross-fade (and in the last chunk the fade-out), go to the host and are handed to `on_chunk` in time order while the
 * next groups run.  `samples` points into library-owned pinned memory, valid until the next call on the handle.
 * `out` (may be NULL) receives the whole waveform as wrnn_generate would write it: the chunks concatenated are equal to
 * it bit for bit.  Returns after the last chunk. */
typedef void (*wrnn_chunk_fn)(void *user, int64_t first_sample, const double *samples, int64_t n, int32_t last);
int wrnn_generate_stream(wrnn_handle *h, const float *mel, int32_t T, int32_t target, int32_t overlap, int32_t mu_law,
                         const wrnn_noise *noise, uint64_t seed, int32_t chunk_folds, wrnn_chunk_fn on_chunk, void *user,
                         double *out, int64_t *out_len, void *cuda_stream);

/* melspectrogram(y) of utils/dsp.py:72-75, the input branch of gen_wavernn.py:44-47 for a .wav file: librosa 0.6.3's
 * stft (centre, reflect padding, periodic Hann window of win_length inside an n_fft frame), |D| through the Slaney
 * area-normalised mel basis (librosa.filters.mel(sr, n_fft, n_mels, fmin)), 20 log10(max(1e-5, .)), then
 * clip((S - min_level_db) / -min_level_db, 0, 1).  hparams.py:20-27 gives the reference's values
 * {22050, 2048, 275, 1100, 80, 40.0, -100.0}.
 * wav: n host floats in [-1, 1].  mel_out: [num_mels][frames] host floats, frames = 1 + n / hop_length, written when
 * frames_cap >= frames; mel_out == NULL only returns *frames_out.  Direct DFT in double on the GPU: within 1e-4 of the
 * float64 statement of the same pipeline (oracle/melspec.py) on the normalised scale. */
typedef struct {
    int32_t sample_rate, n_fft, hop_length, win_length, num_mels;
    float fmin, min_level_db;
} wrnn_mel_params;
int wrnn_melspectrogram(wrnn_handle *h, const float *wav, int64_t n, const wrnn_mel_params *mp, float *mel_out,
                        int64_t frames_cap, int64_t *frames_out, void *cuda_stream);

/* ------------------------------------------------------------------ the path
 * wrnn_generate == WaveRNN.generate for `n_utt` independent utterances (the reference takes one;
 * fatchord_version.py:169-264).  mels[u] is (feat_dims, T[u]) row-major float32.  out[u] must hold
 * (T[u]-1)*hop_length float64; out_len[u] receives that length.  Host buffers; the H2D/D2H
 * copies are part of the call.  `cuda_stream` is a cudaStream_t (NULL = default stream); the call
 * returns after the result is on the host.  Folds of all utterances run as one fold batch. */
int wrnn_generate(wrnn_handle *h, const float *const *mels, const int32_t *T, int32_t n_utt,
                  int32_t batched, int32_t target, int32_t overlap, int32_t mu_law,
                  const wrnn_noise *noise, uint64_t seed,
                  double *const *out, int64_t *out_len, void *cuda_stream);

/* Same work split in three so a benchmark can time the device part with inputs resident in HBM:
 * stage (H2D of mels) -> run (all kernels, asynchronous on the stream) -> fetch (D2H of waves). */
int wrnn_stage(wrnn_handle *h, const float *const *mels, const int32_t *T, int32_t n_utt,
               int32_t batched, int32_t target, int32_t overlap, int32_t mu_law,
               const wrnn_noise *noise, void *cuda_stream);
int wrnn_run_staged(wrnn_handle *h, uint64_t seed, void *cuda_stream);
int wrnn_fetch(wrnn_handle *h, double *const *out, int64_t *out_len, void *cuda_stream);
/* Device time of the last run's phases in ms (CUDA events on the launch stream): conditioning
 * (+fold), per-sample loop, output tail; plus loop extent.  Valid after the stream is idle. */
int wrnn_last_timing(wrnn_handle *h, float *ms_cond, float *ms_loop, float *ms_tail,
                     int64_t *steps, int64_t *folds, int64_t *kernel_launches);
/* Raw per-fold samples of the last run: (folds, steps) float32 (RAW mode: 2k/(C-1)-1). */
int wrnn_fetch_samples(wrnn_handle *h, float *samples, void *cuda_stream);
/* Fold-range sharding of ONE long utterance over several GPUs (folds never communicate: cold-start state per fold,
 * fatchord_version.py:194-196).  After wrnn_set_fold_range(h, lo, hi) the next batched
 * single-utterance wrnn_stage / wrnn_run_staged computes conditioning only for the frames folds [lo, hi) cover and runs
 * the loop for those folds only; injected noise and Philox stay addressed by the GLOBAL fold index, so the samples equal
 * those of an unsharded call.  The result is read with wrnn_fetch_samples ((hi - lo, steps) float32); wrnn_fetch fails,
 * because the float64 cross-fade / unfold / fade-out (:243-258) needs every shard: the gathering host does it
 * (wavernn_b200/shard.py: combine_fold_samples).  lo = hi = 0 switches the range off. */
int wrnn_set_fold_range(wrnn_handle *h, int64_t lo, int64_t hi);

/* ------------------------------------------------------------------ components (parity tests) */
/* UpsampleNetwork.forward on the zero-padded mel (fatchord_version.py:82-89,185-186).
 * mel (feat_dims, T) -> mels_out (T*hop, feat_dims), aux_out (T*hop, res_out_dims), time-major. */
int wrnn_upsample(wrnn_handle *h, const float *mel, int32_t T, float *mels_out, float *aux_out);
/* fold_with_overlap (fatchord_version.py:293-340): x (L, F) -> out (folds, target+2*overlap, F).
 * wrnn_num_folds gives the fold count for sizing `out`. */
int64_t wrnn_num_folds(int64_t total_len, int64_t target, int64_t overlap);
int wrnn_fold(wrnn_handle *h, const float *x, int64_t L, int32_t F, int64_t target, int64_t overlap,
              float *out);
/* xfade_and_unfold (fatchord_version.py:342-405): y (folds, target+2*overlap) float64 ->
 * out (folds*(target+overlap)+overlap) float64. */
int wrnn_unfold(wrnn_handle *h, const double *y, int64_t folds, int64_t target, int64_t overlap,
                double *out);
/* wrnn_unfold without any allocation.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Same
 * arguments, arithmetic and result as wrnn_unfold (xfade_and_unfold, fatchord_version.py:342-405); the device scratch
 * (device copy of y, fade tables, device output) is a file-scope static __device__ array of WRNN_NOALLOC_MAX_DOUBLES
 * doubles (4 MiB, held by every GPU that loads the library until exit), so the call performs no cudaMalloc / cudaFree /
 * host allocation.  Needs folds*(target+2*overlap) + folds*(target+overlap)+overlap + 2*overlap <= WRNN_NOALLOC_MAX_DOUBLES
 * and overlap <= WRNN_NOALLOC_MAX_OVERLAP, else WRNN_ERR_ARG with a message.  All calls on a device share the array, which
 * keeps the last call's contents (nothing in it must start at zero); calls that overlap on the GPU race: serialise them.
 * `cuda_stream` as in wrnn_generate; returns after the result is on the host. */
#define WRNN_NOALLOC_MAX_DOUBLES 524288
#define WRNN_NOALLOC_MAX_OVERLAP 4096
int wrnn_unfold_noalloc(wrnn_handle *h, const double *y, int64_t folds, int64_t target, int64_t overlap, double *out,
                        void *cuda_stream);
/* decode_mu_law(y, mu, from_labels=False) (utils/dsp.py:98-103) for samples y = 2k/(mu-1)-1 given
 * as float32; out float64.  wrnn_mulaw_lut (host only, no GPU) fills the mu-entry table. */
int wrnn_decode_mu_law(wrnn_handle *h, const float *y, int64_t n, int32_t mu, double *out);
int wrnn_mulaw_lut(int32_t mu, double *lut);
/* numpy's float64 `power` is a SIMD routine that differs from libm `pow` by 1 ulp on some inputs
 * (and between CPUs), so a host that must match the reference bit-for-bit evaluates the reference's
 * expression itself and installs the table here (mu entries, copied).  Without it the library uses
 * wrnn_mulaw_lut (libm, within 1 ulp). */
int wrnn_set_mulaw_lut(wrnn_handle *h, int32_t mu, const double *lut);
/* Host tables of the cross-fade / fade-out gains (numpy linspace semantics), host only. */
int wrnn_xfade_tables(int64_t overlap, double *fade_in, double *fade_out);
/* Teacher-forced logits == WaveRNN.forward (fatchord_version.py:131-167): x (n_seq, S) previous
 * samples, mels (n_seq, feat_dims, S/hop + 2*pad) ALREADY padded -> logits (n_seq, S, n_classes). */
int wrnn_forward_logits(wrnn_handle *h, const float *x, const float *mels, int32_t n_seq, int64_t S,
                        float *logits);
/* Which loop kernel the next run will use, for reports: "fma", "fma-pipelined", "fma-resident".
 * "fma-resident" (register-resident weights, TMA exchanges; csrc/loop_res.cu) takes rnn_dims = fc_dims = 512,
 * aux 32, feat_dims a multiple of 8 up to 96, MoL or RAW with 512 classes, on a GPU with >= 129 SMs. */
const char *wrnn_loop_kernel_name(wrnn_handle *h, int64_t folds);
/* Force a loop kernel: 0 = auto (fastest kernel that supports the dims and the fold count), 1 = fma (grid-barrier
 * kernel, any dims and fold count, reference order of operations), 2 = fma-pipelined (tagged exchange, shared-memory
 * weights), 3 = fma-resident (register-resident weights, groups of <= 20 folds, whatever the fold count),
 * 4 = "tcgen05-bf16x3" (csrc/loop_mma.cu: the contractions of fatchord_version.py:208-221 on the tensor
 * cores with bf16 hi/lo split operands and fp32 TMEM accumulators; default dims, MoL; auto from 41 folds up). */
int wrnn_set_loop_kernel(wrnn_handle *h, int32_t which);
/* What the register-resident kernel measured at its first run on this handle: the L2 round trip in SM clocks (a
 * dependent `ld.global.cg` chain) and the factor by which it scaled the delays before the first look at an exchange
 * (tuned at 1965 MHz on a B200).  Zero / one before the first run.  WRNN_NO_CALIBRATION=1 pins the scale to 1. */
int wrnn_exchange_calibration(wrnn_handle *h, float *l2_round_trip_clk, float *delay_scale);
/* Kept for callers of earlier builds: the tensor-core kernel's fold tile is the M of its MMAs (128 folds = the 128 TMEM
 * lanes) and no longer a choice.  The argument is validated (0, or a multiple of 16 <= 128) and otherwise ignored. */
int wrnn_set_mma_fold_tile(wrnn_handle *h, int32_t nt);
/* Bring-up aid: recurrent state of the last tensor-core run, (folds, 512) float32: which = 0 h1, 1 h2, 2 fc2 output.
 * With the environment variable WRNN_MMA_STEPS=n the kernel stops after n loop steps so the state can be compared
 * with the oracle's after the same steps. */
int wrnn_debug_mma_state(wrnn_handle *h, int32_t which, int64_t folds, float *out);

/* ------------------------------------------------------------------ deepmind variant */
typedef struct wrnn_dm_handle wrnn_dm_handle;
int wrnn_dm_create(int32_t hidden_size, int32_t quantisation, int device, wrnn_dm_handle **out);
void wrnn_dm_destroy(wrnn_dm_handle *h);
const char *wrnn_dm_last_error(const wrnn_dm_handle *h);
int wrnn_dm_load_weights(wrnn_dm_handle *h, const wrnn_tensor *tensors, int32_t n);
/* deepmind_version.WaveRNN.generate (models/deepmind_version.py:75-165) for n_seq independent
 * sequences (the reference runs one).  q_coarse/q_fine: optional Exp(1) noise
 * [seq_len][n_seq][quantisation]; NULL => Philox(seed).  Outputs int64 (n_seq, seq_len):
 * output = coarse*256 + fine - 32768 (utils/dsp.py:33-34). */
int wrnn_dm_generate(wrnn_dm_handle *h, int64_t seq_len, int32_t n_seq, const float *q_coarse,
                     const float *q_fine, uint64_t seed, int64_t *output, int64_t *coarse,
                     int64_t *fine, void *cuda_stream);
int wrnn_dm_last_timing(wrnn_dm_handle *h, float *ms_loop, int64_t *kernel_launches);

/* ------------------------------------------------------------------ Tacotron decoder loop */
typedef struct wrnn_taco_handle wrnn_taco_handle;
/* decoder_dims must be 256: PreNet's widths are fixed (models/tacotron.py:137) and the attention GRU's input is
 * decoder_dims + decoder_dims // 2 (:218).  lstm_dims <= 512 (hparams.py:72), n_mels <= 128. */
int wrnn_taco_create(int32_t decoder_dims, int32_t lstm_dims, int32_t n_mels, int device, wrnn_taco_handle **out);
void wrnn_taco_destroy(wrnn_taco_handle *h);
const char *wrnn_taco_last_error(const wrnn_taco_handle *h);
/* Tensors of `Tacotron.state_dict()` by checkpoint key: every `decoder.*` parameter (models/tacotron.py:211-224) and
 * `encoder_proj.weight` (:301); other keys are kept and ignored. */
int wrnn_taco_load_weights(wrnn_taco_handle *h, const wrnn_tensor *tensors, int32_t n);
/* The decoder loop of Tacotron.generate (models/tacotron.py:438-465) for one utterance: enc_seq (L, decoder_dims) is
 * the CBHG encoder's output (host memory), 1 <= L <= 512 characters (the attention weights of a step travel one word per
 * thread; longer texts return WRNN_ERR_ARG, and so does a text whose per-position tables do not fit the shared memory
 * beside the weights: about 500 characters at r = 7, fewer at larger r), r the reduction factor (`model.r`), steps the frame budget
 * (`for t in range(0, steps, r)`), stop_threshold the model's buffer of that name (a step whose r frames are all below
 * it ends the loop once t > 10, :455).  Outputs (host memory): mel (n_mels, n_iter * r) row-major, capacity
 * n_mels * ceil(steps / r) * r floats; attn (n_iter, L), capacity ceil(steps / r) * L floats; *n_iter = decoder steps
 * run.  The postnet and the linear projection after the loop (:458-459) stay with the caller. */
int wrnn_taco_decode(wrnn_taco_handle *h, const float *enc_seq, int32_t L, int32_t r, int32_t steps, float stop_threshold,
                     float *mel, float *attn, int32_t *n_iter, void *cuda_stream);
int wrnn_taco_last_timing(wrnn_taco_handle *h, float *ms_loop, int64_t *kernel_launches);

#ifdef __cplusplus
}
#endif
#endif /* WAVERNN_B200_H */

// melspectrogram (utils/dsp.py:72-75): the step in front of the path when `gen_from_file` is given a .wav
// (gen_wavernn.py:44-47: `mel = melspectrogram(wav)`), which the reference delegates to librosa 0.6.3:
//
//   D = librosa.stft(y, n_fft, hop_length, win_length)         centre = True: reflect-pad n_fft / 2 on both sides, periodic
//                                                              Hann window of win_length zero-padded to n_fft (centred),
//                                                              frames 1 + len(y) / hop, bins 0 .. n_fft / 2
//   S = amp_to_db(linear_to_mel(|D|))                          mel basis: librosa.filters.mel(sr, n_fft, n_mels, fmin),
//                                                              Slaney scale, area-normalised; 20 log10(max(1e-5, .))
//   normalize(S) = clip((S - min_level_db) / -min_level_db, 0, 1)
//
// Two kernels.  `stft_mag_kernel`: one CTA per frame; the windowed frame (only win_length samples are non-zero) and a
// table of the n_fft roots of unity sit in shared memory (double), every thread forms whole bins by a direct DFT with two
// double accumulator pairs (1 100 x 1 025 complex MACs per frame: 3.6 G DFMA for a 10 s utterance, well under a
// millisecond of B200 fp64; an FFT would save arithmetic nobody is waiting for).  `mel_db_kernel`: the 80 x 1 025 projection, dB, normalisation, written [n_mels][T]
// as the model wants it.  The basis and the window are formed on the host in double, once per parameter set.
#include <math.h>
#include <string.h>

#include <vector>

#include "internal.h"

namespace wrnn {

__global__ void __launch_bounds__(256) stft_mag_kernel(const float *__restrict__ y, int64_t n, int n_fft, int hop, int win, const double *__restrict__ window,
                                                       const double2 *__restrict__ roots, float *__restrict__ mag, int bins) {
    extern __shared__ double smd[];
    double *frame = smd;                                        // [win] windowed samples
    double2 *tw = reinterpret_cast<double2 *>(smd + ((win + 1) & ~1));   // [n_fft] (cos, -sin)(2 pi j / n_fft)
    const int64_t t = blockIdx.x;
    const int off = (n_fft - win) / 2;                          // where the window sits in the n_fft frame
    for (int i = threadIdx.x; i < win; i += blockDim.x) {
        int64_t p = t * hop + off + i - n_fft / 2;              // index into y before padding
        if (p < 0) p = -p;                                      // numpy 'reflect': no repeat of the edge sample
        if (p >= n) p = 2 * (n - 1) - p;
        p = p < 0 ? 0 : (p >= n ? n - 1 : p);                   // (signals shorter than n_fft / 2 are rejected on the host)
        frame[i] = (double)y[p] * window[i];
    }
    for (int i = threadIdx.x; i < n_fft; i += blockDim.x) tw[i] = roots[i];
    __syncthreads();
    const int mask = n_fft - 1;                                  // n_fft is a power of two (checked on the host)
    // Double accumulation and double roots: bins far from a strong partial are sums that cancel to 1e-5 of their terms,
    // and the reference (librosa: float64 FFT, then a cast) resolves them; fp32 left 1.5e-4 of the normalised scale there.
    for (int k = threadIdx.x; k < bins; k += blockDim.x) {
        double re0 = 0.0, im0 = 0.0, re1 = 0.0, im1 = 0.0;
        int ph = (k * off) & mask;                               // k * (off + i) mod n_fft, advanced by k per sample
        int i = 0;
        for (; i + 1 < win; i += 2) {
            const double2 a = tw[ph];
            ph = (ph + k) & mask;
            const double2 b = tw[ph];
            ph = (ph + k) & mask;
            const double x0 = frame[i], x1 = frame[i + 1];
            re0 = fma(x0, a.x, re0);
            im0 = fma(x0, a.y, im0);
            re1 = fma(x1, b.x, re1);
            im1 = fma(x1, b.y, im1);
        }
        if (i < win) {
            const double2 a = tw[ph];
            re0 = fma(frame[i], a.x, re0);
            im0 = fma(frame[i], a.y, im0);
        }
        const double re = re0 + re1, im = im0 + im1;
        mag[(size_t)t * bins + k] = (float)sqrt(re * re + im * im);
    }
}

__global__ void mel_db_kernel(const float *__restrict__ mag, const float *__restrict__ basis, int64_t T, int bins, int n_mels,
                              float min_level_db, float *__restrict__ out) {
    // one warp per (frame, mel band)
    const int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (w >= T * n_mels) return;
    const int64_t t = w / n_mels;
    const int m = (int)(w - t * n_mels);
    float acc = 0.f;
    for (int k = lane; k < bins; k += 32) acc = fmaf(basis[(size_t)m * bins + k], mag[(size_t)t * bins + k], acc);
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) {
        const float db = 20.f * log10f(fmaxf(1e-5f, acc));
        out[(size_t)m * T + t] = fminf(fmaxf((db - min_level_db) / -min_level_db, 0.f), 1.f);
    }
}

// librosa.filters.mel (0.6.3: htk = False, norm = 1), in double
static double hz_to_mel(double f) {
    const double f_sp = 200.0 / 3, min_log_hz = 1000.0, min_log_mel = min_log_hz / f_sp, logstep = log(6.4) / 27.0;
    return f >= min_log_hz ? min_log_mel + log(f / min_log_hz) / logstep : f / f_sp;
}
static double mel_to_hz(double m) {
    const double f_sp = 200.0 / 3, min_log_hz = 1000.0, min_log_mel = min_log_hz / f_sp, logstep = log(6.4) / 27.0;
    return m >= min_log_mel ? min_log_hz * exp(logstep * (m - min_log_mel)) : f_sp * m;
}
static void mel_basis(int sr, int n_fft, int n_mels, double f_min, std::vector<float> &out) {
    const int bins = 1 + n_fft / 2;
    const double f_max = sr / 2.0, lo = hz_to_mel(f_min), hi = hz_to_mel(f_max);
    std::vector<double> mel_f(n_mels + 2);
    for (int i = 0; i < n_mels + 2; i++) mel_f[i] = mel_to_hz(lo + (hi - lo) * i / (n_mels + 1));      // np.linspace
    mel_f[n_mels + 1] = mel_to_hz(hi);
    out.assign((size_t)n_mels * bins, 0.f);
    for (int i = 0; i < n_mels; i++) {
        const double enorm = 2.0 / (mel_f[i + 2] - mel_f[i]);
        for (int k = 0; k < bins; k++) {
            const double f = (double)k * sr / n_fft;                                                   // fft_frequencies
            const double lower = (f - mel_f[i]) / (mel_f[i + 1] - mel_f[i]), upper = (mel_f[i + 2] - f) / (mel_f[i + 2] - mel_f[i + 1]);
            const double wv = lower < upper ? lower : upper;
            out[(size_t)i * bins + k] = (float)((wv > 0 ? wv : 0.0) * enorm);
        }
    }
}

}  // namespace wrnn

using namespace wrnn;

extern "C" int wrnn_melspectrogram(wrnn_handle *h, const float *wav, int64_t n, const wrnn_mel_params *mp, float *mel_out,
                                   int64_t frames_cap, int64_t *frames_out, void *cuda_stream) {
    if (!h || !wav || !mp || !frames_out) return fail(h, WRNN_ERR_ARG, "bad argument");
    const int n_fft = mp->n_fft, hop = mp->hop_length, win = mp->win_length, n_mels = mp->num_mels, bins = 1 + n_fft / 2;
    if (n_fft < 16 || (n_fft & (n_fft - 1)) || n_fft > 8192 || win < 1 || win > n_fft || hop < 1 || n_mels < 1 || n_mels > 1024 ||
        mp->sample_rate < 1 || !(mp->min_level_db < 0))
        return fail(h, WRNN_ERR_ARG, "melspectrogram: n_fft must be a power of two in [16, 8192], 1 <= win_length <= n_fft, hop >= 1, min_level_db < 0");
    if (n <= n_fft / 2) return fail(h, WRNN_ERR_ARG, "melspectrogram: %lld samples, reflect padding needs more than n_fft / 2 = %d", (long long)n, n_fft / 2);
    const int64_t T = 1 + n / hop;
    *frames_out = T;
    if (!mel_out) return 0;                                     // size query
    if (frames_cap < T) return fail(h, WRNN_ERR_ARG, "melspectrogram: output holds %lld frames, %lld needed", (long long)frames_cap, (long long)T);
    WRNN_DEVICE(h);
    cudaStream_t st = (cudaStream_t)cuda_stream;
    // tables: window [win] double | roots [n_fft] double2 | basis [n_mels][bins] float  (host in double, rebuilt when the parameters change)
    if (win & 1) return fail(h, WRNN_ERR_ARG, "melspectrogram: win_length must be even");
    const size_t tab_doubles = (size_t)win + 2 * (size_t)n_fft, tab_bytes = tab_doubles * sizeof(double) + (size_t)n_mels * bins * sizeof(float);
    if (!h->mel_tab.p || memcmp(&h->mel_tab_params, mp, sizeof(*mp)) != 0) {
        std::vector<double> tab(tab_doubles);
        const double two_pi = 6.283185307179586476925286766559;
        for (int i = 0; i < win; i++) tab[i] = 0.5 - 0.5 * cos(two_pi * i / win);                      // scipy get_window('hann', fftbins=True)
        for (int j = 0; j < n_fft; j++) {
            tab[win + 2 * j] = cos(two_pi * j / n_fft);
            tab[win + 2 * j + 1] = -sin(two_pi * j / n_fft);
        }
        std::vector<float> basis;
        mel_basis(mp->sample_rate, n_fft, n_mels, mp->fmin, basis);
        WRNN_TRY(ws_reserve(h, h->mel_tab, tab_bytes + 16));
        WRNN_CUDA(h, cudaMemcpyAsync(h->mel_tab.p, tab.data(), tab_doubles * sizeof(double), cudaMemcpyHostToDevice, st));
        WRNN_CUDA(h, cudaMemcpyAsync((char *)h->mel_tab.p + tab_doubles * sizeof(double), basis.data(), basis.size() * sizeof(float),
                                     cudaMemcpyHostToDevice, st));
        WRNN_CUDA(h, cudaStreamSynchronize(st));                // the host tables go out of scope
        h->mel_tab_params = *mp;
    }
    const double *d_window = (const double *)h->mel_tab.p;
    const double2 *d_roots = (const double2 *)(d_window + win);
    const float *d_basis = (const float *)(d_window + tab_doubles);
    WRNN_TRY(ws_reserve(h, h->ws_tmp0, (size_t)n * sizeof(float)));
    WRNN_TRY(ws_reserve(h, h->ws_tmp1, ((size_t)T * bins + (size_t)T * n_mels) * sizeof(float)));
    float *d_y = (float *)h->ws_tmp0.p, *d_mag = (float *)h->ws_tmp1.p, *d_out = d_mag + (size_t)T * bins;
    WRNN_CUDA(h, cudaMemcpyAsync(d_y, wav, (size_t)n * sizeof(float), cudaMemcpyHostToDevice, st));
    const size_t smem = (size_t)((win + 1) & ~1) * sizeof(double) + (size_t)n_fft * sizeof(double2);
    if (smem > 48 * 1024) WRNN_CUDA(h, cudaFuncSetAttribute(stft_mag_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    stft_mag_kernel<<<(unsigned)T, 256, smem, st>>>(d_y, n, n_fft, hop, win, d_window, d_roots, d_mag, bins);
    mel_db_kernel<<<(unsigned)((T * n_mels * 32 + 255) / 256), 256, 0, st>>>(d_mag, d_basis, T, bins, n_mels, mp->min_level_db, d_out);
    h->launches += 2;
    WRNN_CUDA(h, cudaGetLastError());
    WRNN_CUDA(h, cudaMemcpyAsync(mel_out, d_out, (size_t)T * n_mels * sizeof(float), cudaMemcpyDeviceToHost, st));
    WRNN_CUDA(h, cudaStreamSynchronize(st));
    return 0;
}

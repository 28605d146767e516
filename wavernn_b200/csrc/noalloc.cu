// wrnn_unfold_noalloc: xfade_and_unfold (fatchord_version.py:342-405) for callers that cannot call cudaMalloc.
//
// NO COUNTERPART IN THE REFERENCE: added for callers under CUDA graph capture,
// tight serving loops and harnesses that forbid allocation.  It is a variant of `wrnn_unfold` (api_run.cu), whose device
// scratch (the device copy of the folds, the two fade tables and the device output) comes from `ws_reserve` = cudaMalloc;
// here that scratch is ONE file-scope `static __device__` array of a fixed, modest size, so the call allocates nothing:
// no cudaMalloc / cudaFree / cudaMallocHost, no host heap, no allocating library routine.  Every GPU that loads the
// library holds the array (4 MiB) until the process exits.
//
// Limits: folds * (target + 2 * overlap)  +  folds * (target + overlap) + overlap  +  2 * overlap  <=  NOALLOC_MAX_DOUBLES
// (524 288 doubles: BASELINE config 2, 20 folds x 12 100 samples -> 231 550 samples, fits), overlap <= NOALLOC_MAX_OVERLAP.
// Larger calls are rejected with WRNN_ERR_ARG and a message; use wrnn_unfold for them.
//
// Sharing: every call on a device uses the same array, and the array keeps the last call's contents.  Nothing in it has
// to start at zero (every output element is written by the kernel, every input element by the copies in front of it),
// but two calls that overlap ON THE GPU (different streams, or different host threads on one device) race on it: callers
// serialise them.  The call itself synchronises its stream before it returns, so back-to-back calls from one thread are safe.
#include <math.h>
#include <stdio.h>

#include <mutex>

#include "internal.h"

namespace wrnn {

constexpr int64_t NOALLOC_MAX_DOUBLES = WRNN_NOALLOC_MAX_DOUBLES;      // 4 MiB of scratch per device
constexpr int64_t NOALLOC_MAX_OVERLAP = WRNN_NOALLOC_MAX_OVERLAP;
constexpr int NOALLOC_MAX_DEVICES = 64;

// The scratch.  In HOST code this name is a host-side stand-in, not a device address: host code below only ever hands it
// to cudaGetSymbolAddress; the kernel receives the address that call returns.
static __device__ __align__(256) double g_noalloc_scratch[NOALLOC_MAX_DOUBLES];

static double *g_scratch_addr[NOALLOC_MAX_DEVICES] = {nullptr};
static std::mutex g_scratch_mutex;

// Once per device, with that device current: the device address of the scratch array.
static cudaError_t scratch_for_device(int device, double **out) {
    std::lock_guard<std::mutex> lock(g_scratch_mutex);
    if (!g_scratch_addr[device]) {
        double *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void **)&p, g_noalloc_scratch);
        if (e != cudaSuccess) return e;
        g_scratch_addr[device] = p;
    }
    *out = g_scratch_addr[device];
    return cudaSuccess;
}

// Same arithmetic as unfold_kernel (cond.cu): at most two addends per output sample, products and sums rounded one by
// one (__dmul_rn / __dadd_rn, never an FMA), so the result equals numpy's bit for bit.  `s` = [y | fade_in | fade_out | out].
__global__ void unfold_noalloc_kernel(double *__restrict__ s, int64_t folds, int64_t target, int64_t overlap) {
    const int64_t stride = target + overlap, len = target + 2 * overlap, total = folds * stride + overlap;
    const double *y = s, *fade_in = s + folds * len, *fade_out = fade_in + overlap;
    double *out = s + folds * len + 2 * overlap;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const int64_t b = i / stride, pos = i - b * stride;
    auto gain = [&](double v, int64_t p) {
        if (p < overlap) v = __dmul_rn(v, fade_in[p]);                          // y[:, :overlap] *= fade_in
        if (p >= len - overlap) v = __dmul_rn(v, fade_out[p - (len - overlap)]);    // y[:, -overlap:] *= fade_out
        return v;
    };
    double v = 0.0;
    if (b >= 1 && pos < overlap) v = __dadd_rn(v, gain(y[(size_t)(b - 1) * len + pos + stride], pos + stride));
    if (b < folds) v = __dadd_rn(v, gain(y[(size_t)b * len + pos], pos));
    out[i] = v;
}

}  // namespace wrnn

using namespace wrnn;

extern "C" int wrnn_unfold_noalloc(wrnn_handle *h, const double *y, int64_t folds, int64_t target, int64_t overlap, double *out,
                                   void *cuda_stream) {
    if (!h || !y || !out || folds < 1 || target < 0 || overlap < 0) return fail(h, WRNN_ERR_ARG, "bad argument");
    const int64_t len = target + 2 * overlap, total = folds * (target + overlap) + overlap;
    if (overlap > NOALLOC_MAX_OVERLAP || len > NOALLOC_MAX_DOUBLES || folds > NOALLOC_MAX_DOUBLES ||
        folds * len + total + 2 * overlap > NOALLOC_MAX_DOUBLES)
        return fail(h, WRNN_ERR_ARG,
                    "wrnn_unfold_noalloc: %lld folds x %lld samples (overlap %lld) need %lld doubles of scratch, the static "
                    "array holds %lld (overlap <= %lld): use wrnn_unfold",
                    (long long)folds, (long long)len, (long long)overlap, (long long)(folds * len + total + 2 * overlap),
                    (long long)NOALLOC_MAX_DOUBLES, (long long)NOALLOC_MAX_OVERLAP);
    if (h->device < 0 || h->device >= NOALLOC_MAX_DEVICES) return fail(h, WRNN_ERR_ARG, "device index out of range");
    WRNN_DEVICE(h);
    cudaStream_t st = (cudaStream_t)cuda_stream;
    double *scratch = nullptr;
    WRNN_CUDA(h, scratch_for_device(h->device, &scratch));
    // fade tables (fatchord_version.py:379-391), numpy semantics, on the stack: no heap
    double tab[2 * NOALLOC_MAX_OVERLAP];
    {
        const int64_t silence = overlap / 2, fade = overlap - silence;
        double *fade_in = tab, *fade_out = tab + overlap;
        for (int64_t i = 0; i < silence; i++) { fade_in[i] = 0.0; fade_out[i] = 1.0; }
        double t_lin[NOALLOC_MAX_OVERLAP];
        host_linspace(-1.0, 1.0, fade, t_lin);                                   // api.cu: numpy.linspace semantics, no allocation
        for (int64_t i = 0; i < fade; i++) {
            const double t = t_lin[i];
            volatile double a = 1 + t, b = 1 - t;
            volatile double ha = 0.5 * a, hb = 0.5 * b;
            fade_in[silence + i] = sqrt(ha);
            fade_out[silence + i] = sqrt(hb);
        }
    }
    WRNN_CUDA(h, cudaMemcpyAsync(scratch, y, (size_t)(folds * len) * sizeof(double), cudaMemcpyHostToDevice, st));
    if (overlap > 0)
        WRNN_CUDA(h, cudaMemcpyAsync(scratch + folds * len, tab, (size_t)(2 * overlap) * sizeof(double), cudaMemcpyHostToDevice, st));
    // the table copy reads a stack buffer: it must have left the host before this frame goes away (it has: the stream is
    // synchronised below, before any return that follows a successful enqueue)
    unfold_noalloc_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(scratch, folds, target, overlap);
    h->launches++;
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess)
        e = cudaMemcpyAsync(out, scratch + folds * len + 2 * overlap, (size_t)total * sizeof(double), cudaMemcpyDeviceToHost, st);
    const cudaError_t es = cudaStreamSynchronize(st);
    if (e != cudaSuccess || es != cudaSuccess)
        return fail(h, WRNN_ERR_CUDA, "wrnn_unfold_noalloc: %s", cudaGetErrorString(e != cudaSuccess ? e : es));
    return 0;
}

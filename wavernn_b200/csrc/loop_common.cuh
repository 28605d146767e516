// Device helpers shared by the per-sample loop kernels: grid barrier, Philox, samplers, gates.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace wrnn {

// ---------------------------------------------------------------------------- grid barrier
// Monotonic-counter barrier between the CTAs of one cooperative launch (all co-resident).
// `gen` counts barriers passed by this CTA; the counter reaches gen*gridDim.x when all arrived.
// A watchdog turns a would-be hang into an error flag the host reports (a hung GPU box is a strike).
struct GridSync {
    unsigned long long *counter;
    int *abort_flag;
};

__device__ __forceinline__ unsigned long long ld_acquire_u64(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_add_u64(unsigned long long *p, unsigned long long v) {
    asm volatile("red.release.gpu.global.add.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ int ld_volatile_s32(const int *p) {
    int v;
    asm volatile("ld.volatile.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ void grid_barrier(const GridSync &gs, unsigned long long &gen) {
    __syncthreads();
    gen++;
    if (threadIdx.x == 0) {
        red_release_add_u64(gs.counter, 1ull);
        const unsigned long long want = gen * gridDim.x;
        long long t0 = clock64();
        unsigned spins = 0;
        while (ld_acquire_u64(gs.counter) < want) {
            // the watchdog and the abort flag every 64 polls only: a second global load per poll doubled the time to see
            // the last arrival
            if ((++spins & 63u) == 0u && (clock64() - t0 > 4000000000ll || ld_volatile_s32(gs.abort_flag))) {   // ~2 s
                atomicExch(gs.abort_flag, 1);
                break;
            }
        }
    }
    __syncthreads();
}

// ---------------------------------------------------------------------------- Philox4x32-10
__device__ __forceinline__ uint4 philox4x32(uint4 ctr, uint2 key) {
#pragma unroll
    for (int i = 0; i < 10; i++) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, ctr.x), lo0 = 0xD2511F53u * ctr.x;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, ctr.z), lo1 = 0xCD9E8D57u * ctr.z;
        ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
        key.x += 0x9E3779B9u;
        key.y += 0xBB67AE85u;
    }
    return ctr;
}
__device__ __forceinline__ float u01(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }   // [0,1)
// uniform_(1e-5, 1 - 1e-5) (utils/distribution.py:106,118)
__device__ __forceinline__ float u_clip(uint32_t x) { return u01(x) * ((1.0f - 1e-5f) - 1e-5f) + 1e-5f; }
// Exp(1): -log(u), u in (0,1]
__device__ __forceinline__ float exp1(uint32_t x) { return -logf(1.0f - u01(x)); }

// ---------------------------------------------------------------------------- math
__device__ __forceinline__ float sigmoidf_acc(float x) { return 1.0f / (1.0f + expf(-x)); }

// GRUCell in ATen's CPU evaluation order (aten/src/ATen/native/RNN.cpp, what nn.GRUCell at
// fatchord_version.py:210,214 runs): ir/hr etc. already include their biases.
__device__ __forceinline__ float gru_update(float ir, float iz, float in_, float hr, float hz, float hn, float h) {
    const float r = sigmoidf_acc(hr + ir);
    const float z = sigmoidf_acc(hz + iz);
    const float n = tanhf(in_ + hn * r);
    return (h - n) * z + n;
}

// The same cell with the special-function unit (ex2.approx through __expf, rcp through __fdividef): |error| ~ 2e-7
// per gate, below the 3.6e-7 the tagged exchange words already carry.  For the resident kernel, where the 80 gate
// threads of a CTA sit on the sample-to-sample chain twice per step; the exact-order kernels keep gru_update.
__device__ __forceinline__ float sigmoidf_fast(float x) { return __fdividef(1.0f, 1.0f + __expf(-x)); }
__device__ __forceinline__ float tanhf_fast(float x) {
    const float e = __expf(-2.0f * fabsf(x));                 // in (0, 1]: no overflow
    return copysignf(__fdividef(1.0f - e, 1.0f + e), x);
}
__device__ __forceinline__ float gru_update_fast(float ir, float iz, float in_, float hr, float hz, float hn, float h) {
    const float r = sigmoidf_fast(hr + ir);
    const float z = sigmoidf_fast(hz + iz);
    const float n = tanhf_fast(in_ + hn * r);
    return (h - n) * z + n;
}

// ---------------------------------------------------------------------------- samplers (one warp per fold)
// sample_from_discretized_mix_logistic (utils/distribution.py:87-123) for one fold.  `logits` (3*nmix,
// possibly written by another CTA: .cg loads), lane < nmix holds its own u1.  Returns the sample on
// every lane; *pick = mixture index.
__device__ __forceinline__ float mol_sample_warp(const float *logits, int nmix, float myu1, float u2,
                                                 int lane, int *pick) {
    float t = -INFINITY;
    if (lane < nmix) t = __ldcg(logits + lane) - logf(-logf(myu1));
    int idx = lane;
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) {       // argmax, lowest index on ties (torch.max semantics)
        const float ot = __shfl_xor_sync(0xffffffffu, t, o);
        const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
        if (ot > t || (ot == t && oi < idx)) { t = ot; idx = oi; }
    }
    const float mean = __ldcg(logits + nmix + idx);
    const float ls = fmaxf(__ldcg(logits + 2 * nmix + idx), -32.23619130191664f);   // log(1e-14)
    float x = mean + expf(ls) * (logf(u2) - logf(1.f - u2));
    x = fminf(fmaxf(x, -1.f), 1.f);
    *pick = idx;
    return x;
}

// RAW branch (fatchord_version.py:231-237): softmax -> Categorical(probs).sample() == argmax(p/q),
// q ~ Exp(1) -> 2k/(C-1)-1.  qsrc: injected noise row or null (Philox).  Returns class index.
__device__ __forceinline__ int categorical_warp(const float *logits, int C, const float *qsrc, uint2 key,
                                                uint32_t c0, uint32_t c1, uint32_t stream, int lane) {
    float m = -INFINITY;
    for (int i = lane; i < C; i += 32) m = fmaxf(m, logits[i]);
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    float s = 0.f;
    for (int i = lane; i < C; i += 32) s += expf(logits[i] - m);
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    float ps = 0.f;   // Categorical renormalises: probs / probs.sum(-1)
    for (int i = lane; i < C; i += 32) ps += expf(logits[i] - m) / s;
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) ps += __shfl_xor_sync(0xffffffffu, ps, o);
    float best = -INFINITY;
    int bi = 0x7fffffff;
    for (int i0 = 0; i0 < C; i0 += 128) {          // 4 draws per lane per Philox call
        uint4 r = make_uint4(0, 0, 0, 0);
        if (!qsrc) r = philox4x32(make_uint4(c0, c1, (uint32_t)(i0 / 128 * 32 + lane), stream), key);
        const uint32_t rr[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int i = i0 + j * 32 + lane;
            if (i < C) {
                const float p = (expf(logits[i] - m) / s) / ps;
                const float q = qsrc ? qsrc[i] : exp1(rr[j]);
                const float v = p / q;
                if (v > best) { best = v; bi = i; }    // ascending i per lane: first max kept
            }
        }
    }
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
    }
    return bi;
}
__device__ __forceinline__ float raw_value(int k, int C) {   // 2 * k.float() / (n_classes - 1.) - 1.
    return __fsub_rn(__fdiv_rn(2.f * (float)k, (float)C - 1.f), 1.f);
}

// Tree reduce-scatter of N per-lane partials across the warp with N-1-ish shuffles instead of 5N:
// at each level lanes swap half of their values.  Afterwards v[0..N/32) hold complete sums of the
// flat indices  base(lane) + i,  base = sum over bits b of lane: bit_b ? N >> (5 - b) : 0.
template <int N>
__device__ __forceinline__ void warp_reduce_scatter(float (&v)[N], int lane) {
    static_assert(N % 32 == 0, "N must be a multiple of 32");
#pragma unroll
    for (int lvl = 0; lvl < 5; lvl++) {
        const int mask = 16 >> lvl;
        const int half = (N / 2) >> lvl;
        const bool up = (lane & mask) != 0;
#pragma unroll
        for (int i = 0; i < half; i++) {
            const float send = up ? v[i] : v[i + half];
            const float keep = up ? v[i + half] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, mask);
        }
    }
}
template <int N>
__device__ __forceinline__ int reduce_scatter_base(int lane) {
    return ((lane >> 4) & 1) * (N / 2) + ((lane >> 3) & 1) * (N / 4) + ((lane >> 2) & 1) * (N / 8) +
           ((lane >> 1) & 1) * (N / 16) + (lane & 1) * (N / 32);
}

__device__ __forceinline__ float ld_cg(const float *p) { return __ldcg(p); }
__device__ __forceinline__ float4 ld_cg4(const float4 *p) { return __ldcg(p); }

}  // namespace wrnn

// Tagged-word exchange through L2 between the CTAs of one persistent launch (see loop_pipe.cu header).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace wrnn {

// ---------------------------------------------------------------------------- PTX access helpers
__device__ __forceinline__ uint4 ld_relaxed_v4(const void *p) {
    uint4 v;
    asm volatile("ld.relaxed.gpu.global.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t ld_relaxed_u32(const void *p) {
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long ld_relaxed_u64(const void *p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_v4(void *p, uint4 v) {
    asm volatile("st.relaxed.gpu.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void st_relaxed_u32(void *p, uint32_t v) {
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_relaxed_u64(void *p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

__device__ __forceinline__ uint32_t tag_word(float v, uint32_t tag) {      // round to 22 bits, insert tag
    return ((__float_as_uint(v) + 2u) & ~3u) | tag;
}
__device__ __forceinline__ float untag(uint32_t w) { return __uint_as_float(w & ~3u); }
__device__ __forceinline__ bool tags_ok(const uint4 &v, uint32_t tag) {
    return (((v.x ^ tag) | (v.y ^ tag) | (v.z ^ tag) | (v.w ^ tag)) & 3u) == 0u;
}

constexpr long long WATCHDOG_CLOCKS = 4000000000ll;     // ~2 s: a lost exchange becomes an error, not a hang

// named barriers (sub-CTA thread groups)
__device__ __forceinline__ void bar_sync(int id, int nthreads) {
    // barrier.sync, not its .aligned form bar.sync: lanes leave poll loops at different times, and a warp that arrives
    // in two fragments is counted twice by the aligned barrier (seen as a hang when the lane that issues a gather
    // sits in the group that then synchronises on a named barrier)
    asm volatile("barrier.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ void bar_arrive(int id, int nthreads) {
    asm volatile("barrier.arrive %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

}  // namespace wrnn

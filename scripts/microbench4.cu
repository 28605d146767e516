// All-to-all exchange through L2 as the loop kernels do it: G CTAs each publish 20 x 16 B tagged words per round,
// then every CTA gathers all G x 20 slots (40 KB at G = 128).  Measures clocks per round for several gather styles.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
__device__ __forceinline__ uint4 ld_relaxed_v4(const void *p) {
    uint4 v;
    asm volatile("ld.relaxed.gpu.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_v4(void *p, uint4 v) {
    asm volatile("st.relaxed.gpu.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_u64(const void *p) {
    unsigned long long v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned ld_acquire_u32(const void *p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_u64(void *p) {
    asm volatile("red.release.gpu.global.add.u64 [%0], 1;" ::"l"(p) : "memory");
}
__device__ __forceinline__ void st_release_u32(void *p, unsigned v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ bool ok4(const uint4 &v, uint32_t tag) { return (((v.x ^ tag) | (v.y ^ tag) | (v.z ^ tag) | (v.w ^ tag)) & 3u) == 0u; }
__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gsrc) {
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gsrc) : "memory");
}
// buf: [2][20][G] uint4.  style 0: every thread polls all its slots through registers (batches of 5)
// style 1: probe first slot, cp.async the rest, validate.   nth = threads that take part in the gather.
__global__ void __launch_bounds__(512, 1) k(uint4 *buf, long long *out, int rounds, int style, int nth, int work) {
    extern __shared__ uint4 sm[];
    const int G = gridDim.x, tid = threadIdx.x, n4 = 20 * G;
    long long t0 = clock64();
    float acc = tid;
    for (int r = 0; r < rounds; r++) {
        const uint32_t tag = (r + 1) & 3;
        uint4 *b = buf + (size_t)(r & 1) * n4;
        if (tid < 20) st_relaxed_v4(b + tid * G + blockIdx.x, make_uint4(tag | (r << 2), tag, tag, tag));
        for (int i = 0; i < work; i++) acc = fmaf(acc, 1.0001f, 1e-7f);      // optional compute between publish and gather
        if ((tid < nth || style == 5) && tid < n4) {
            if (style == 0) {
                for (int base = tid; base < n4; base += 5 * nth) {
                    uint4 v[5];
#pragma unroll
                    for (int j = 0; j < 5; j++) if (base + j * nth < n4) v[j] = ld_relaxed_v4(b + base + j * nth);
#pragma unroll
                    for (int j = 0; j < 5; j++) if (base + j * nth < n4) { for (int sp = 0; !ok4(v[j], tag) && sp < 2000000; sp++) v[j] = ld_relaxed_v4(b + base + j * nth); sm[base + j * nth] = v[j]; }
                }
            } else if (style == 2 || style == 3) {
                // flags: style 2 = one counter (red.release / ld.acquire), style 3 = one flag word per producer
                unsigned long long *ctr = reinterpret_cast<unsigned long long *>(buf + 3 * 4096);
                unsigned *flags = reinterpret_cast<unsigned *>(buf + 3 * 4096 + 64);
                if (tid < 32) {
                    __syncwarp();
                    if (style == 2) {
                        if (tid == 0) { red_release_u64(ctr); for (int sp = 0; ld_acquire_u64(ctr) < (unsigned long long)G * (r + 1) && sp < 2000000; sp++) {} }
                    } else {
                        if (tid == 0) st_release_u32(flags + blockIdx.x * 32, r + 1);      // one flag per 128 B line
                    }
                }
                if (style == 3 && tid < G) for (int sp = 0; ld_acquire_u32(flags + tid * 32) != (unsigned)(r + 1) && sp < 2000000; sp++) {}
                __syncthreads();
                for (int base = tid; base < n4; base += 5 * nth) {
                    uint4 v[5];
#pragma unroll
                    for (int j = 0; j < 5; j++) if (base + j * nth < n4) v[j] = ld_relaxed_v4(b + base + j * nth);
#pragma unroll
                    for (int j = 0; j < 5; j++) if (base + j * nth < n4) sm[base + j * nth] = v[j];
                }
            } else if (style == 4) {
                // dense relaxed flags (no fence): warp 0 polls 128 flags (4 per lane), then one gather pass; tags validate
                unsigned *flags = reinterpret_cast<unsigned *>(buf + 3 * 4096 + 64);
                if (tid < 32) {
                    __syncwarp();
                    if (tid == 0) asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(flags + blockIdx.x), "r"(r + 1) : "memory");
                    if (4 * tid < G) {
                        uint4 fv = ld_relaxed_v4(flags + 4 * tid);
                        for (int sp = 0; (fv.x != r + 1 || fv.y != r + 1 || fv.z != r + 1 || fv.w != r + 1) && sp < 2000000; sp++) fv = ld_relaxed_v4(flags + 4 * tid);
                    }
                }
                __syncthreads();
                for (int base = tid; base < n4; base += 5 * nth) {
                    uint4 v[5];
#pragma unroll
                    for (int j = 0; j < 5; j++) if (base + j * nth < n4) v[j] = ld_relaxed_v4(b + base + j * nth);
#pragma unroll
                    for (int j = 0; j < 5; j++) if (base + j * nth < n4) { for (int sp = 0; !ok4(v[j], tag) && sp < 2000000; sp++) v[j] = ld_relaxed_v4(b + base + j * nth); sm[base + j * nth] = v[j]; }
                }
            } else if (style == 5) {
                // spread probe: thread p < G polls producer p's slot of fold p % 20 (distinct lines -> distinct L2 slices),
                // then one gather pass by nth threads, tags validate
                if (tid < G) {
                    const uint4 *ps = b + (tid % 20) * G + tid;
                    uint4 v = ld_relaxed_v4(ps);
                    for (int sp = 0; !ok4(v, tag) && sp < 2000000; sp++) v = ld_relaxed_v4(ps);
                }
                __syncthreads();
                if (tid < nth)
                for (int base = tid; base < n4; base += 5 * nth) {
                    uint4 v[5];
#pragma unroll
                    for (int j = 0; j < 5; j++) if (base + j * nth < n4) v[j] = ld_relaxed_v4(b + base + j * nth);
#pragma unroll
                    for (int j = 0; j < 5; j++) if (base + j * nth < n4) { for (int sp = 0; !ok4(v[j], tag) && sp < 2000000; sp++) v[j] = ld_relaxed_v4(b + base + j * nth); sm[base + j * nth] = v[j]; }
                }
            } else {
                uint4 v = ld_relaxed_v4(b + tid);
                for (int sp = 0; !ok4(v, tag) && sp < 2000000; sp++) v = ld_relaxed_v4(b + tid);
                sm[tid] = v;
                for (int i = tid + nth; i < n4; i += nth) cp_async16(sm + i, b + i);
                asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
                for (int i = tid + nth; i < n4; i += nth) { uint4 w = sm[i]; for (int sp = 0; !ok4(w, tag) && sp < 2000000; sp++) w = ld_relaxed_v4(b + i); sm[i] = w; }
            }
        }
        __syncthreads();
    }
    long long t1 = clock64();
    if (tid == 0 && blockIdx.x == 0) out[0] = (t1 - t0) / rounds;
    if (acc == 1.2345f) out[1] = 1;
}
int main() {
    uint4 *buf; long long *out, h;
    cudaMalloc(&buf, 1 << 20); cudaMalloc(&out, 64);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    for (int G : {128})
        for (int style : {0, 5})
            for (int nth : {128, 256, 512})
                for (int work : {0, 500}) {
                    cudaMemset(buf, 0, 1 << 20);
                    int rounds = 2000;
                    void *args[] = {&buf, &out, &rounds, &style, &nth, &work};
                    cudaLaunchCooperativeKernel((void *)k, dim3(G), dim3(512), args, 48 * 1024, 0);
                    cudaDeviceSynchronize();
                    cudaMemcpy(&h, out, 8, cudaMemcpyDeviceToHost);
                    printf("G=%3d style=%s nth=%3d work=%4d FMA: %6lld clk per round  %s\n", G, style == 0 ? "register poll  " : style == 1 ? "probe+cp.async" : style == 2 ? "counter+gather " : style == 3 ? "flags+gather   " : style == 4 ? "dense flags    " : "spread probe   ", nth, work, h,
                           cudaGetErrorString(cudaGetLastError()));
                }
    return 0;
}

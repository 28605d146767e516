// Latency / throughput calibration for the per-sample loop kernels on B200 (clocks are SM cycles).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o microbench2 microbench2.cu && ./microbench2
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

__device__ __forceinline__ uint4 ld_relaxed_v4(const void *p) {
    uint4 v;
    asm volatile("ld.relaxed.gpu.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long ld_relaxed_u64(const void *p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u64(void *p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// (a) dependent FMA chain, (b) dependent LDS chain, (c) dependent shuffle chain, (d) __syncthreads
__global__ void k_basic(long long *out, float *sink, int threads_active) {
    __shared__ float sm[1024];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) sm[i] = (float)((i * 7 + 1) & 1023);
    __syncthreads();
    float a = threadIdx.x * 1e-3f, b = 1.0001f;
    long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < 1024; i++) a = fmaf(a, b, 1e-7f);
    long long t1 = clock64();
    int idx = threadIdx.x & 1023;
#pragma unroll 1
    for (int i = 0; i < 1024; i++) idx = (int)sm[idx] & 1023;
    long long t2 = clock64();
    float s = a;
#pragma unroll 1
    for (int i = 0; i < 1024; i++) s += __shfl_xor_sync(0xffffffffu, s, 1);
    long long t3 = clock64();
#pragma unroll 1
    for (int i = 0; i < 1024; i++) __syncthreads();
    long long t4 = clock64();
    float l = a + 2.f;
#pragma unroll 1
    for (int i = 0; i < 256; i++) l = logf(l) + 3.f;
    long long t5 = clock64();
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        out[0] = (t1 - t0);
        out[1] = (t2 - t1);
        out[2] = (t3 - t2);
        out[3] = (t4 - t3);
        out[4] = (t5 - t4) * 4;
    }
    if (a + s + idx + l == 1.2345f) sink[0] = a;
}

// (e) L2 round trip of ld.relaxed.gpu (dependent pointer chase in a small L2-resident buffer)
__global__ void k_l2lat(const unsigned long long *chain, long long *out, unsigned long long *sink) {
    unsigned long long p = 0;
    long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < 1024; i++) p = ld_relaxed_u64(chain + p);
    long long t1 = clock64();
    if (threadIdx.x == 0) out[0] = t1 - t0;
    if (p == 0xdeadbeef) sink[0] = p;
}

// (f) ping-pong between two CTAs through L2 with relaxed stores/loads: one-way signalling latency
__global__ void k_pingpong(unsigned long long *flags, long long *out, int iters) {
    if (threadIdx.x != 0) return;
    unsigned long long *mine = flags + blockIdx.x * 16, *other = flags + (1 - blockIdx.x) * 16;
    long long t0 = clock64();
    for (int i = 1; i <= iters; i++) {
        if (blockIdx.x == 0) {
            st_relaxed_u64(mine, i);
            while (ld_relaxed_u64(other) != (unsigned long long)i) {}
        } else {
            while (ld_relaxed_u64(other) != (unsigned long long)i) {}
            st_relaxed_u64(mine, i);
        }
    }
    long long t1 = clock64();
    if (blockIdx.x == 0) out[0] = (t1 - t0);
}

// (g) all-gather bandwidth: every CTA reads the same `bytes` from L2 into smem with all threads (5 loads in flight each)
__global__ void k_gather(const uint4 *src, int n4, long long *out, float *sink, int iters) {
    extern __shared__ uint4 smv[];
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
        for (int base = threadIdx.x; base < n4; base += 5 * blockDim.x) {
            uint4 v[5];
#pragma unroll
            for (int j = 0; j < 5; j++)
                if (base + j * blockDim.x < n4) v[j] = ld_relaxed_v4(src + base + j * blockDim.x);
#pragma unroll
            for (int j = 0; j < 5; j++)
                if (base + j * blockDim.x < n4) smv[base + j * blockDim.x] = v[j];
        }
        __syncthreads();
    }
    long long t1 = clock64();
    if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = (t1 - t0);
    if (smv[threadIdx.x % n4].x == 0x12345u) sink[0] = 1.f;
}

// (h) LDS.128 throughput: all lanes distinct addresses vs all lanes the same address
__global__ void k_lds(long long *out, float *sink, int bcast) {
    __shared__ float4 sm[2048];
    for (int i = threadIdx.x; i < 2048; i += blockDim.x) sm[i] = make_float4(i, 1, 2, 3);
    __syncthreads();
    float acc = 0.f;
    const int base = bcast ? (threadIdx.x >> 5) : threadIdx.x;
    long long t0 = clock64();
#pragma unroll 4
    for (int i = 0; i < 1024; i++) {
        const float4 v = sm[(base + i * 32) & 2047];
        acc += v.x + v.w;
    }
    long long t1 = clock64();
    __syncthreads();
    if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = t1 - t0;
    if (acc == 1.2345f) sink[0] = acc;
}

// (i) a 4-warp FMA + LDS job timed alone and while `nspin` other warps poll a flag in L2 that never changes
__global__ void k_interfere(const unsigned long long *flag, long long *out, float *sink, int nspin_warps, volatile int *stop) {
    __shared__ float4 sm[2560];
    __shared__ int done;
    for (int i = threadIdx.x; i < 2560; i += blockDim.x) sm[i] = make_float4(i * 1e-3f, 1, 2, 3);
    if (threadIdx.x == 0) done = 0;
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp < 4) {
        float acc[12];
        for (int i = 0; i < 12; i++) acc[i] = 0.f;
        const float w = 1.0001f + lane;
        long long t0 = clock64();
        for (int rep = 0; rep < 50; rep++)
            for (int f = 0; f < 20; f++) {
                const float4 xa = sm[f * 128 + warp * 32 + lane];
#pragma unroll
                for (int i = 0; i < 12; i++) {
                    acc[i] = fmaf(w, xa.x, acc[i]);
                    acc[i] = fmaf(w, xa.y, acc[i]);
                    acc[i] = fmaf(w, xa.z, acc[i]);
                    acc[i] = fmaf(w, xa.w, acc[i]);
                }
            }
        long long t1 = clock64();
        float s = 0;
        for (int i = 0; i < 12; i++) s += acc[i];
        if (s == 1.2345f) sink[0] = s;
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (threadIdx.x == 0) { out[0] = (t1 - t0) / 50; atomicExch(&done, 1); }
    } else if (warp < 4 + nspin_warps) {
        while (ld_relaxed_u64(flag + lane) != 77ull && *((volatile int *)&done) == 0) {}
    }
}

int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    const int G = p.multiProcessorCount;
    long long *out, h[8];
    float *sink;
    cudaMalloc(&out, 64);
    cudaMalloc(&sink, 64);
    for (int th : {32, 128, 512}) {
        k_basic<<<1, th>>>(out, sink, th);
        cudaDeviceSynchronize();
        cudaMemcpy(h, out, 40, cudaMemcpyDeviceToHost);
        printf("threads=%3d: dependent FMA %.1f clk, dependent LDS %.1f clk, SHFL+FADD %.1f clk, __syncthreads %.1f clk, logf+add %.1f clk\n",
               th, h[0] / 1024.0, h[1] / 1024.0, h[2] / 1024.0, h[3] / 1024.0, h[4] / 1024.0);
    }
    {   // pointer chase over 64 KB
        const int n = 8192;
        unsigned long long *hc = new unsigned long long[n], *dc, *ds;
        for (int i = 0; i < n; i++) hc[i] = (unsigned long long)((i * 1237 + 517) % n);
        cudaMalloc(&dc, n * 8);
        cudaMalloc(&ds, 64);
        cudaMemcpy(dc, hc, n * 8, cudaMemcpyHostToDevice);
        k_l2lat<<<1, 32>>>(dc, out, ds);
        k_l2lat<<<1, 32>>>(dc, out, ds);
        cudaDeviceSynchronize();
        cudaMemcpy(h, out, 8, cudaMemcpyDeviceToHost);
        printf("ld.relaxed.gpu dependent chain (L2 hit): %.1f clk per load\n", h[0] / 1024.0);
    }
    {
        unsigned long long *fl;
        cudaMalloc(&fl, 4096);
        for (int rep = 0; rep < 3; rep++) {
            cudaMemset(fl, 0, 4096);
            int iters = 2000;
            void *args[] = {&fl, &out, &iters};
            cudaLaunchCooperativeKernel((void *)k_pingpong, dim3(2), dim3(32), args, 0, 0);
            cudaDeviceSynchronize();
            cudaMemcpy(h, out, 8, cudaMemcpyDeviceToHost);
            printf("ping-pong through L2 (st.relaxed -> ld.relaxed poll), two CTAs: %.1f clk per one-way hop\n", h[0] / (2.0 * iters));
        }
    }
    {
        uint4 *src;
        cudaMalloc(&src, 1 << 20);
        cudaMemset(src, 1, 1 << 20);
        for (int kb : {2, 10, 40, 80}) {
            for (int grid : {1, 32, G}) {
                const int n4 = kb * 1024 / 16, iters = 200;
                cudaFuncSetAttribute(k_gather, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
                k_gather<<<grid, 512, kb * 1024>>>(src, n4, out, sink, iters);
                cudaDeviceSynchronize();
                cudaMemcpy(h, out, 8, cudaMemcpyDeviceToHost);
                printf("gather %2d KB from L2 into smem, %3d CTAs x 512 thr: %.0f clk per gather (%.1f B/clk/SM, %.0f B/clk chip)\n", kb,
                       grid, h[0] / (double)iters, kb * 1024.0 * iters / h[0], kb * 1024.0 * iters / h[0] * grid);
            }
        }
    }
    {
        unsigned long long *fl;
        cudaMalloc(&fl, 4096);
        cudaMemset(fl, 0, 4096);
        for (int ns : {0, 4, 12}) {
            k_interfere<<<1, 512>>>(fl, out, sink, ns, nullptr);
            cudaDeviceSynchronize();
            cudaMemcpy(h, out, 8, cudaMemcpyDeviceToHost);
            printf("4-warp job (20 LDS.128 + 960 FFMA per thread) with %2d warps polling L2: %lld clk\n", ns, h[0]);
        }
    }
    for (int bc : {0, 1})
        for (int th : {128, 512}) {
            k_lds<<<1, th>>>(out, sink, bc);
            cudaDeviceSynchronize();
            cudaMemcpy(h, out, 8, cudaMemcpyDeviceToHost);
            printf("LDS.128 %s, %3d threads: %.2f clk per warp-load (SM-wide %.2f)\n", bc ? "broadcast" : "distinct ", th,
                   h[0] / 1024.0, h[0] / 1024.0 / (th / 32));
        }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}

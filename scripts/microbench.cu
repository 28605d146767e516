// Microbenchmarks that size the per-sample loop design on B200 (run: nvcc ... && ./microbench).
//  1. grid barrier latency (monotonic counter, 148 CTAs)      2. mma.sync m16n8k16 bf16 issue rate
//  3. producer->consumer exchange through L2 (flag + data)
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdio.h>
#include <stdint.h>

__device__ __forceinline__ unsigned long long ld_acq(const unsigned long long *p) {
    unsigned long long v; asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ void red_rel(unsigned long long *p) {
    asm volatile("red.release.gpu.global.add.u64 [%0], %1;" ::"l"(p), "l"(1ull) : "memory"); }

__global__ void k_barrier(unsigned long long *ctr, int iters, long long *out) {
    unsigned long long gen = 0;
    long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
        __syncthreads();
        gen++;
        if (threadIdx.x == 0) { red_rel(ctr); while (ld_acq(ctr) < gen * gridDim.x) {} }
        __syncthreads();
    }
    if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = (clock64() - t0) / iters;
}

// barrier + every CTA publishes 2 KB and then reads all 148 slices (all-gather through L2)
__global__ void k_allgather(unsigned long long *ctr, float4 *buf, int iters, int per_cta_f4, long long *out, float *sink) {
    extern __shared__ float4 sm[];
    unsigned long long gen = 0;
    float acc = 0.f;
    long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
        float4 *mine = buf + ((size_t)(i & 1) * gridDim.x + blockIdx.x) * per_cta_f4;
        for (int j = threadIdx.x; j < per_cta_f4; j += blockDim.x) mine[j] = make_float4(i, j, 0, 1);
        __syncthreads();
        gen++;
        if (threadIdx.x == 0) { red_rel(ctr); while (ld_acq(ctr) < gen * gridDim.x) {} }
        __syncthreads();
        const float4 *all = buf + (size_t)(i & 1) * gridDim.x * per_cta_f4;
        const int n = gridDim.x * per_cta_f4;
        for (int j = threadIdx.x; j < n; j += blockDim.x) sm[j] = __ldcg(all + j);
        __syncthreads();
        acc += sm[(threadIdx.x * 7) % n].x;
    }
    if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = (clock64() - t0) / iters;
    if (acc == 12345.f) sink[0] = acc;
}

__global__ void k_mma(int iters, int nacc, long long *out, float *sink) {
    uint32_t a[4] = {0x3f803f80u, 0x3f803f80u, 0x3f803f80u, 0x3f803f80u}, b[2] = {0x3f803f80u, 0x3f803f80u};
    float c[8][4];
    for (int i = 0; i < 8; i++) for (int j = 0; j < 4; j++) c[i][j] = 0.f;
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (i < nacc)
                asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3])
                             : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
        }
    }
    long long t1 = clock64();
    __syncthreads();
    if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = t1 - t0;
    float s = 0; for (int i = 0; i < 8; i++) s += c[i][0];
    if (s == 1.2345f) sink[0] = s;
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int G = p.multiProcessorCount;
    unsigned long long *ctr; long long *out; float *sink; float4 *buf;
    cudaMalloc(&ctr, 64); cudaMalloc(&out, 64); cudaMalloc(&sink, 64); cudaMalloc(&buf, 64 << 20);
    long long h;
    for (int threads : {256, 512}) {
        cudaMemset(ctr, 0, 64);
        int iters = 20000;
        void *args[] = {&ctr, &iters, &out};
        cudaLaunchCooperativeKernel((void *)k_barrier, dim3(G), dim3(threads), args, 0, 0);
        cudaDeviceSynchronize(); cudaMemcpy(&h, out, 8, cudaMemcpyDeviceToHost);
        printf("grid barrier  G=%d threads=%d : %lld clk  (%s)\n", G, threads, h, cudaGetErrorString(cudaGetLastError()));
    }
    for (int f4 : {32, 80, 160}) {       // 0.5, 1.25, 2.5 KB per CTA -> 74..370 KB gathered per CTA
        cudaMemset(ctr, 0, 64);
        int iters = 5000; size_t smem = (size_t)G * f4 * 16;
        cudaFuncSetAttribute(k_allgather, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        void *args[] = {&ctr, &buf, &iters, &f4, &out, &sink};
        cudaLaunchCooperativeKernel((void *)k_allgather, dim3(G), dim3(512), args, smem, 0);
        cudaDeviceSynchronize(); cudaMemcpy(&h, out, 8, cudaMemcpyDeviceToHost);
        printf("publish+barrier+allgather %5zu B/CTA -> %6zu B per CTA: %lld clk  (%s)\n", (size_t)f4 * 16, smem, h,
               cudaGetErrorString(cudaGetLastError()));
    }
    for (int warps : {1, 4, 8, 16}) for (int nacc : {1, 2, 4, 8}) {
        int iters = 2000;
        k_mma<<<G, warps * 32>>>(iters, nacc, out, sink);
        cudaDeviceSynchronize(); cudaMemcpy(&h, out, 8, cudaMemcpyDeviceToHost);
        printf("mma.sync bf16 m16n8k16 warps/SM=%2d indep acc=%d : %.2f clk per mma per warp, %.2f clk per mma per SM\n", warps, nacc,
               (double)h / (iters * nacc), (double)h / (iters * nacc * warps));
    }
    return 0;
}

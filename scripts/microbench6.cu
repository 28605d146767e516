// LDS.128 / LDS.32 throughput per SM with unrolled, independent loads (no address arithmetic in the loop).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o microbench6 microbench6.cu && ./microbench6
#include <cuda_runtime.h>
#include <stdio.h>
template <int W>
__global__ void k(long long *out, float *sink, int iters) {
    extern __shared__ float4 sm[];
    for (int i = threadIdx.x; i < 2560; i += blockDim.x) sm[i] = make_float4(i, 1, 2, 3);
    __syncthreads();
    const float4 *p = sm + (threadIdx.x & 127);
    const float *q = reinterpret_cast<const float *>(sm) + (threadIdx.x & 127);
    float acc = 0.f;
    const unsigned pa = (unsigned)__cvta_generic_to_shared(p), qa = (unsigned)__cvta_generic_to_shared(q);
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int f = 0; f < 20; f++) {
            if (W == 4) {
                float4 v;
                asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(pa + f * 2048));
                acc += (v.x + v.y) + (v.z + v.w);
            } else {
                float v;
                asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(qa + f * 2048));
                acc += v;
            }
        }
    }
    long long t1 = clock64();
    __syncthreads();
    if (threadIdx.x == 0) out[0] = t1 - t0;
    if (acc == 1.2345f) sink[0] = acc;
}
int main() {
    long long *out, h; float *sink;
    cudaMalloc(&out, 64); cudaMalloc(&sink, 64);
    cudaFuncSetAttribute(k<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 48 * 1024);
    cudaFuncSetAttribute(k<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 48 * 1024);
    for (int th : {128, 256, 512}) {
        const int iters = 200;
        k<4><<<1, th, 41 * 1024>>>(out, sink, iters); cudaDeviceSynchronize();
        cudaMemcpy(&h, out, 8, cudaMemcpyDeviceToHost);
        printf("LDS.128 %3d threads: %.2f clk per warp-load SM-wide (%.1f B/clk)\n", th, h / (double)(iters * 20 * (th / 32)), 512.0 * iters * 20 * (th / 32) / h);
        k<1><<<1, th, 41 * 1024>>>(out, sink, iters); cudaDeviceSynchronize();
        cudaMemcpy(&h, out, 8, cudaMemcpyDeviceToHost);
        printf("LDS.32  %3d threads: %.2f clk per warp-load SM-wide (%.1f B/clk)\n", th, h / (double)(iters * 20 * (th / 32)), 128.0 * iters * 20 * (th / 32) / h);
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}

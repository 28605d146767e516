// SHFL / LDS / STS issue rates per SM (independent operations, volatile asm), 4..16 warps.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o microbench8 microbench8.cu && ./microbench8
#include <cuda_runtime.h>
#include <stdio.h>
template <int MODE>
__global__ void k(long long *out, float *sink, int iters) {
    extern __shared__ float4 sm[];
    for (int i = threadIdx.x; i < 2560; i += blockDim.x) sm[i] = make_float4(i, 1, 2, 3);
    __syncthreads();
    const unsigned base = (unsigned)__cvta_generic_to_shared(sm) + threadIdx.x * 16;
    float v[8];
    for (int i = 0; i < 8; i++) v[i] = threadIdx.x + i;
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (MODE == 0) {
                asm volatile("shfl.sync.bfly.b32 %0, %0, 1, 0x1f, 0xffffffff;" : "+f"(v[i]));
            } else if (MODE == 1) {
                float4 x;
                asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(x.x), "=f"(x.y), "=f"(x.z), "=f"(x.w) : "r"(base + (i & 3) * 8192));
                v[i] += x.x + x.w;
            } else if (MODE == 2) {
                float x;
                asm volatile("ld.shared.f32 %0, [%1];" : "=f"(x) : "r"(base / 4 * 1 + 0 + ((unsigned)__cvta_generic_to_shared(sm)) * 0 + (i & 3) * 8192 + threadIdx.x * 0));
                v[i] += x;
            } else {
                asm volatile("st.shared.f32 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(sm) + threadIdx.x * 4 + (i & 3) * 8192), "f"(v[i]));
            }
        }
    }
    long long t1 = clock64();
    __syncthreads();
    float s = 0;
    for (int i = 0; i < 8; i++) s += v[i];
    if (threadIdx.x == 0) out[0] = t1 - t0;
    if (s == 1.2345f) sink[0] = s;
}
int main() {
    long long *out, h; float *sink;
    cudaMalloc(&out, 64); cudaMalloc(&sink, 64);
    const int iters = 2000;
    const char *names[4] = {"SHFL.BFLY", "LDS.128 (distinct)", "LDS.32", "STS.32"};
    for (int mode = 0; mode < 4; mode++)
        for (int th : {32, 128, 256, 512}) {
            void (*kern)(long long *, float *, int) = mode == 0 ? k<0> : mode == 1 ? k<1> : mode == 2 ? k<2> : k<3>;
            cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 48 * 1024);
            kern<<<1, th, 41 * 1024>>>(out, sink, iters); cudaDeviceSynchronize();
            cudaMemcpy(&h, out, 8, cudaMemcpyDeviceToHost);
            printf("%-20s %3d threads: %.2f clk per warp-instr per warp, %.2f warp-instr per clk per SM\n", names[mode], th,
                   h / (double)(iters * 8), (double)iters * 8 * (th / 32) / h);
        }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}

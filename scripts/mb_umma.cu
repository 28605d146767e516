// Microbenchmark / bring-up harness for the tcgen05 path of csrc/loop_mma.cu (run on a B200 through gpurun):
//   1. descriptor check: D[128 x N] = A[128 x K] . B[N x K]^T with bf16 operands in the K-major, no-swizzle
//      "[k-chunk][row][8 elements]" shared-memory image the loop kernel uses (cp.async.bulk of pre-tiled global
//      memory, no tensor maps), for both readings of the LBO / SBO descriptor fields, against a CPU product;
//   2. the same with hi/lo split operands (3 MMAs per k-step) against the fp32 product (what bf16x3 buys);
//   3. a 96-row A image addressed as M = 128 (rows 96..127 alias the next chunk: garbage rows, never read);
//   4. tcgen05.mma issue rate for M = 128, N in {16, 32, 80, 128, 256} (clock64 around a chain of MMAs);
//   5. L2 -> shared-memory streaming rate with cp.async.bulk rings on 1 .. 148 CTAs (bytes / clk / SM).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o scripts/mb_umma scripts/mb_umma.cu
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include <vector>

#define CK(x)                                                                                     \
    do {                                                                                          \
        cudaError_t e_ = (x);                                                                     \
        if (e_ != cudaSuccess) {                                                                  \
            printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__);       \
            exit(1);                                                                              \
        }                                                                                         \
    } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(void *b, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(void *b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(void *b) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(void *b, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(smem_u32(b)), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ bool mbar_wait(void *b, uint32_t parity) {      // bounded: a bring-up bug must not hang the box
    const long long t0 = clock64();
    while (!mbar_try_wait(b, parity))
        if (clock64() - t0 > 200000000ll) return false;
    return true;
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, void *b) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(b)) : "memory");
}
// ---- tcgen05
__device__ __forceinline__ void tmem_alloc(uint32_t *dst_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accum) : "memory");
}
__device__ __forceinline__ void umma_commit(void *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
          "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
          "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
          "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
// K-major, no swizzle: element (row r, k) at start + (r % 8) * 16 + (r / 8) * SBO + (k / 8) * LBO + (k % 8) * 2 bytes
__host__ __device__ inline uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3fff);
    d |= (uint64_t)((lbo >> 4) & 0x3fff) << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3fff) << 32;
    d |= (uint64_t)1 << 46;           // descriptor version (Blackwell)
    return d;                         // base offset 0, lbo mode 0, layout type 0 = no swizzle
}
__host__ __device__ inline uint32_t make_idesc(int M, int N) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);   // f32 acc, bf16 x bf16, K-major both
}

struct TestArgs {
    const __nv_bfloat16 *Ahi, *Alo, *Bhi, *Blo;     // images [chunk][rows][8]
    float *D;                                       // [128][N]
    int rowsA, N, nchunks, passes, swap_lbo_sbo, reps;
    long long *clk;                                 // [2]: mma clocks, status
};

// one CTA, staged with CTA barriers (load -> MMA chain -> TMEM read): warp 0 loads, warp 1 issues, warps 2..5 read TMEM
__global__ void __launch_bounds__(192, 1) umma_test(TestArgs a) {
    extern __shared__ __align__(128) uint8_t sm[];
    __shared__ __align__(8) uint64_t bar_load, bar_mma;
    __shared__ uint32_t tmem_base;
    __shared__ int s_ok;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t bytesA = (uint32_t)a.nchunks * a.rowsA * 16, bytesB = (uint32_t)a.nchunks * a.N * 16;
    uint8_t *sAhi = sm, *sAlo = sAhi + bytesA, *sBhi = sAlo + bytesA, *sBlo = sBhi + bytesB;   // + 1 KB slack behind for the aliased rows
    if (threadIdx.x == 0) {
        mbar_init(&bar_load, 1);
        mbar_init(&bar_mma, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        s_ok = 1;
    }
    __syncthreads();
    if (warp == 1) tmem_alloc(&tmem_base, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tb = tmem_base;
    if (warp == 0 && lane == 0) {
        mbar_expect_tx(&bar_load, 2 * bytesA + 2 * bytesB);
        bulk_g2s(sAhi, a.Ahi, bytesA, &bar_load);
        bulk_g2s(sAlo, a.Alo, bytesA, &bar_load);
        bulk_g2s(sBhi, a.Bhi, bytesB, &bar_load);
        bulk_g2s(sBlo, a.Blo, bytesB, &bar_load);
        if (!mbar_wait(&bar_load, 0)) s_ok = 0;
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        if (lane == 0) {
            const uint32_t idesc = make_idesc(128, a.N);
            uint32_t lboA = (uint32_t)a.rowsA * 16, lboB = (uint32_t)a.N * 16, sbo = 128;
            const long long t0 = clock64();
            for (int rep = 0; rep < a.reps; rep++) {
                for (int ks = 0; ks < a.nchunks / 2; ks++) {
                    const uint32_t offA = (uint32_t)ks * 2 * lboA, offB = (uint32_t)ks * 2 * lboB;
                    uint64_t dAh, dAl, dBh, dBl;
                    if (!a.swap_lbo_sbo) {
                        dAh = make_desc(smem_u32(sAhi) + offA, lboA, sbo); dAl = make_desc(smem_u32(sAlo) + offA, lboA, sbo);
                        dBh = make_desc(smem_u32(sBhi) + offB, lboB, sbo); dBl = make_desc(smem_u32(sBlo) + offB, lboB, sbo);
                    } else {
                        dAh = make_desc(smem_u32(sAhi) + offA, sbo, lboA); dAl = make_desc(smem_u32(sAlo) + offA, sbo, lboA);
                        dBh = make_desc(smem_u32(sBhi) + offB, sbo, lboB); dBl = make_desc(smem_u32(sBlo) + offB, sbo, lboB);
                    }
                    umma_bf16(tb, dAh, dBh, idesc, (ks > 0 || rep > 0) ? 1u : 0u);
                    if (a.passes == 3) {
                        umma_bf16(tb, dAh, dBl, idesc, 1u);
                        umma_bf16(tb, dAl, dBh, idesc, 1u);
                    }
                }
            }
            umma_commit(&bar_mma);
            if (!mbar_wait(&bar_mma, 0)) s_ok = 0;
            const long long t1 = clock64();
            a.clk[0] = t1 - t0;
        }
        __syncwarp();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (threadIdx.x == 0) a.clk[1] = s_ok ? 1 : -1;
    if (warp >= 2) {
        const int q = warp & 3, row = q * 32 + lane;
        for (int c0 = 0; c0 < a.N; c0 += 32) {
            uint32_t r[32];
            tmem_ld32(tb + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, r);
            for (int j = 0; j < 32; j++)
                if (c0 + j < a.N) a.D[(size_t)row * a.N + c0 + j] = __uint_as_float(r[j]);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tb, 512);
}

// Staged probe with progress markers in mapped host memory: which tcgen05 step does not come back?
//   mode bit 0: TMEM alloc / dealloc, bit 1: bulk copies, bit 2: one MMA + commit, bit 3: tcgen05.ld
__global__ void __launch_bounds__(192, 1) probe(TestArgs a, int mode, volatile int *mark) {
    extern __shared__ __align__(128) uint8_t sm[];
    __shared__ __align__(8) uint64_t bar_load, bar_mma;
    __shared__ uint32_t tmem_base;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t bytesA = (uint32_t)a.nchunks * a.rowsA * 16, bytesB = (uint32_t)a.nchunks * a.N * 16;
    uint8_t *sAhi = sm, *sBhi = sm + 2 * bytesA;
    if (threadIdx.x == 0) {
        mbar_init(&bar_load, 1);
        mbar_init(&bar_mma, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        mark[0] = 1;
    }
    if (threadIdx.x == 0) tmem_base = 0;
    __syncthreads();
    if ((mode & 1) && warp == 1) {
        tmem_alloc(&tmem_base, 512);
        if (lane == 0) mark[1] = 1;
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tb = tmem_base;
    if (threadIdx.x == 0) { mark[2] = (int)tb; mark[0] = 2; }
    if ((mode & 2) && warp == 0 && lane == 0) {
        mbar_expect_tx(&bar_load, bytesA + bytesB);
        bulk_g2s(sAhi, a.Ahi, bytesA, &bar_load);
        bulk_g2s(sBhi, a.Bhi, bytesB, &bar_load);
        mark[3] = 1;
        const bool ok = mbar_wait(&bar_load, 0);
        mark[3] = ok ? 2 : -1;
    }
    __syncthreads();
    if ((mode & 4) && warp == 1) {
        tc_fence_after();
        if (lane == 0) {
            const uint32_t idesc = make_idesc(128, a.N);
            const uint32_t lboA = (uint32_t)a.rowsA * 16, lboB = (uint32_t)a.N * 16;
            mark[4] = 1;
            for (int ks = 0; ks < a.nchunks / 2; ks++)
                umma_bf16(tb, make_desc(smem_u32(sAhi) + ks * 2 * lboA, lboA, 128), make_desc(smem_u32(sBhi) + ks * 2 * lboB, lboB, 128),
                          idesc, ks > 0 ? 1u : 0u);
            mark[4] = 2;
            umma_commit(&bar_mma);
            mark[4] = 3;
            const bool ok = mbar_wait(&bar_mma, 0);
            mark[4] = ok ? 4 : -1;
        }
        __syncwarp();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if ((mode & 8) && warp >= 2) {
        const int q = warp & 3;
        uint32_t r[32];
        if (lane == 0) mark[5 + q] = 1;
        tmem_ld32(tb + ((uint32_t)(q * 32) << 16), r);
        if (lane == 0) mark[5 + q] = 2;
        a.D[(size_t)(q * 32 + lane) * a.N] = __uint_as_float(r[0]);
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x == 0) mark[0] = 3;
    if ((mode & 1) && warp == 1) {
        tmem_dealloc(tb, 512);
        if (lane == 0) mark[9] = 1;
    }
    __syncthreads();
    if (threadIdx.x == 0) mark[0] = 4;
}

// L2 -> smem streaming: every CTA walks the whole buffer (same bytes for all: L2-resident) through a ring of bulk copies
__global__ void __launch_bounds__(64, 1) l2_stream(const uint8_t *buf, size_t buf_bytes, uint32_t chunk, int nring, int nchunks_total,
                                                   long long *clk) {
    extern __shared__ __align__(128) uint8_t sm[];
    __shared__ __align__(8) uint64_t bars[16];
    if (threadIdx.x == 0) {
        for (int i = 0; i < nring; i++) mbar_init(&bars[i], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const size_t per = buf_bytes / chunk;
        size_t pos = ((size_t)blockIdx.x * 37) % per;
        const long long t0 = clock64();
        int issued = 0, done = 0;
        bool ok = true;
        for (; issued < nring && issued < nchunks_total; issued++) {
            mbar_expect_tx(&bars[issued], chunk);
            bulk_g2s(sm + (size_t)issued * chunk, buf + pos * chunk, chunk, &bars[issued]);
            pos = (pos + 1) % per;
        }
        while (done < nchunks_total && ok) {
            const int slot = done % nring;
            ok = mbar_wait(&bars[slot], (done / nring) & 1);
            done++;
            if (issued < nchunks_total) {
                mbar_expect_tx(&bars[slot], chunk);
                bulk_g2s(sm + (size_t)slot * chunk, buf + pos * chunk, chunk, &bars[slot]);
                pos = (pos + 1) % per;
                issued++;
            }
        }
        clk[blockIdx.x] = ok ? clock64() - t0 : -1;
    }
}


// Streaming with `nissue` issuing warps (lane 0 of each), each its own ring: is the per-copy cost in the issuing thread or in the TMA unit?
// Also measures the issue cost of one UBLKCP (clock around the instruction) and the latency of an isolated copy.
__global__ void __launch_bounds__(256, 1) l2_stream2(const uint8_t *buf, size_t buf_bytes, uint32_t chunk, int nring, int per_warp, int nissue,
                                                     long long *out) {
    extern __shared__ __align__(128) uint8_t sm[];
    __shared__ __align__(8) uint64_t bars[8][8];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int w = 0; w < 8; w++)
            for (int i = 0; i < 8; i++) mbar_init(&bars[w][i], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (warp < nissue && lane == 0) {
        uint8_t *ring = sm + (size_t)warp * nring * chunk;
        const size_t per = buf_bytes / chunk;
        size_t pos = ((size_t)blockIdx.x * 37 + warp * 101) % per;
        // isolated copy latency and issue cost
        long long c0 = clock64();
        mbar_expect_tx(&bars[warp][0], chunk);
        bulk_g2s(ring, buf + pos * chunk, chunk, &bars[warp][0]);
        long long c1 = clock64();
        bool ok = mbar_wait(&bars[warp][0], 0);
        long long c2 = clock64();
        long long issue_sum = 0;
        const long long t0 = clock64();
        int issued = 0, done = 0;
        for (; issued < nring && issued < per_warp; issued++) {
            mbar_expect_tx(&bars[warp][issued], chunk);
            bulk_g2s(ring + (size_t)issued * chunk, buf + pos * chunk, chunk, &bars[warp][issued]);
            pos = (pos + 1) % per;
        }
        while (done < per_warp && ok) {
            const int slot = done % nring;
            // slot 0 has already completed one phase (the isolated copy)
            const uint32_t par = (uint32_t)(((done / nring) + (slot == 0 ? 1 : 0)) & 1);
            ok = mbar_wait(&bars[warp][slot], par);
            done++;
            if (issued < per_warp) {
                const long long i0 = clock64();
                mbar_expect_tx(&bars[warp][slot], chunk);
                bulk_g2s(ring + (size_t)slot * chunk, buf + pos * chunk, chunk, &bars[warp][slot]);
                issue_sum += clock64() - i0;
                pos = (pos + 1) % per;
                issued++;
            }
        }
        const long long t1 = clock64();
        if (blockIdx.x == 0 || warp == 0) {
            long long *o = out + ((size_t)blockIdx.x * 8 + warp) * 4;
            o[0] = ok ? t1 - t0 : -1;
            o[1] = c1 - c0;
            o[2] = c2 - c1;
            o[3] = issue_sum / (per_warp > nring ? per_warp - nring : 1);
        }
    }
}

// MMA chain timing with a minimal issue loop: `nacc` independent accumulators used round-robin (1 = one dependent chain),
// `nmma` MMAs in total on resident operands.  Separates issue overhead, dependent-accumulate latency and the tensor floor.
__global__ void __launch_bounds__(64, 1) mma_chain(int N, int nacc, int nmma, long long *out) {
    extern __shared__ __align__(128) uint8_t sm[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_base;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 16384; i += 64) reinterpret_cast<uint32_t *>(sm)[i] = 0;     // 64 KB of zero operands
    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (warp == 1) tmem_alloc(&tmem_base, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tb = tmem_base;
    if (warp == 1 && lane == 0) {
        const uint32_t idesc = make_idesc(128, N);
        const uint64_t dA = make_desc(smem_u32(sm), 128 * 16, 128), dB = make_desc(smem_u32(sm) + 8192, (uint32_t)N * 16, 128);
        const long long t0 = clock64();
        int acc = 0;
        for (int i = 0; i < nmma; i += 4) {
#pragma unroll
            for (int j = 0; j < 4; j++) {
                umma_bf16(tb + (uint32_t)(acc * N), dA + (uint64_t)(j * 2), dB, idesc, 1u);
                acc = acc + 1 == nacc ? 0 : acc + 1;
            }
        }
        const long long t1 = clock64();
        umma_commit(&bar);
        const bool ok = mbar_wait(&bar, 0);
        const long long t2 = clock64();
        out[0] = ok ? t2 - t0 : -1;
        out[1] = t1 - t0;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tb, 512);
}

// Same streaming test with the issuers as LANES OF ONE WARP (mode 0) or with mbarrier.test_wait polling instead of try_wait
// (mode 1, issuers = warps): what is the ~680 clk per copy per issuing thread made of?
__device__ __forceinline__ bool mbar_test_wait(void *b, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(smem_u32(b)), "r"(parity) : "memory");
    return ok != 0;
}
__global__ void __launch_bounds__(256, 1) l2_stream3(const uint8_t *buf, size_t buf_bytes, uint32_t chunk, int nring, int per_issuer, int nissue,
                                                     int mode, long long *out) {
    extern __shared__ __align__(128) uint8_t sm[];
    __shared__ __align__(8) uint64_t bars[8][8];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int w = 0; w < 8; w++)
            for (int i = 0; i < 8; i++) mbar_init(&bars[w][i], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const int me = mode == 0 ? lane : warp;
    const bool active = mode == 0 ? (warp == 0 && lane < nissue) : (warp < nissue && lane == 0);
    if (active) {
        uint8_t *ring = sm + (size_t)me * nring * chunk;
        const size_t per = buf_bytes / chunk;
        size_t pos = ((size_t)blockIdx.x * 37 + me * 101) % per;
        const long long t0 = clock64();
        int issued = 0, done = 0;
        bool ok = true;
        for (; issued < nring && issued < per_issuer; issued++) {
            mbar_expect_tx(&bars[me][issued], chunk);
            bulk_g2s(ring + (size_t)issued * chunk, buf + pos * chunk, chunk, &bars[me][issued]);
            pos = (pos + 1) % per;
        }
        while (done < per_issuer && ok) {
            const int slot = done % nring;
            const uint32_t par = (uint32_t)((done / nring) & 1);
            if (mode == 1) {
                const long long w0 = clock64();
                while (!mbar_test_wait(&bars[me][slot], par))
                    if (clock64() - w0 > 200000000ll) { ok = false; break; }
            } else {
                ok = mbar_wait(&bars[me][slot], par);
            }
            done++;
            if (issued < per_issuer) {
                mbar_expect_tx(&bars[me][slot], chunk);
                bulk_g2s(ring + (size_t)slot * chunk, buf + pos * chunk, chunk, &bars[me][slot]);
                pos = (pos + 1) % per;
                issued++;
            }
        }
        if (me == 0) out[blockIdx.x] = ok ? clock64() - t0 : -1;
    }
}

static float bf16_round(float x) { return __bfloat162float(__float2bfloat16(x)); }

int main(int argc, char **argv) {
    setvbuf(stdout, nullptr, _IONBF, 0);
    if (argc > 1 && !strcmp(argv[1], "probe")) {
        const int mode = argc > 2 ? atoi(argv[2]) : 15;
        CK(cudaSetDeviceFlags(cudaDeviceMapHost));
        CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        volatile int *mark;
        int *dmark;
        CK(cudaHostAlloc((void **)&mark, 64 * sizeof(int), cudaHostAllocMapped));
        CK(cudaHostGetDevicePointer((void **)&dmark, (void *)mark, 0));
        for (int i = 0; i < 64; i++) mark[i] = 0;
        const int N = 64, K = 64, nch = K / 8;
        std::vector<__nv_bfloat16> A((size_t)nch * 128 * 8, __float2bfloat16(1.f)), B((size_t)nch * N * 8, __float2bfloat16(0.5f));
        __nv_bfloat16 *dA, *dB;
        float *dD;
        CK(cudaMalloc(&dA, A.size() * 2)); CK(cudaMalloc(&dB, B.size() * 2)); CK(cudaMalloc(&dD, 128 * N * 4));
        CK(cudaMemcpy(dA, A.data(), A.size() * 2, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(dB, B.data(), B.size() * 2, cudaMemcpyHostToDevice));
        CK(cudaMemset(dD, 0, 128 * N * 4));
        TestArgs ta{dA, dA, dB, dB, dD, 128, N, nch, 1, 0, 1, nullptr};
        probe<<<1, 192, 2 * nch * 128 * 16 + 2 * nch * N * 16 + 1024>>>(ta, mode, dmark);
        cudaError_t le = cudaGetLastError();
        printf("probe mode %d launched: %s\n", mode, cudaGetErrorString(le));
        for (int i = 0; i < 50; i++) {
            if (cudaStreamQuery(0) != cudaErrorNotReady) break;
            usleep(100000);
        }
        cudaError_t qe = cudaStreamQuery(0);
        printf("probe mode %d: query=%s marks: stage=%d alloc=%d tmem_base=0x%x tma=%d mma=%d ld=[%d %d %d %d] dealloc=%d\n", mode,
               cudaGetErrorString(qe), mark[0], mark[1], mark[2], mark[3], mark[4], mark[5], mark[6], mark[7], mark[8], mark[9]);
        if (qe == cudaSuccess) {
            float d[4];
            CK(cudaMemcpy(d, dD, sizeof(d), cudaMemcpyDeviceToHost));
            float d2;
            CK(cudaMemcpy(&d2, dD + 64 * N, 4, cudaMemcpyDeviceToHost));
            printf("probe D[0][0]=%g (expect K*0.5=%g when MMA ran) D[64][0]=%g\n", d[0], K * 0.5, d2);
            return 0;
        }
        fflush(stdout);
        _exit(3);
    }
    int dev = 0;
    CK(cudaSetDevice(dev));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, dev));
    printf("device %s, %d SMs, cc %d.%d\n", prop.name, prop.multiProcessorCount, prop.major, prop.minor);

    // ---- streaming: issuers as lanes of one warp / test_wait polling
    if (argc > 1 && !strcmp(argv[1], "stream3")) {
        CK(cudaFuncSetAttribute(l2_stream3, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        const size_t bytes = 16u << 20;
        uint8_t *buf;
        long long *dout;
        CK(cudaMalloc(&buf, bytes));
        CK(cudaMemset(buf, 1, bytes));
        CK(cudaMalloc(&dout, 148 * sizeof(long long)));
        struct Cfg { uint32_t chunk; int nring, nissue, mode; };
        const Cfg cfgs[] = {{12288, 4, 1, 0}, {12288, 4, 2, 0}, {12288, 4, 4, 0}, {12288, 2, 8, 0}, {12288, 4, 1, 1}, {12288, 4, 4, 1},
                            {24576, 2, 4, 0}, {24576, 4, 2, 0}, {24576, 2, 4, 1}, {4096, 4, 4, 0}, {4096, 4, 8, 0}};
        for (int grid : {1, 148})
            for (const Cfg &c : cfgs) {
                const int per = (int)((4u << 20) / c.chunk / c.nissue);
                const size_t smem = (size_t)c.nring * c.chunk * c.nissue;
                if (smem > 200 * 1024) continue;
                for (int rep = 0; rep < 2; rep++) {
                    l2_stream3<<<grid, 256, smem>>>(buf, bytes, c.chunk, c.nring, per, c.nissue, c.mode, dout);
                    CK(cudaDeviceSynchronize());
                }
                std::vector<long long> o(grid);
                CK(cudaMemcpy(o.data(), dout, grid * sizeof(long long), cudaMemcpyDeviceToHost));
                long long mx = 0;
                for (auto v : o) mx = v > mx ? v : mx;
                printf("l2 stream3 grid=%3d chunk=%5u ring=%d issuers=%d (%s): %.1f B/clk/SM, %.0f clk per copy per issuer\n", grid, c.chunk,
                       c.nring, c.nissue, c.mode == 0 ? "lanes of one warp, try_wait" : "warps, test_wait poll",
                       (double)per * c.nissue * c.chunk / mx, (double)mx / per);
            }
        return 0;
    }
    CK(cudaFuncSetAttribute(umma_test, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CK(cudaFuncSetAttribute(l2_stream, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    long long *dclk;
    CK(cudaMalloc(&dclk, 256 * sizeof(long long)));

    struct Case { int rowsA, N, K, passes, swap, split; };
    const Case cases[] = {
        {128, 64, 64, 1, 0, 0},   {128, 64, 64, 1, 1, 0},        // descriptor reading
        {128, 128, 128, 1, 0, 0}, {128, 128, 128, 1, 1, 0},
        {128, 80, 128, 1, 0, 0},
        {96, 80, 256, 1, 0, 0},                                   // aliased rows
        {128, 16, 128, 1, 0, 0},  {128, 32, 256, 1, 0, 0},  {128, 256, 64, 1, 0, 0},
        {128, 128, 64, 3, 0, 1}, {96, 80, 128, 3, 0, 1},         // hi/lo split
    };
    srand(1234);
    for (const Case &c : cases) {
        const int nch = c.K / 8;
        std::vector<float> A((size_t)128 * c.K), B((size_t)c.N * c.K);
        for (auto &v : A) v = (float)rand() / RAND_MAX * 2.f - 1.f;
        for (auto &v : B) v = (float)rand() / RAND_MAX * 2.f - 1.f;
        std::vector<__nv_bfloat16> Ahi((size_t)nch * c.rowsA * 8 + 4096), Alo(Ahi.size()), Bhi((size_t)nch * c.N * 8), Blo(Bhi.size());
        for (int r = 0; r < c.rowsA; r++)
            for (int k = 0; k < c.K; k++) {
                const float v = A[(size_t)r * c.K + k], hi = bf16_round(v);
                const size_t o = ((size_t)(k / 8) * c.rowsA + r) * 8 + k % 8;
                Ahi[o] = __float2bfloat16(hi);
                Alo[o] = __float2bfloat16(v - hi);
            }
        for (int r = 0; r < c.N; r++)
            for (int k = 0; k < c.K; k++) {
                const float v = B[(size_t)r * c.K + k], hi = bf16_round(v);
                const size_t o = ((size_t)(k / 8) * c.N + r) * 8 + k % 8;
                Bhi[o] = __float2bfloat16(hi);
                Blo[o] = __float2bfloat16(v - hi);
            }
        __nv_bfloat16 *dAhi, *dAlo, *dBhi, *dBlo;
        float *dD;
        CK(cudaMalloc(&dAhi, Ahi.size() * 2)); CK(cudaMalloc(&dAlo, Alo.size() * 2));
        CK(cudaMalloc(&dBhi, Bhi.size() * 2)); CK(cudaMalloc(&dBlo, Blo.size() * 2));
        CK(cudaMalloc(&dD, (size_t)128 * c.N * 4));
        CK(cudaMemcpy(dAhi, Ahi.data(), Ahi.size() * 2, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(dAlo, Alo.data(), Alo.size() * 2, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(dBhi, Bhi.data(), Bhi.size() * 2, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(dBlo, Blo.data(), Blo.size() * 2, cudaMemcpyHostToDevice));
        CK(cudaMemset(dD, 0xff, (size_t)128 * c.N * 4));
        TestArgs ta{dAhi, dAlo, dBhi, dBlo, dD, c.rowsA, c.N, nch, c.passes, c.swap, 1, dclk};
        const size_t smem = 2 * (size_t)nch * c.rowsA * 16 + 2 * (size_t)nch * c.N * 16 + 1024;
        printf("launch rowsA=%d N=%d K=%d passes=%d swap=%d smem=%zu\n", c.rowsA, c.N, c.K, c.passes, c.swap, smem);
        umma_test<<<1, 192, smem>>>(ta);
        CK(cudaDeviceSynchronize());
        std::vector<float> D((size_t)128 * c.N);
        long long clk[2];
        CK(cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(clk, dclk, sizeof(clk), cudaMemcpyDeviceToHost));
        double max_err_bf = 0, max_err_f32 = 0, max_ref = 0;
        for (int r = 0; r < c.rowsA; r++)
            for (int n = 0; n < c.N; n++) {
                double sb = 0, sf = 0;
                for (int k = 0; k < c.K; k++) {
                    sb += (double)bf16_round(A[(size_t)r * c.K + k]) * bf16_round(B[(size_t)n * c.K + k]);
                    sf += (double)A[(size_t)r * c.K + k] * B[(size_t)n * c.K + k];
                }
                const double got = D[(size_t)r * c.N + n];
                max_err_bf = fmax(max_err_bf, fabs(got - sb));
                max_err_f32 = fmax(max_err_f32, fabs(got - sf));
                max_ref = fmax(max_ref, fabs(sf));
            }
        printf("case rowsA=%d N=%d K=%d passes=%d swap=%d: status %lld, %lld clk, max|D - bf16 product| %.3e, max|D - fp32 product| %.3e (max |ref| %.2f)\n",
               c.rowsA, c.N, c.K, c.passes, c.swap, clk[1], clk[0], max_err_bf, max_err_f32, max_ref);
        cudaFree(dAhi); cudaFree(dAlo); cudaFree(dBhi); cudaFree(dBlo); cudaFree(dD);
    }

    // ---- MMA issue rate: reps x (K = 256 -> 16 k-steps) x passes on resident operands
    {
        const int K = 128, nch = K / 8;
        for (int N : {16, 32, 64, 80, 128, 256}) {
            for (int passes : {1, 3}) {
                std::vector<__nv_bfloat16> Z((size_t)nch * 256 * 8 + 4096, __float2bfloat16(0.f));
                __nv_bfloat16 *dZ;
                float *dD;
                CK(cudaMalloc(&dZ, Z.size() * 2));
                CK(cudaMemcpy(dZ, Z.data(), Z.size() * 2, cudaMemcpyHostToDevice));
                CK(cudaMalloc(&dD, (size_t)128 * N * 4));
                const int reps = 64;
                TestArgs ta{dZ, dZ, dZ, dZ, dD, 96, N, nch, passes, 0, reps, dclk};
                const size_t smem = 2 * (size_t)nch * 96 * 16 + 2 * (size_t)nch * N * 16 + 1024;
                umma_test<<<1, 192, smem>>>(ta);
                CK(cudaDeviceSynchronize());
                long long clk[2];
                CK(cudaMemcpy(clk, dclk, sizeof(clk), cudaMemcpyDeviceToHost));
                const double n_mma = (double)reps * (nch / 2) * passes;
                printf("mma rate M=128 N=%3d passes=%d: %lld clk for %.0f MMAs = %.1f clk/MMA (floor 128*N/256 = %.1f)\n", N, passes,
                       clk[0], n_mma, clk[0] / n_mma, 128.0 * N / 256);
                cudaFree(dZ); cudaFree(dD);
            }
        }
    }

    // ---- L2 -> smem streaming
    {
        const size_t bytes = 16u << 20;
        uint8_t *buf;
        CK(cudaMalloc(&buf, bytes));
        CK(cudaMemset(buf, 1, bytes));
        for (int grid : {1, 8, 32, 74, 148}) {
            for (uint32_t chunk : {2048u, 8192u, 32768u}) {
                const int nring = chunk == 32768u ? 4 : 8;
                const int total = (int)((8u << 20) / chunk);     // 8 MB per CTA
                l2_stream<<<grid, 64, (size_t)nring * chunk>>>(buf, bytes, chunk, nring, total, dclk);   // warm L2
                CK(cudaDeviceSynchronize());
                l2_stream<<<grid, 64, (size_t)nring * chunk>>>(buf, bytes, chunk, nring, total, dclk);
                CK(cudaDeviceSynchronize());
                std::vector<long long> clk(grid);
                CK(cudaMemcpy(clk.data(), dclk, grid * sizeof(long long), cudaMemcpyDeviceToHost));
                long long mx = 0, mn = 1ll << 60;
                for (auto v : clk) { mx = v > mx ? v : mx; mn = v < mn ? v : mn; }
                printf("l2 stream grid=%3d chunk=%5u ring=%d: %.1f .. %.1f B/clk/SM (8 MB per CTA, max %lld clk)\n", grid, chunk, nring,
                       (double)total * chunk / mx, (double)total * chunk / mn, mx);
            }
        }
        cudaFree(buf);
    }


    // ---- MMA chains: dependent vs independent accumulators
    {
        CK(cudaFuncSetAttribute(mma_chain, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
        long long *dout;
        CK(cudaMalloc(&dout, 2 * sizeof(long long)));
        for (int N : {16, 32, 64, 80, 128}) {
            for (int nacc : {1, 2, 3, 4}) {
                if (nacc * N > 512) continue;
                const int nmma = 4096;
                mma_chain<<<1, 64, 64 * 1024>>>(N, nacc, nmma, dout);
                CK(cudaDeviceSynchronize());
                long long o[2];
                CK(cudaMemcpy(o, dout, sizeof(o), cudaMemcpyDeviceToHost));
                printf("mma chain N=%3d accumulators=%d: %.1f clk/MMA to completion, %.1f clk/MMA issue only (floor %.1f)\n", N, nacc,
                       (double)o[0] / nmma, (double)o[1] / nmma, 128.0 * N / 256);
            }
        }
        cudaFree(dout);
    }
    // ---- streaming, several issuing warps
    {
        CK(cudaFuncSetAttribute(l2_stream2, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        const size_t bytes = 16u << 20;
        uint8_t *buf;
        long long *dout;
        CK(cudaMalloc(&buf, bytes));
        CK(cudaMemset(buf, 1, bytes));
        CK(cudaMalloc(&dout, 148 * 8 * 4 * sizeof(long long)));
        struct Cfg { uint32_t chunk; int nring, nissue; };
        const Cfg cfgs[] = {{2048, 8, 1}, {2048, 8, 4}, {2048, 8, 8}, {4096, 4, 1}, {4096, 8, 4}, {8192, 2, 1}, {8192, 4, 1}, {8192, 8, 1}, {8192, 4, 4},
                            {16384, 2, 1}, {16384, 4, 1}, {16384, 4, 2}, {16384, 2, 4}, {32768, 2, 1}, {32768, 4, 1}, {32768, 2, 2}, {49152, 4, 1}, {65536, 3, 1}};
        for (int grid : {1, 148}) {
            for (const Cfg &c : cfgs) {
                const int per_warp = (int)((4u << 20) / c.chunk / c.nissue);
                const size_t smem = (size_t)c.nring * c.chunk * c.nissue;
                if (smem > 200 * 1024) continue;
                for (int rep = 0; rep < 2; rep++) {
                    CK(cudaMemset(dout, 0, 148 * 8 * 4 * sizeof(long long)));
                    l2_stream2<<<grid, 256, smem>>>(buf, bytes, c.chunk, c.nring, per_warp, c.nissue, dout);
                    CK(cudaDeviceSynchronize());
                }
                std::vector<long long> o((size_t)grid * 8 * 4);
                CK(cudaMemcpy(o.data(), dout, o.size() * sizeof(long long), cudaMemcpyDeviceToHost));
                long long mx = 0;
                for (int b = 0; b < grid; b++) mx = o[(size_t)b * 32] > mx ? o[(size_t)b * 32] : mx;
                printf("l2 stream2 grid=%3d chunk=%5u ring=%d issuers=%d: %.1f B/clk/SM, %.0f clk per copy per issuer; isolated copy: issue %lld clk, latency %lld clk; steady issue %lld clk\n",
                       grid, c.chunk, c.nring, c.nissue, (double)per_warp * c.nissue * c.chunk / mx, (double)mx / per_warp, o[1], o[2], o[3]);
            }
        }
        cudaFree(buf);
        cudaFree(dout);
    }
    printf("done\n");
    return 0;
}

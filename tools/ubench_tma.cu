// ubench_tma.cu — how fast can one SM pull a contiguous byte range from HBM into shared memory?
// grid = 148 x ctas CTAs, each streams its own contiguous slice of a 2 GiB buffer (cold HBM) through
// a ring of `depth` slots of `chunk` bytes with cp.async.bulk (UBLKCP) + mbarriers, consumers only
// touch one word per chunk.  Also an LDG.128 variant (registers, no smem).  Development tool; the
// numbers decide the GEMV's stage size / ring depth.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../llama.kotlin_b200/csrc/lk_common.cuh"
using namespace lk;

__global__ void __launch_bounds__(64, 1) tma_stream(const uint8_t *src, size_t bytes_per_cta, int chunk, int depth, int evict_first,
                                                    unsigned long long *out) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t *full = (uint64_t *)smem;          // [depth]
    uint64_t *empty = full + 32;                // [depth]
    uint8_t *ring = smem + 512;
    const int tid = threadIdx.x;
    const uint8_t *base = src + (size_t)blockIdx.x * bytes_per_cta;
    const int n = (int)(bytes_per_cta / chunk);
    if (tid == 0) {
        for (int i = 0; i < depth; i++) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        fence_mbar_init();
    }
    __syncthreads();
    const long long t0 = clock64();
    if (tid == 0) {   // producer
        const uint64_t pol = evict_first ? l2_policy_evict_first() : l2_policy_evict_last();
        int s = 0, round = 0;
        for (int i = 0; i < n; i++) {
            if (round > 0) mbar_wait(&empty[s], (round - 1) & 1);
            mbar_arrive_expect_tx(&full[s], (uint32_t)chunk);
            tma_load_1d(ring + (size_t)s * chunk, base + (size_t)i * chunk, (uint32_t)chunk, &full[s], pol);
            if (++s == depth) { s = 0; round++; }
        }
    } else if (tid == 32) {   // consumer
        int s = 0, par = 0;
        uint32_t acc = 0;
        for (int i = 0; i < n; i++) {
            mbar_wait(&full[s], par);
            acc += *(volatile uint32_t *)(ring + (size_t)s * chunk);
            mbar_arrive(&empty[s]);
            if (++s == depth) { s = 0; par ^= 1; }
        }
        out[blockIdx.x * 2 + 1] = acc;
    }
    __syncthreads();
    if (tid == 0) out[blockIdx.x * 2] = (unsigned long long)(clock64() - t0);
}

__global__ void __launch_bounds__(256, 1) ldg_stream(const uint4 *src, size_t vec_per_cta, int unroll_dummy, unsigned long long *out) {
    const uint4 *base = src + (size_t)blockIdx.x * vec_per_cta;
    uint32_t acc = 0;
    const long long t0 = clock64();
    for (size_t i = threadIdx.x; i + 7 * 256 < vec_per_cta; i += 8 * 256) {
        uint4 v[8];
#pragma unroll
        for (int j = 0; j < 8; j++) asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v[j].x), "=r"(v[j].y), "=r"(v[j].z), "=r"(v[j].w) : "l"(base + i + j * 256));
#pragma unroll
        for (int j = 0; j < 8; j++) acc += v[j].x ^ v[j].y ^ v[j].z ^ v[j].w;
    }
    if (acc == 0x12345) out[blockIdx.x * 2 + 1] = acc;
    __syncthreads();
    if (threadIdx.x == 0) out[blockIdx.x * 2] = (unsigned long long)(clock64() - t0);
}

int main() {
    const size_t total = (size_t)2 << 30;
    uint8_t *buf; unsigned long long *out;
    cudaMalloc(&buf, total); cudaMemset(buf, 1, total);
    cudaMalloc(&out, 148 * 4 * 16);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaFuncSetAttribute(tma_stream, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    const int chunks[] = {2048, 4096, 8192, 16384, 32768, 65536};
    for (int ctas = 1; ctas <= 2; ctas++)
        for (int chunk : chunks)
            for (int inflight_kb : {16, 32, 64, 96, 192}) {
                const int depth = inflight_kb * 1024 / chunk;
                if (depth < 1 || depth > 32 || inflight_kb * ctas > 200) continue;
                const int grid = 148 * ctas;
                const size_t per_cta = (total / 4 / grid / chunk) * chunk;   // 512 MiB per run, rotate the offset
                static int rot = 0;
                const uint8_t *src = buf + (size_t)(rot++ % 4) * (total / 4);
                const int smem = 512 + depth * chunk;
                cudaEventRecord(e0);
                tma_stream<<<grid, 64, smem>>>(src, per_cta, chunk, depth, 1, out);
                cudaEventRecord(e1);
                cudaError_t e = cudaDeviceSynchronize();
                float ms; cudaEventElapsedTime(&ms, e0, e1);
                printf("TMA  ctas/SM %d chunk %6d depth %2d (%3d KB in flight/CTA): %7.1f GB/s  %s\n", ctas, chunk, depth, inflight_kb,
                       (double)per_cta * grid / ms / 1e6, e == cudaSuccess ? "" : cudaGetErrorString(e));
            }
    for (int ctas = 1; ctas <= 4; ctas *= 2) {
        const int grid = 148 * ctas;
        const size_t vec_per_cta = total / 4 / grid / 16;
        cudaEventRecord(e0);
        ldg_stream<<<grid, 256>>>((const uint4 *)buf, vec_per_cta, 0, out);
        cudaEventRecord(e1);
        cudaDeviceSynchronize();
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        printf("LDG.128 x8 unroll, 256 thr, ctas/SM %d: %7.1f GB/s\n", ctas, (double)vec_per_cta * 16 * grid / ms / 1e6);
    }
    return 0;
}

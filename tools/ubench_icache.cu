// ubench_icache.cu — is straight-line code cold at every kernel launch?  A kernel body of
// NI unrolled FFMAs (16 B each) is executed `passes` times inside one launch; we time every
// pass with clock64, for several back-to-back launches.  Development tool.
#include <cstdio>
#include <cuda_runtime.h>

template <int NI>
__global__ void __launch_bounds__(256, 1) k(float *out, long long *clk, int passes) {
    float a[8];
#pragma unroll
    for (int i = 0; i < 8; i++) a[i] = threadIdx.x * 0.001f + i;
    for (int p = 0; p < passes; p++) {
        const long long t0 = clock64();
#pragma unroll
        for (int i = 0; i < NI; i++) a[i & 7] = fmaf(a[i & 7], 1.0001f, 0.5f);
        const long long t1 = clock64();
        if (threadIdx.x == 0) clk[blockIdx.x * 8 + p] = t1 - t0;
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int NI>
void run(float *out, long long *clk) {
    for (int l = 0; l < 3; l++) {
        k<NI><<<148, 256>>>(out, clk, 3);
        cudaDeviceSynchronize();
        long long h[8];
        cudaMemcpy(h, clk, 64, cudaMemcpyDeviceToHost);
        printf("NI %5d (%3d KB) launch %d: pass cycles %lld %lld %lld  (ideal %d at 2 warps/SMSP)\n", NI, NI * 16 / 1024, l, h[0], h[1], h[2], NI * 2);
    }
}

int main() {
    float *out; long long *clk;
    cudaMalloc(&out, 148 * 256 * 4); cudaMalloc(&clk, 148 * 64);
    run<256>(out, clk); run<1024>(out, clk); run<2048>(out, clk); run<4096>(out, clk);
    // interleave two different kernels
    for (int l = 0; l < 2; l++) { run<1024>(out, clk); run<2048>(out, clk); }
    return 0;
}

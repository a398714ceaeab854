// ubench_imma.cu — issue rate and latency of the legacy warp-level int8 MMA (mma.sync.m16n8k32.s8 -> IMMA.16832) on sm_100a,
// the instruction the MMA form of the decode GEMV would use for its plane sums.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 tools/ubench_imma.cu -o tools/ubench_imma
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); return 1; } } while (0)

__device__ __forceinline__ void imma(int (&c)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

template <int ILP>
__global__ void k_imma(int iters, int *out, long long *cyc) {
    uint32_t a[4] = {threadIdx.x * 0x01010101u, 0x02030405u, 0x0a0b0c0du, threadIdx.x}, b[2] = {0x01ff02feu, threadIdx.x * 3u};
    int c[ILP][4] = {};
    __syncthreads();
    const long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int j = 0; j < ILP; j++) imma(c[j], a, b);
    }
    const long long t1 = clock64();
    int s = 0;
#pragma unroll
    for (int j = 0; j < ILP; j++) s += c[j][0] + c[j][1] + c[j][2] + c[j][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int ILP>
int run(int warps, int iters, int *out, long long *cyc) {
    k_imma<ILP><<<148, warps * 32>>>(iters, out, cyc);
    CK(cudaDeviceSynchronize());
    k_imma<ILP><<<148, warps * 32>>>(iters, out, cyc);
    CK(cudaDeviceSynchronize());
    long long h[148]; CK(cudaMemcpy(h, cyc, sizeof h, cudaMemcpyDeviceToHost));
    double m = 0; for (int i = 0; i < 148; i++) m += (double)h[i]; m /= 148;
    const double per_sm = (double)warps * iters * ILP / m;
    printf("warps/SM %2d  ILP %d: %.1f clk per loop pass, %.3f MMA/clk/SM (%.3f per SMSP) = %.0f dense int8 TOPS at 1.9 GHz over 148 SMs\n", warps, ILP, m / iters, per_sm,
           per_sm / 4, per_sm * 16 * 8 * 32 * 2 * 1.9e9 * 148 / 1e12);
    return 0;
}

int main() {
    int *out; long long *cyc;
    CK(cudaMalloc(&out, 148 * 1024 * 4)); CK(cudaMalloc(&cyc, 148 * 8));
    const int iters = 4096;
    run<1>(1, iters, out, cyc);      // latency of a dependent chain
    run<4>(1, iters, out, cyc);
    run<1>(4, iters, out, cyc);
    run<4>(4, iters, out, cyc);
    run<2>(8, iters, out, cyc);
    run<4>(8, iters, out, cyc);
    run<4>(16, iters, out, cyc);
    return 0;
}

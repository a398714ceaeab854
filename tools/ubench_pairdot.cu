// ubench_pairdot.cu — compute-only microbenchmark of the GEMV inner product (pair_dot from
// lk_gemv.cu) on shared-memory resident weights: cycles per warp-level pair-dot per SMSP as a
// function of the number of warps per SM.  Development tool (numbers quoted in DESIGN.md).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I llama.kotlin_b200/csrc tools/ubench_pairdot.cu -o tools/ubench_pairdot
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "lk_gemv.cu"

using namespace lk;

template <int TYPE, int UNROLL>
__global__ void __launch_bounds__(512, 1) k(int rows, int reps, int rg, float *out, long long *clk) {
    extern __shared__ __align__(128) uint8_t smem[];
    constexpr int PW = PairWords<TYPE>::value;
    constexpr int NP = 64;                      // K = 4096
    const int row_bytes = NP * PW * 4;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    float *red = (float *)smem;                 // [rows][64]
    uint32_t *w = (uint32_t *)(smem + rows * NP * 4);
    for (int i = tid; i < rows * NP * PW; i += blockDim.x) w[i] = (i * 2654435761u) ^ (i >> 3);
    XRegs x;
#pragma unroll
    for (int pl = 0; pl < 3; pl++)
#pragma unroll
        for (int b = 0; b < 2; b++)
#pragma unroll
            for (int j = 0; j < 8; j++) x.q[pl][b][j] = (tid * 31 + pl * 7 + b * 3 + j) * 2246822519u;
    x.s = 1e-3f; x.c0 = 0.5f; x.c1 = 0.25f;
    const int cw = warp & 1, rgi = warp >> 1;
    const int col = cw * 32 + lane;
    __syncthreads();
    const long long t0 = clock64();
    for (int r = 0; r < reps; r++) {
#pragma unroll UNROLL
        for (int row = rgi; row < rows; row += rg) {
            const float v = pair_dot<TYPE>((const uint32_t *)((const uint8_t *)w + (size_t)row * row_bytes) + col * PW, x);
            red[row * NP + col] = v;
        }
    }
    const long long t1 = clock64();
    __syncthreads();
    if (tid == 0) clk[blockIdx.x] = t1 - t0;
    out[blockIdx.x * blockDim.x + tid] = red[(tid % rows) * NP + (tid & 63)];
}

template <int TYPE, int UNROLL>
void run(const char *name, int warps, float *out, long long *clk) {
    const int rg = warps / 2, rows = 24, reps = 40;
    const int smem = rows * 64 * 4 + rows * 64 * PairWords<TYPE>::value * 4;
    cudaFuncSetAttribute(k<TYPE, UNROLL>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    k<TYPE, UNROLL><<<148, warps * 32, smem>>>(rows, reps, rg, out, clk);
    cudaDeviceSynchronize();
    k<TYPE, UNROLL><<<148, warps * 32, smem>>>(rows, reps, rg, out, clk);
    cudaError_t e = cudaDeviceSynchronize();
    long long h;
    cudaMemcpy(&h, clk, 8, cudaMemcpyDeviceToHost);
    const double wpd = (double)rows * 2 * reps;       // warp-level pair-dots per SM
    printf("%-6s unroll %d warps %2d: %8lld clk, %.1f clk per warp-pair-dot per SMSP (%.2f clk per 8-element word)  %s\n", name, UNROLL, warps, h,
           (double)h / (wpd / 4), (double)h / (wpd / 4) / 8, cudaGetErrorString(e));
}

int main() {
    float *out; long long *clk;
    cudaMalloc(&out, 148 * 512 * 4); cudaMalloc(&clk, 148 * 8);
    for (int w : {4, 8, 12, 16}) { run<T_Q4_0, 1>("Q4_0", w, out, clk); run<T_Q4_0, 2>("Q4_0", w, out, clk); }
    for (int w : {8, 16}) { run<T_Q4_1, 1>("Q4_1", w, out, clk); run<T_Q8_0, 1>("Q8_0", w, out, clk); }
    return 0;
}

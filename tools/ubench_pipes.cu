// ubench_pipes.cu — issue-rate microbenchmark of the SASS pipes the GEMV inner loop uses
// (development tool; results recorded in DESIGN.md).  For each op: 8 independent chains per
// thread, 1024 threads per CTA (8 warps per SMSP), one CTA per SM; reports warp-instructions
// per clock per SMSP (1.0 = full issue rate, 0.5 = a 16-lane pipe).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 tools/ubench_pipes.cu -o tools/ubench_pipes
#include <cstdio>
#include <cstdlib>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#define ITERS 512
#define CH 8

enum Op { FFMA, FFMA2, PRMT, LOP3, DP4A, CVT_H2F, HFMA2, I2F, IMAD, SHF, MIX_PRMT_FFMA2, MIX_DP4A_LOP3, MIX_DP4A_FFMA,
          MIX_PRMT_DP4A, FADD2, MIX_CVT_LOP3, I2F_S8, MIX_LOP3_FFMA, FMUL2, MIX_PRMT_FFMA, MIX_HFMA2_LOP3, NOPS };
const char *names[] = {"FFMA", "FFMA2(f32x2)", "PRMT", "LOP3", "DP4A", "CVT.F32.F16", "HFMA2", "I2F.S32", "IMAD", "SHF",
                       "PRMT+FFMA2 1:1", "DP4A+LOP3 1:1", "DP4A+FFMA 1:1", "PRMT+DP4A 1:1", "FADD2(f32x2)", "CVT.F32.F16+LOP3 1:1",
                       "I2F.S8(byte)", "LOP3+FFMA 1:1", "FMUL2", "PRMT+FFMA 1:1", "HFMA2+LOP3 1:1"};
const int per_iter[] = {1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 2, 2, 2, 2, 1, 2, 1, 2, 1, 2, 2};

template <int OP>
__global__ void __launch_bounds__(1024, 1) k(unsigned long long *out, uint32_t seed, float fs) {
    uint32_t r[CH], q[CH];
    float f[CH], g[CH];
    unsigned long long d[CH];
#pragma unroll
    for (int i = 0; i < CH; i++) {
        r[i] = seed + threadIdx.x * 7 + i; q[i] = seed * 3 + i; f[i] = fs + i; g[i] = fs * 0.5f + i;
        d[i] = ((unsigned long long)__float_as_uint(f[i]) << 32) | __float_as_uint(g[i]);
    }
    const unsigned long long dm = ((unsigned long long)__float_as_uint(fs) << 32) | __float_as_uint(fs);
    __syncthreads();
    const long long t0 = clock64();
#pragma unroll 4
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < CH; i++) {
            if (OP == FFMA || OP == MIX_DP4A_FFMA || OP == MIX_LOP3_FFMA || OP == MIX_PRMT_FFMA) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f[i]) : "f"(fs), "f"(g[i]));
            if (OP == FFMA2 || OP == MIX_PRMT_FFMA2) asm volatile("fma.rn.f32x2 %0, %0, %1, %1;" : "+l"(d[i]) : "l"(dm));
            if (OP == FADD2) asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(d[i]) : "l"(dm));
            if (OP == FMUL2) asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(d[i]) : "l"(dm));
            if (OP == PRMT || OP == MIX_PRMT_FFMA2 || OP == MIX_PRMT_DP4A || OP == MIX_PRMT_FFMA) asm volatile("prmt.b32 %0, %0, %1, 0x7044;" : "+r"(r[i]) : "r"(seed));
            if (OP == LOP3 || OP == MIX_DP4A_LOP3 || OP == MIX_CVT_LOP3 || OP == MIX_LOP3_FFMA || OP == MIX_HFMA2_LOP3) asm volatile("lop3.b32 %0, %0, %1, %2, 0xEA;" : "+r"(q[i]) : "r"(seed), "r"(0x64006400));
            if (OP == DP4A || OP == MIX_DP4A_LOP3 || OP == MIX_DP4A_FFMA || OP == MIX_PRMT_DP4A) asm volatile("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(r[i]) : "r"(q[i]), "r"(seed));
            if (OP == CVT_H2F || OP == MIX_CVT_LOP3) { unsigned short h = (unsigned short)r[i]; asm volatile("cvt.f32.f16 %0, %1;" : "=f"(f[i]) : "h"(h)); r[i] = __float_as_uint(f[i]) + i; }
            if (OP == HFMA2 || OP == MIX_HFMA2_LOP3) asm volatile("fma.rn.f16x2 %0, %0, %1, %1;" : "+r"(r[i]) : "r"(seed));
            if (OP == I2F) { asm volatile("cvt.rn.f32.s32 %0, %1;" : "=f"(f[i]) : "r"(r[i])); r[i] = __float_as_uint(f[i]); }
            if (OP == I2F_S8) { asm volatile("{ .reg .b8 b0, b1, b2, b3; mov.b32 {b0, b1, b2, b3}, %1; cvt.rn.f32.s8 %0, b1; }" : "=f"(f[i]) : "r"(r[i])); r[i] = __float_as_uint(f[i]); }
            if (OP == IMAD) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(r[i]) : "r"(seed), "r"(q[i]));
            if (OP == SHF) asm volatile("shf.r.wrap.b32 %0, %0, %1, 5;" : "+r"(r[i]) : "r"(q[i]));
        }
    }
    const long long t1 = clock64();
    uint32_t acc = 0;
#pragma unroll
    for (int i = 0; i < CH; i++) acc += r[i] + q[i] + __float_as_uint(f[i]) + (uint32_t)d[i] + (uint32_t)(d[i] >> 32);
    if (threadIdx.x == 0) out[blockIdx.x * 2] = (unsigned long long)(t1 - t0);
    if (acc == 0x12345678u) out[blockIdx.x * 2 + 1] = acc;
}

template <int OP>
void run(unsigned long long *dout) {
    k<OP><<<148, 1024>>>(dout, 12345u, 1.0001f);
    cudaDeviceSynchronize();
    k<OP><<<148, 1024>>>(dout, 12345u, 1.0001f);
    cudaError_t e = cudaDeviceSynchronize();
    unsigned long long h[2];
    cudaMemcpy(h, dout, 16, cudaMemcpyDeviceToHost);
    const double winstr = (double)ITERS * CH * per_iter[OP] * 8;  // per SMSP: 8 warps
    printf("%-22s %8llu clk  %.3f warp-instr/clk/SMSP  (%s)\n", names[OP], h[0], winstr / (double)h[0], cudaGetErrorString(e));
}

int main() {
    unsigned long long *dout;
    cudaMalloc(&dout, 148 * 16);
    run<FFMA>(dout); run<FFMA2>(dout); run<FADD2>(dout); run<FMUL2>(dout); run<PRMT>(dout); run<LOP3>(dout); run<DP4A>(dout);
    run<CVT_H2F>(dout); run<HFMA2>(dout); run<I2F>(dout); run<I2F_S8>(dout); run<IMAD>(dout); run<SHF>(dout);
    run<MIX_PRMT_FFMA2>(dout); run<MIX_PRMT_FFMA>(dout); run<MIX_DP4A_LOP3>(dout); run<MIX_DP4A_FFMA>(dout); run<MIX_PRMT_DP4A>(dout);
    run<MIX_CVT_LOP3>(dout); run<MIX_LOP3_FFMA>(dout); run<MIX_HFMA2_LOP3>(dout);
    return 0;
}

// lk_elementwise.cu — K0 block dequantisation, the generic (any-shape) matmul,
// and the F32 ops around the matmul: RMSNorm (K3), SiLU / SiLU*up (K4),
// softmax (K5), add / mul (R4).  All HBM-bound streaming kernels.
//
// Reference paths relative to
//   /root/reference/src/nativeMain/kotlin/ai/solace/llamakotlin/
#include "lk_common.cuh"
#include "lk_kernels.h"

namespace lk {

static inline int grid_for(int64_t work_items, int threads, int max_blocks = 148 * 16) {
    int64_t g = (work_items + threads - 1) / threads;
    if (g < 1) g = 1;
    if (g > max_blocks) g = max_blocks;
    return (int)g;
}

// ---------------------------------------------------------------------------
// K0 dequantize: y = d*q | d*(nib-8) | d*nib + m, each F32 op rounded once
// (__fmul_rn/__fadd_rn: no FMA contraction) => bit-exact with
// dequantizeTensor, core/GGMLComputeOps.kt:932-964.  Nibble order is the
// reference's: element e -> (qs[e/2] >> 4*(e&1)) & 15 (core/GGMLTypes.kt:647-651).
// Each thread produces 8 consecutive elements (two coalesced 16-byte stores);
// the packed bytes are read with 2-byte loads (blocks are only 2-byte aligned).
// Algorithmic bytes per element: 4 (store) + blockBytes/32 (load).
// ---------------------------------------------------------------------------
template <int TYPE>
__global__ void __launch_bounds__(256) dequant_kernel(const uint8_t *__restrict__ src, int64_t n,
                                                      float *__restrict__ dst, int nib) {
    constexpr int BB = block_bytes(TYPE);
    const int64_t ngroups = n / 8;  // n % 32 == 0
    for (int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; g < ngroups;
         g += (int64_t)gridDim.x * blockDim.x) {
        const int64_t blk = g >> 2;
        const int sub = (int)(g & 3);  // which 8 elements of the block
        const uint8_t *p = src + blk * BB;
        const float d = h2f(*(const uint16_t *)p);
        float v[8];
        if (TYPE == T_Q8_0) {
            const uint16_t *q = (const uint16_t *)(p + 2 + sub * 8);
#pragma unroll
            for (int i = 0; i < 4; i++) {
                uint16_t w = q[i];
                v[2 * i] = __fmul_rn(d, (float)(int8_t)(w & 0xFF));
                v[2 * i + 1] = __fmul_rn(d, (float)(int8_t)(w >> 8));
            }
        } else {
            constexpr int QOFF = (TYPE == T_Q4_0) ? 2 : 4;
            uint32_t w;
            if (nib) {   // ggml order: elements 8*sub..+7 are the low (sub < 2) / high nibbles of bytes 8*(sub&1)..+7
                const uint16_t *q = (const uint16_t *)(p + QOFF + (sub & 1) * 8);
                const int sh = (sub >> 1) * 4;
                w = 0;
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    const uint32_t two = q[i];
                    w |= (((two >> sh) & 15u) | (((two >> (8 + sh)) & 15u) << 4)) << (8 * i);
                }
            } else {
                const uint16_t *q = (const uint16_t *)(p + QOFF + sub * 4);
                w = (uint32_t)q[0] | ((uint32_t)q[1] << 16);
            }
            float m = 0.f;
            if (TYPE == T_Q4_1) m = h2f(*(const uint16_t *)(p + 2));
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const float nib = (float)((w >> (4 * i)) & 15u);
                if (TYPE == T_Q4_0)
                    v[i] = __fmul_rn(d, __fsub_rn(nib, 8.0f));
                else
                    v[i] = __fadd_rn(__fmul_rn(d, nib), m);
            }
        }
        float4 *o = (float4 *)(dst + g * 8);
        o[0] = make_float4(v[0], v[1], v[2], v[3]);
        o[1] = make_float4(v[4], v[5], v[6], v[7]);
    }
}

cudaError_t launch_dequantize(int type, const void *src, int64_t n, float *dst, cudaStream_t s, int nib) {
    if (n == 0) return cudaSuccess;
    const int threads = 256;
    const int grid = grid_for(n / 8, threads);
    switch (type) {
        case T_Q8_0: dequant_kernel<T_Q8_0><<<grid, threads, 0, s>>>((const uint8_t *)src, n, dst, nib); break;
        case T_Q4_0: dequant_kernel<T_Q4_0><<<grid, threads, 0, s>>>((const uint8_t *)src, n, dst, nib); break;
        case T_Q4_1: dequant_kernel<T_Q4_1><<<grid, threads, 0, s>>>((const uint8_t *)src, n, dst, nib); break;
        default: return cudaErrorNotSupported;
    }
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------
// Generic matmul: any K, M, N; a in {F32, Q4_0, Q4_1, Q8_0}; reference layouts
// (b element (j,k) at j + k*N, dst element (j,i) at j + i*N).  One warp per
// (row i, tile of up to 8 columns); lanes stride over k element-wise with the
// reference's flat block addressing (flat = i*K + k, block = flat/32,
// core/GGMLComputeOps.kt:58-62) so K need not be a multiple of 32.
// The dequantised weight uses the reference's exact per-element formula; only
// the summation order differs (lane-strided + shuffle tree), which is why the
// matmul bar is a tolerance, not bit-exactness.
// This is the correctness fallback, not a tuned kernel.
// ---------------------------------------------------------------------------
template <int TYPE>
__device__ __forceinline__ float load_weight(const void *a, int64_t flat, int nib) {
    if (TYPE == T_F32) return ((const float *)a)[flat];
    constexpr int BB = block_bytes(TYPE);
    const uint8_t *p = (const uint8_t *)a + (flat >> 5) * BB;
    const int item = (int)(flat & 31);
    const float d = h2f(*(const uint16_t *)p);
    if (TYPE == T_Q8_0) return __fmul_rn(d, (float)(int8_t)p[2 + item]);
    constexpr int QOFF = (TYPE == T_Q4_0) ? 2 : 4;
    const uint8_t byte = p[QOFF + (nib ? (item & 15) : (item >> 1))];      // nib: ggml order (byte j = elements j, j+16)
    const float q = (float)((nib ? (item >> 4) : (item & 1)) ? (byte >> 4) : (byte & 15));
    if (TYPE == T_Q4_0) return __fmul_rn(d, __fsub_rn(q, 8.0f));
    return __fadd_rn(__fmul_rn(d, q), h2f(*(const uint16_t *)(p + 2)));
}

template <int TYPE>
__global__ void __launch_bounds__(128) matmul_generic_kernel(const void *__restrict__ a, int64_t K, int64_t M,
                                                             const float *__restrict__ b, int64_t N,
                                                             float *__restrict__ dst, int64_t ntiles_n, int nib) {
    constexpr int TN = 8;
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t w = warp; w < M * ntiles_n; w += nwarps) {
        const int64_t i = w / ntiles_n;
        const int64_t j0 = (w % ntiles_n) * TN;
        const int nj = (int)((N - j0) < TN ? (N - j0) : TN);
        float acc[TN];
#pragma unroll
        for (int j = 0; j < TN; j++) acc[j] = 0.f;
        for (int64_t k = lane; k < K; k += 32) {
            const float wv = load_weight<TYPE>(a, i * K + k, nib);
            const float *bp = b + j0 + k * N;
#pragma unroll
            for (int j = 0; j < TN; j++)
                if (j < nj) acc[j] = __fadd_rn(acc[j], __fmul_rn(wv, bp[j]));
        }
#pragma unroll
        for (int j = 0; j < TN; j++) acc[j] = warp_sum(acc[j]);
        if (lane == 0)
            for (int j = 0; j < nj; j++) dst[j0 + j + i * N] = acc[j];
    }
}

cudaError_t launch_matmul_generic(int type, const void *a, int64_t K, int64_t M, const float *b, int64_t N,
                                  float *dst, cudaStream_t s, int nib) {
    if (M == 0 || N == 0) return cudaSuccess;
    const int64_t ntn = (N + 7) / 8;
    const int threads = 128;
    const int grid = grid_for(M * ntn * 32, threads, 148 * 32);
    switch (type) {
        case T_F32: matmul_generic_kernel<T_F32><<<grid, threads, 0, s>>>(a, K, M, b, N, dst, ntn, nib); break;
        case T_Q8_0: matmul_generic_kernel<T_Q8_0><<<grid, threads, 0, s>>>(a, K, M, b, N, dst, ntn, nib); break;
        case T_Q4_0: matmul_generic_kernel<T_Q4_0><<<grid, threads, 0, s>>>(a, K, M, b, N, dst, ntn, nib); break;
        case T_Q4_1: matmul_generic_kernel<T_Q4_1><<<grid, threads, 0, s>>>(a, K, M, b, N, dst, ntn, nib); break;
        default: return cudaErrorNotSupported;
    }
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------
// R4 add / mul, K4 SiLU and SiLU*up: one pass, 16-byte vector accesses.
// Algorithmic bytes per element: 12 (binary ops, silu_mul), 8 (silu).
// ---------------------------------------------------------------------------
enum { OP_ADD = 0, OP_MUL = 1, OP_SILU = 2, OP_SILU_MUL = 3, OP_SCALE = 4 };

// silu(x) = x * (1 / (1 + exp(-x))) — nativeTest core/GGMLComputeOpsTest.kt:491-499.
// expf (not __expf) and IEEE division keep it within 2 ulp of the libm path.
__device__ __forceinline__ float silu1(float x) { return __fmul_rn(x, __fdiv_rn(1.0f, __fadd_rn(1.0f, expf(-x)))); }

template <int OP>
__device__ __forceinline__ float ew_apply(float a, float b) {
    if (OP == OP_ADD) return __fadd_rn(a, b);
    if (OP == OP_MUL) return __fmul_rn(a, b);
    if (OP == OP_SILU) return silu1(a);
    if (OP == OP_SCALE) return __fmul_rn(a, b);     // LlamaAttention.scale: value * factor (model/LlamaAttention.kt:190-191)
    return __fmul_rn(silu1(a), b);
}

template <int OP>
__global__ void __launch_bounds__(256) ew_kernel(const float *__restrict__ a, const float *__restrict__ b, int64_t n,
                                                 float *__restrict__ dst, int vec_ok, float p0) {
    constexpr bool BIN = (OP != OP_SILU && OP != OP_SCALE);
    pdl_wait();                // inputs may be the previous kernel's output
    pdl_launch_dependents();   // the next kernel may start its prologue (a GEMV: its weight prefetch) under this one
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t nthr = (int64_t)gridDim.x * blockDim.x;
    int64_t done = 0;
    if (vec_ok) {
        const int64_t n4 = n / 4;
        for (int64_t i = tid; i < n4; i += nthr) {
            const float4 x = ((const float4 *)a)[i];
            float4 y = make_float4(p0, p0, p0, p0);
            if (BIN) y = ((const float4 *)b)[i];
            float4 r;
            r.x = ew_apply<OP>(x.x, y.x);
            r.y = ew_apply<OP>(x.y, y.y);
            r.z = ew_apply<OP>(x.z, y.z);
            r.w = ew_apply<OP>(x.w, y.w);
            ((float4 *)dst)[i] = r;
        }
        done = n4 * 4;
    }
    for (int64_t i = done + tid; i < n; i += nthr) dst[i] = ew_apply<OP>(a[i], BIN ? b[i] : p0);
}

template <int OP>
static cudaError_t launch_ew(const float *a, const float *b, int64_t n, float *dst, cudaStream_t s, bool pdl, float p0 = 0.f) {
    if (n == 0) return cudaSuccess;
    const uintptr_t al = (uintptr_t)a | (uintptr_t)dst | (uintptr_t)(b ? b : a);
    const int vec_ok = (al & 15) == 0;
    const int threads = 256;
    return launch_kernel(ew_kernel<OP>, dim3(grid_for((n + 3) / 4, threads)), dim3(threads), 0, s, pdl, a, b, n, dst, vec_ok, p0);
}
cudaError_t launch_add(const float *a, const float *b, int64_t n, float *d, cudaStream_t s, bool pdl) { return launch_ew<OP_ADD>(a, b, n, d, s, pdl); }
cudaError_t launch_mul(const float *a, const float *b, int64_t n, float *d, cudaStream_t s, bool pdl) { return launch_ew<OP_MUL>(a, b, n, d, s, pdl); }
cudaError_t launch_silu(const float *x, int64_t n, float *d, cudaStream_t s, bool pdl) { return launch_ew<OP_SILU>(x, nullptr, n, d, s, pdl); }
cudaError_t launch_scale(const float *x, float f, int64_t n, float *d, cudaStream_t s, bool pdl) { return launch_ew<OP_SCALE>(x, nullptr, n, d, s, pdl, f); }
cudaError_t launch_silu_mul(const float *g, const float *u, int64_t n, float *d, cudaStream_t s, bool pdl) { return launch_ew<OP_SILU_MUL>(g, u, n, d, s, pdl); }

// ---------------------------------------------------------------------------
// block-wide reductions (256 threads)
// ---------------------------------------------------------------------------
template <typename T, typename F>
__device__ __forceinline__ T block_reduce(T v, F op, T *scratch /* >= 32 */) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = op(v, __shfl_xor_sync(0xffffffffu, v, o));
    __syncthreads();  // scratch reuse across calls
    if (lane == 0) scratch[wid] = v;
    __syncthreads();
    T r = scratch[0];
    for (int w = 1; w < nw; w++) r = op(r, scratch[w]);
    return r;
}

// ---------------------------------------------------------------------------
// K3 RMSNorm — spec: manualRMSNorm, nativeTest core/GGMLComputeOpsTest.kt:502-521:
// sum of squares accumulated in F64, rms = (float)sqrt(mean + eps), y = x / rms
// (IEEE division), optional * weight (model/LlamaModel.kt:59-62).
// B200's FP64 pipe is vestigial (16 DFMA + 16 F2F per thread made this a 13 us kernel for 4096 elements,
// profiles/r01i), so each thread accumulates its share as an unevaluated float pair with error-free
// transformations (x*x = p + e exactly by FMA; TwoSum for the running sum): ~2^-45 relative, i.e. the
// F64 result to within a few ulp of F64; warps combine their pairs the same way through shuffles and
// a single thread does the remaining F64 work (sum of <= 16 warp totals, divide, sqrt).  The result differs from the oracle only where (float)sqrt sits on a rounding
// boundary.  One CTA per row; bytes per element 8 (+4 weight).
// ---------------------------------------------------------------------------
// float-pair (hi, lo) += float-pair, error-free up to second order (TwoSum on the high parts)
__device__ __forceinline__ void pair_add(float &s, float &c, float s2, float c2) {
    const float t = __fadd_rn(s, s2), bp = __fsub_rn(t, s);
    const float err = __fadd_rn(__fsub_rn(s, __fsub_rn(t, bp)), __fsub_rn(s2, bp));
    s = t;
    c = __fadd_rn(__fadd_rn(c, c2), err);
}

__global__ void __launch_bounds__(512) rms_norm_kernel(const float *__restrict__ x, const float *__restrict__ w,
                                                       float eps, int64_t ne0, float *__restrict__ dst) {
    __shared__ float2 part[16];
    __shared__ float rms_s;
    pdl_wait();                // inputs may be the previous kernel's output
    pdl_launch_dependents();   // the next kernel may start its prologue (a GEMV: its weight prefetch) under this one
    const float *xr = x + (int64_t)blockIdx.x * ne0;
    float *dr = dst + (int64_t)blockIdx.x * ne0;
    float s = 0.f, c = 0.f;
    for (int64_t i = threadIdx.x; i < ne0; i += blockDim.x) {
        const float v = xr[i];
        const float p = __fmul_rn(v, v), e = __fmaf_rn(v, v, -p);          // v*v = p + e exactly
        pair_add(s, c, p, e);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) pair_add(s, c, __shfl_xor_sync(0xffffffffu, s, o), __shfl_xor_sync(0xffffffffu, c, o));
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = make_float2(s, c);
    __syncthreads();
    if (threadIdx.x == 0) {    // the only F64 arithmetic of the kernel: <= 32 DADD, one divide, one sqrt, one thread
        double sumsq = 0.0;
        for (int wi = 0; wi < (int)(blockDim.x >> 5); wi++) sumsq += (double)part[wi].x + (double)part[wi].y;
        rms_s = (float)sqrt(sumsq / (double)ne0 + (double)eps);
    }
    __syncthreads();
    const float rms = rms_s;
    const bool zero = (rms == 0.0f) || (rms != rms);
    for (int64_t i = threadIdx.x; i < ne0; i += blockDim.x) {
        float v = zero ? 0.0f : __fdiv_rn(xr[i], rms);
        if (w != nullptr && !zero) v = __fmul_rn(v, w[i]);
        dr[i] = v;
    }
}

// Rows of up to 16 floats per thread (ne0 <= 8192 at 512 threads, 16-byte aligned): ONE pass over HBM — the row is loaded
// with 16-byte loads into registers, squared and summed from there, scaled and stored (round 1's kernel read every row
// twice with 4-byte loads: 0.46 of the measured HBM rate on 268 MB tensors).  Same arithmetic: float-pair accumulation of the
// squares, one thread does the F64 tail, IEEE division per element.
template <int V>      // float4 chunks per thread
__global__ void __launch_bounds__(512) rms_norm_row_kernel(const float *__restrict__ x, const float *__restrict__ w, float eps,
                                                           int64_t ne0, float *__restrict__ dst) {
    __shared__ float2 part[16];
    __shared__ float rms_s;
    pdl_wait();
    pdl_launch_dependents();
    const float4 *xr = (const float4 *)(x + (int64_t)blockIdx.x * ne0);
    float4 *dr = (float4 *)(dst + (int64_t)blockIdx.x * ne0);
    const int n4 = (int)(ne0 >> 2);
    float4 v[V];
    float s = 0.f, c = 0.f;
#pragma unroll
    for (int j = 0; j < V; j++) {
        const int i = threadIdx.x + j * blockDim.x;
        v[j] = i < n4 ? xr[i] : make_float4(0.f, 0.f, 0.f, 0.f);
        const float e4[4] = {v[j].x, v[j].y, v[j].z, v[j].w};
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const float pq = __fmul_rn(e4[q], e4[q]), eq = __fmaf_rn(e4[q], e4[q], -pq);
            pair_add(s, c, pq, eq);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) pair_add(s, c, __shfl_xor_sync(0xffffffffu, s, o), __shfl_xor_sync(0xffffffffu, c, o));
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = make_float2(s, c);
    __syncthreads();
    if (threadIdx.x == 0) {
        double sumsq = 0.0;
        for (int wi = 0; wi < (int)(blockDim.x >> 5); wi++) sumsq += (double)part[wi].x + (double)part[wi].y;
        rms_s = (float)sqrt(sumsq / (double)ne0 + (double)eps);
    }
    __syncthreads();
    const float rms = rms_s;
    const bool zero = (rms == 0.0f) || (rms != rms);
#pragma unroll
    for (int j = 0; j < V; j++) {
        const int i = threadIdx.x + j * blockDim.x;
        if (i >= n4) continue;
        float4 r = make_float4(0.f, 0.f, 0.f, 0.f);
        if (!zero) {
            r = make_float4(__fdiv_rn(v[j].x, rms), __fdiv_rn(v[j].y, rms), __fdiv_rn(v[j].z, rms), __fdiv_rn(v[j].w, rms));
            if (w != nullptr) {
                const float4 g = ((const float4 *)w)[i];
                r = make_float4(__fmul_rn(r.x, g.x), __fmul_rn(r.y, g.y), __fmul_rn(r.z, g.z), __fmul_rn(r.w, g.w));
            }
        }
        dr[i] = r;
    }
}

cudaError_t launch_rms_norm(const float *x, const float *w, float eps, int64_t ne0, int64_t nrows, float *dst,
                            cudaStream_t s, bool pdl) {
    if (nrows == 0 || ne0 == 0) return cudaSuccess;
    const bool vec = (ne0 % 4) == 0 && ne0 <= 8192 && ((((uintptr_t)x | (uintptr_t)dst | (uintptr_t)(w ? w : x)) & 15) == 0);
    if (vec) {
        const int n4 = (int)(ne0 / 4);
        const int thr = n4 <= 256 ? 256 : 512;
        const int per = (n4 + thr - 1) / thr;
        if (per <= 1) return launch_kernel(rms_norm_row_kernel<1>, dim3((unsigned)nrows), dim3(thr), 0, s, pdl, x, w, eps, ne0, dst);
        if (per <= 2) return launch_kernel(rms_norm_row_kernel<2>, dim3((unsigned)nrows), dim3(thr), 0, s, pdl, x, w, eps, ne0, dst);
        return launch_kernel(rms_norm_row_kernel<4>, dim3((unsigned)nrows), dim3(thr), 0, s, pdl, x, w, eps, ne0, dst);
    }
    const int threads = ne0 >= 2048 ? 512 : 256;
    return launch_kernel(rms_norm_kernel, dim3((unsigned)nrows), dim3(threads), 0, s, pdl, x, w, eps, ne0, dst);
}

// ---------------------------------------------------------------------------
// K5 softmax over ne[0] — LlamaAttention.softmax, model/LlamaAttention.kt:242-288:
// max, exp(x - max), F32 sum, divide.  One CTA per row, three passes over a row
// that stays in L1/L2 (8 algorithmic bytes per element).
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) softmax_kernel(const float *__restrict__ x, int64_t ne0,
                                                      float *__restrict__ dst) {
    __shared__ float scratch[32];
    pdl_wait();                // inputs may be the previous kernel's output
    pdl_launch_dependents();   // the next kernel may start its prologue (a GEMV: its weight prefetch) under this one
    const float *xr = x + (int64_t)blockIdx.x * ne0;
    float *dr = dst + (int64_t)blockIdx.x * ne0;
    float m = -INFINITY;
    for (int64_t i = threadIdx.x; i < ne0; i += blockDim.x) m = fmaxf(m, xr[i]);
    m = block_reduce(m, [](float a, float b) { return fmaxf(a, b); }, scratch);
    float sum = 0.f;
    for (int64_t i = threadIdx.x; i < ne0; i += blockDim.x) {
        const float e = expf(__fsub_rn(xr[i], m));
        dr[i] = e;
        sum += e;
    }
    sum = block_reduce(sum, [](float a, float b) { return a + b; }, scratch);
    for (int64_t i = threadIdx.x; i < ne0; i += blockDim.x) dr[i] = __fdiv_rn(dr[i], sum);
}

// Register-resident rows (same limits as rms_norm_row_kernel): one HBM pass; max, expf(x - max), F32 sum, IEEE division
// as LlamaAttention.softmax (model/LlamaAttention.kt:242-288).
template <int V>
__global__ void __launch_bounds__(512) softmax_row_kernel(const float *__restrict__ x, int64_t ne0, float *__restrict__ dst) {
    __shared__ float scratch[32];
    pdl_wait();
    pdl_launch_dependents();
    const float4 *xr = (const float4 *)(x + (int64_t)blockIdx.x * ne0);
    float4 *dr = (float4 *)(dst + (int64_t)blockIdx.x * ne0);
    const int n4 = (int)(ne0 >> 2);
    float4 v[V];
    float m = -INFINITY;
#pragma unroll
    for (int j = 0; j < V; j++) {
        const int i = threadIdx.x + j * blockDim.x;
        v[j] = i < n4 ? xr[i] : make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
        m = fmaxf(fmaxf(m, fmaxf(v[j].x, v[j].y)), fmaxf(v[j].z, v[j].w));
    }
    m = block_reduce(m, [](float a, float b) { return fmaxf(a, b); }, scratch);
    float sum = 0.f;
#pragma unroll
    for (int j = 0; j < V; j++) {
        const int i = threadIdx.x + j * blockDim.x;
        if (i < n4) {
            v[j] = make_float4(expf(__fsub_rn(v[j].x, m)), expf(__fsub_rn(v[j].y, m)), expf(__fsub_rn(v[j].z, m)), expf(__fsub_rn(v[j].w, m)));
            sum += (v[j].x + v[j].y) + (v[j].z + v[j].w);
        }
    }
    sum = block_reduce(sum, [](float a, float b) { return a + b; }, scratch);
#pragma unroll
    for (int j = 0; j < V; j++) {
        const int i = threadIdx.x + j * blockDim.x;
        if (i < n4) dr[i] = make_float4(__fdiv_rn(v[j].x, sum), __fdiv_rn(v[j].y, sum), __fdiv_rn(v[j].z, sum), __fdiv_rn(v[j].w, sum));
    }
}

cudaError_t launch_softmax(const float *x, int64_t ne0, int64_t nrows, float *dst, cudaStream_t s, bool pdl) {
    if (nrows == 0 || ne0 == 0) return cudaSuccess;
    if ((ne0 % 4) == 0 && ne0 <= 8192 && ((((uintptr_t)x | (uintptr_t)dst) & 15) == 0)) {
        const int n4 = (int)(ne0 / 4);
        const int thr = n4 <= 256 ? 256 : 512;
        const int per = (n4 + thr - 1) / thr;
        if (per <= 1) return launch_kernel(softmax_row_kernel<1>, dim3((unsigned)nrows), dim3(thr), 0, s, pdl, x, ne0, dst);
        if (per <= 2) return launch_kernel(softmax_row_kernel<2>, dim3((unsigned)nrows), dim3(thr), 0, s, pdl, x, ne0, dst);
        return launch_kernel(softmax_row_kernel<4>, dim3((unsigned)nrows), dim3(thr), 0, s, pdl, x, ne0, dst);
    }
    return launch_kernel(softmax_kernel, dim3((unsigned)nrows), dim3(256), 0, s, pdl, x, ne0, dst);
}

// ---------------------------------------------------------------------------
// A11 quantizeTensor — core/GGMLComputeOps.kt:1040-1203 (Q8_0 :1058-1071, Q4_0 :1072-1087, Q4_1 :1162-1203) with
// floatToHalf, core/NumericConversions.kt:61-124, INCLUDING its half-denormal off-by-one (:83): bit-exact
// with the reference, so blocks quantized on the device are the blocks the Kotlin code would have produced.
// One warp per 32-element block (lane = element): HBM-bound, 4 B read + blockBytes/32 written per element.
//   * every F32 operation is one IEEE operation in the reference's order (no FMA contraction);
//   * kotlin.math.round = ties-to-even (rintf), Float.toInt() saturates, NaN -> 0, then coerceIn;
//   * maxOf / minOf return NaN if either operand is NaN; among equal candidates (+0 / -0) the sequential scans
//     `m = if (m < x) m else x` keep the LAST one, which the shuffle reduction reproduces with an index tie-break.
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint16_t kt_float_to_half(float f) {
    const uint32_t bits = __float_as_uint(f);
    const uint32_t sign = (bits >> 16) & 0x8000u;
    const int32_t a = (int32_t)(bits & 0x7FFFFFFFu);
    if (a > 0x47FFEFFF) return (uint16_t)(sign | 0x7C00u | ((a & 0x007FFFFF) ? 0x0200u : 0u));   // overflow / inf / NaN (:67-73)
    if (a < 0x38800000) {                                                                         // half denormal range (:75-100)
        const uint32_t mant = ((uint32_t)a & 0x007FFFFFu) | 0x00800000u;
        const int32_t shift = 127 - (a >> 23);                       // sic: the reference's off-by-one (should be 126 - exp)
        const uint32_t hm = shift < 24 ? (mant >> (shift & 31)) : 0u;                             // Kotlin masks shift counts to 5 bits
        const int32_t rb = (int32_t)(mant & ((1u << (shift & 31)) - 1u));
        const int32_t half = (int32_t)(1u << ((shift - 1) & 31));
        if (rb > half || (rb == half && (hm & 1u))) {
            const uint32_t t = hm + 1u;
            return (uint16_t)(sign | (t == 0x0400u ? (1u << 10) : t));
        }
        return (uint16_t)(sign | hm);
    }
    const uint32_t he = (((uint32_t)a >> 23) - 112u) << 10;                                       // normal range (:102-122)
    uint32_t hm = ((uint32_t)a & 0x007FFFFFu) >> 13;
    if ((a & 0x00001000) && ((a & 0x00000FFF) || (hm & 1u))) {
        hm++;
        if (hm == 0x0400u) return (uint16_t)(sign | (he + (1u << 10)));
    }
    return (uint16_t)(sign | he | hm);
}

__device__ __forceinline__ int kt_round_clamp(float v, int lo, int hi) {   // round(v).toInt().coerceIn(lo, hi)
    const float r = rintf(v);
    if (r != r) return 0 < lo ? lo : (0 > hi ? hi : 0);
    if (r <= (float)lo) return lo;
    if (r >= (float)hi) return hi;
    return (int)r;
}

// warp-wide min / max with the sequential scan's semantics (see above); every lane gets the result
template <bool MAX>
__device__ __forceinline__ float kt_warp_extreme(float v, int lane) {
    int idx = lane;
    bool nan = v != v;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, v, o);
        const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
        const bool better = MAX ? (ov > v) : (ov < v);
        if (better || (ov == v && oi > idx)) { v = ov; idx = oi; }
    }
    nan = __any_sync(0xffffffffu, nan);
    return nan ? __uint_as_float(0x7FC00000u) : v;
}

template <int TYPE>
__global__ void __launch_bounds__(256) quantize_kernel(const float *__restrict__ x, int64_t nblocks, uint8_t *__restrict__ out, int nib) {
    constexpr int BB = block_bytes(TYPE);
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t b = warp; b < nblocks; b += nwarps) {
        const float v = x[b * 32 + lane];
        uint8_t *ob = out + b * BB;
        int q;
        if (TYPE == T_Q4_1) {
            const float fmin = kt_warp_extreme<false>(v, lane), fmax = kt_warp_extreme<true>(v, lane);
            float d = __fdiv_rn(__fsub_rn(fmax, fmin), 15.0f);
            if (d == 0.0f) d = 1.0f;
            const float inv = __fdiv_rn(1.0f, d);
            if (lane == 0) {
                *(uint16_t *)ob = kt_float_to_half(d);
                *(uint16_t *)(ob + 2) = kt_float_to_half(fmin);
            }
            q = kt_round_clamp(__fmul_rn(__fsub_rn(v, fmin), inv), 0, 15);
        } else {
            // amax = 0f; for each: amax = maxOf(amax, abs(x)): values are >= +0, ties are identical bit patterns
            const float amax = kt_warp_extreme<true>(fabsf(v), lane);
            const float scale = amax == 0.0f ? 1.0f : __fdiv_rn(amax, TYPE == T_Q8_0 ? 127.0f : 8.0f);
            const float inv = (TYPE == T_Q4_0 && scale == 0.0f) ? 0.0f : __fdiv_rn(1.0f, scale);
            if (lane == 0) *(uint16_t *)ob = kt_float_to_half(scale);
            if (TYPE == T_Q8_0) q = kt_round_clamp(__fmul_rn(v, inv), -128, 127);
            else q = kt_round_clamp(__fadd_rn(__fmul_rn(v, inv), 8.0f), 0, 15);
        }
        if (TYPE == T_Q8_0) {
            ob[2 + lane] = (uint8_t)(int8_t)q;
        } else {
            constexpr int QOFF = TYPE == T_Q4_0 ? 2 : 4;
            // Kotlin order: byte j = elements 2j (low nibble), 2j+1 (high); ggml order: byte j = elements j (low), j+16 (high)
            const int partner = __shfl_xor_sync(0xffffffffu, q, nib ? 16 : 1);
            if (nib ? lane < 16 : (lane & 1) == 0)
                ob[QOFF + (nib ? lane : lane >> 1)] = (uint8_t)((q & 15) | ((partner & 15) << 4));
        }
    }
}

cudaError_t launch_quantize(int type, const float *x, int64_t n, void *out, cudaStream_t s, int nib) {
    if (n == 0) return cudaSuccess;
    const int64_t nblocks = n / 32;
    const int grid = grid_for(nblocks * 32, 256);
    switch (type) {
        case T_Q8_0: quantize_kernel<T_Q8_0><<<grid, 256, 0, s>>>(x, nblocks, (uint8_t *)out, nib); break;
        case T_Q4_0: quantize_kernel<T_Q4_0><<<grid, 256, 0, s>>>(x, nblocks, (uint8_t *)out, nib); break;
        case T_Q4_1: quantize_kernel<T_Q4_1><<<grid, 256, 0, s>>>(x, nblocks, (uint8_t *)out, nib); break;
        default: return cudaErrorNotSupported;
    }
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------
// A9 strided F32 tensors (GGMLTensor.getFloat / setFloat through nb[], core/GGMLTypes.kt:328-380): gather a strided
// tensor into the contiguous layout the kernels take, or scatter a contiguous result back through dst's strides.
// Element (i0, i1, i2, i3) lives at byte offset sum i_d * nb[d].
// ---------------------------------------------------------------------------
struct StrideDesc { int64_t ne[4]; int64_t nb[4]; };
__global__ void __launch_bounds__(256) strided_copy_kernel(const uint8_t *__restrict__ src, uint8_t *__restrict__ dst, StrideDesc d, int64_t n, int scatter) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        int64_t r = i, off = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) { const int64_t ik = r % d.ne[k]; r /= d.ne[k]; off += ik * d.nb[k]; }
        if (scatter) *(float *)(dst + off) = ((const float *)src)[i];
        else ((float *)dst)[i] = *(const float *)(src + off);
    }
}
cudaError_t launch_strided_copy(const void *src, void *dst, const int64_t ne[4], const uint64_t nb[4], int scatter, cudaStream_t s) {
    StrideDesc d;
    int64_t n = 1;
    for (int k = 0; k < 4; k++) { d.ne[k] = ne[k] > 0 ? ne[k] : 1; d.nb[k] = (int64_t)nb[k]; n *= d.ne[k]; }
    if (n == 0) return cudaSuccess;
    strided_copy_kernel<<<grid_for(n, 256), 256, 0, s>>>((const uint8_t *)src, (uint8_t *)dst, d, n, scatter);
    return cudaGetLastError();
}

}  // namespace lk

// lk_kquants.cu — the reference's K-quant and BitNet 1.58 paths.
//
// Replaces computeMatMul's Q2_K / Q4_K / Q8_K x F32 branches (core/GGMLComputeOps.kt:1481-1513 ->
// computeDotProductQ2_KF32 :152-236, Q4_KF32 :241-341, Q8_KF32 :382-432), computeDotProductBitNet158F32 (:635-660) and
// the matching dequantizeTensor cases (:966-1036, block routines :2125-2155, :2207-2250, :2356-2368).
//
// These are the REFERENCE'S OWN formats (block sizes of ggml-common.h, simpler field meaning):
//   Q2_K  84 B / 256:  scales[16] @0 (lo nibble scale, hi nibble min), qs[64] @16 (4 x 2 bit), d f16 @80, dmin f16 @82
//                      w = (q/3) * ((qs/15) * d) + (qm * d + dmin)                   core/GGMLTypes.kt:737-798
//   Q4_K  144 B / 256: d f16 @0, dmin f16 @2, scales[12] @4, qs[128] @16 (lo nibble = even element)
//                      w = (q/15) * ((s & 63)/63 * d) + (qmin/63 * d + dmin), qmin = (s >> 6) | (scales[2 sub + 1] & 15) << 2
//                      for sub <= 5                                                  core/GGMLComputeOps.kt:262-306
//   Q8_K  292 B / 256: d f32 @0, qs int8[256] @4, bsums @260;  w = float(q) * d      core/GGMLTypes.kt:886-918
//   BITNET_1_58 10 B / 32: d f16 @0, 7 bytes x 5 base-3 digits @2; w = d * (digit - 1)   core/GGMLTypes.kt:920-1000
//
// Three kernels:
//   kq_dequant_kernel   bit-exact with dequantizeTensor (one IEEE op per reference op, __fmul_rn / __fadd_rn / __fdiv_rn);
//   kq_gemv_kernel      N == 1, rows of whole blocks, for the shapes K1 declines (Q2_K / Q4_K / Q8_K with K % 256 == 0 and
//                       BITNET_1_58 with K % 64 == 0 normally run on the streaming DP4A GEMV of lk_gemv.cu, block_dot<T_Q4_K>
//                       etc., 3-9 x faster: an int->F32 conversion + FMA per element is instruction-bound): one CTA per 4 rows, a 256-element super-block
//                       of each row per warp step (8 elements per lane, 4 / 2 / 8-byte quant loads, 2 x LDG.128 of x
//                       shared by the 4 rows), per sub-block scale * sum(q x) + min * sum(x) in F32, shuffle + shared
//                       reduction;  algorithmic bytes = M K/256 B + 4K + 4M;
//   kq_matmul_generic_kernel  every other shape, with the reference's block addressing restated literally (a whole
//                       256-chunk starting at k0 reads block (i K + k0)/256 from its element 0; a trailing partial chunk
//                       is addressed by flat index, and for Q4_K adds dmin without the sub-block min, :308-336).
// The matmul bar is the 1e-3 tolerance of the main path (summation order differs from the sequential reference).
#include "lk_common.cuh"
#include "lk_kernels.h"

namespace lk {

__host__ __device__ constexpr int kq_block_bytes(int t) {
    return t == T_Q2_K ? 84 : t == T_Q4_K ? 144 : t == T_Q8_K ? 292 : t == T_BITNET_1_58 ? 10 : 0;
}
__host__ __device__ constexpr int kq_block_elems(int t) {
    return (t == T_Q2_K || t == T_Q4_K || t == T_Q8_K) ? 256 : t == T_BITNET_1_58 ? 32 : 0;
}

int kq_type_block_bytes(int type) { return kq_block_bytes(type); }
int kq_type_block_elems(int type) { return kq_block_elems(type); }

__device__ __forceinline__ uint16_t ld_u16(const uint8_t *p) { return *(const uint16_t *)p; }

// one dequantised element, reference operation order (core/GGMLComputeOps.kt:181-200, :262-306, :407-413, :653-655)
template <int TYPE>
__device__ __forceinline__ float kq_value(const uint8_t *a, int64_t blk, int item, bool partial) {
    const uint8_t *p = a + blk * kq_block_bytes(TYPE);
    if (TYPE == T_Q2_K) {
        const float d = h2f(ld_u16(p + 80)), dmin = h2f(ld_u16(p + 82));
        const int sub = item >> 4, sm = p[sub];
        const float scale = __fmul_rn(__fdiv_rn((float)(sm & 15), 15.0f), d);
        const float mn = __fadd_rn(__fmul_rn((float)(sm >> 4), d), dmin);
        const int q = (p[16 + (item >> 2)] >> ((item & 3) * 2)) & 3;
        return __fadd_rn(__fmul_rn(__fdiv_rn((float)q, 3.0f), scale), mn);
    } else if (TYPE == T_Q4_K) {
        const float d = h2f(ld_u16(p)), dmin = h2f(ld_u16(p + 2));
        const int sub = item >> 5, sb = p[4 + sub];
        const int qmin = (sb >> 6) | ((sub <= 5 ? (p[4 + sub * 2 + 1] & 15) : 0) << 2);
        const float scale = __fmul_rn(__fdiv_rn((float)(sb & 63), 63.0f), d);
        const float mn = __fadd_rn(__fmul_rn(__fdiv_rn((float)qmin, 63.0f), d), dmin);
        const int byte = p[16 + (item >> 1)];
        const int q = (item & 1) ? (byte >> 4) : (byte & 15);
        return __fadd_rn(__fmul_rn(__fdiv_rn((float)q, 15.0f), scale), partial ? dmin : mn);
    } else if (TYPE == T_Q8_K) {
        return __fmul_rn((float)(int8_t)p[4 + item], *(const float *)p);
    } else {
        const float d = h2f(ld_u16(p));
        int byte = p[2 + item / 5];
        const int pos = item % 5;
        for (int i = 0; i < pos; i++) byte /= 3;
        const int digit = pos == 4 ? byte : byte % 3;            // the leading digit is 3 for bytes >= 243: the
        if (digit > 2) return __uint_as_float(0x7FC00000u);      // reference throws IllegalStateException there
        return __fmul_rn(d, (float)(digit - 1));
    }
}

template <int TYPE>
__global__ void __launch_bounds__(256) kq_dequant_kernel(const uint8_t *__restrict__ src, int64_t n, float *__restrict__ dst) {
    constexpr int BE = kq_block_elems(TYPE);
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; g < n / 4; g += stride) {
        const int64_t f = g * 4;
        const int64_t blk = f / BE;
        const int item = (int)(f % BE);
        float4 v;
        v.x = kq_value<TYPE>(src, blk, item, false);
        v.y = kq_value<TYPE>(src, blk, item + 1, false);
        v.z = kq_value<TYPE>(src, blk, item + 2, false);
        v.w = kq_value<TYPE>(src, blk, item + 3, false);
        *(float4 *)(dst + f) = v;
    }
}

cudaError_t launch_kq_dequantize(int type, const void *src, int64_t n, float *dst, cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    const int threads = 256;
    int64_t blocks = (n / 4 + threads - 1) / threads;
    if (blocks > 148 * 16) blocks = 148 * 16;
    const int grid = (int)blocks;
    switch (type) {
        case T_Q2_K: kq_dequant_kernel<T_Q2_K><<<grid, threads, 0, s>>>((const uint8_t *)src, n, dst); break;
        case T_Q4_K: kq_dequant_kernel<T_Q4_K><<<grid, threads, 0, s>>>((const uint8_t *)src, n, dst); break;
        case T_Q8_K: kq_dequant_kernel<T_Q8_K><<<grid, threads, 0, s>>>((const uint8_t *)src, n, dst); break;
        case T_BITNET_1_58: kq_dequant_kernel<T_BITNET_1_58><<<grid, threads, 0, s>>>((const uint8_t *)src, n, dst); break;
        default: return cudaErrorNotSupported;
    }
    return cudaGetLastError();
}

// ---- decode GEMV ------------------------------------------------------------------------------------------------------
// One CTA (8 warps) per group of 4 rows; warp w takes super-blocks w, w+8, ... of all 4 rows, so the 8 activations a lane
// loads are used 4 times (x comes through L1 at 4 B per element: without the reuse a 2-bit format would be L1-bound at
// less than half the HBM rate) and >= 64 KB of weight loads are in flight per SM (64 warps x 4 rows x 2 steps).
constexpr int KQ_ROWS = 4;

// lane l owns elements [8l, 8l+8) of the super-block at p; xv = the matching activations
// (scales by reciprocal multiplication, not the reference's IEEE divisions: the matmul bar is a tolerance and a division
// is ~10 instructions per lane per block)
template <int TYPE>
__device__ __forceinline__ float kq_lane_dot(const uint8_t *__restrict__ p, const float (&xv)[8], float sx, int lane) {
    if (TYPE == T_Q8_K) {
        const uint32_t q0 = *(const uint32_t *)(p + 4 + lane * 8), q1 = *(const uint32_t *)(p + 8 + lane * 8);   // 292-byte blocks: 4-byte aligned
        const float d = *(const float *)p;
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < 4; i++) {
            s = fmaf((float)(int8_t)(q0 >> (8 * i)), xv[i], s);
            s = fmaf((float)(int8_t)(q1 >> (8 * i)), xv[4 + i], s);
        }
        return d * s;
    } else if (TYPE == T_Q4_K) {
        const uint32_t hdr = *(const uint32_t *)p;                   // d | dmin << 16
        const uint32_t q = *(const uint32_t *)(p + 16 + lane * 4);   // bytes 4l..4l+3: elements 8l+2j (lo), 8l+2j+1 (hi)
        const int sub = lane >> 2, sb = p[4 + sub];
        const int qmin = (sb >> 6) | ((sub <= 5 ? (p[4 + sub * 2 + 1] & 15) : 0) << 2);
        const float d = h2f((uint16_t)hdr), dmin = h2f((uint16_t)(hdr >> 16));
        const float scale = (float)(sb & 63) * d * (1.0f / (63.0f * 15.0f));
        const float mn = fmaf((float)qmin * (1.0f / 63.0f), d, dmin);
        float sq = 0.f;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            sq = fmaf((float)((q >> (8 * j)) & 15u), xv[2 * j], sq);
            sq = fmaf((float)((q >> (8 * j + 4)) & 15u), xv[2 * j + 1], sq);
        }
        return fmaf(scale, sq, mn * sx);
    } else {   // Q2_K
        const uint32_t tail = *(const uint32_t *)(p + 80);           // d | dmin << 16 (84-byte blocks: 4-byte aligned)
        const uint32_t q = ld_u16(p + 16 + lane * 2);                // bytes 2l, 2l+1: elements 8l+4b+j at bits 2j
        const int sm = p[lane >> 1];
        const float d = h2f((uint16_t)tail), dmin = h2f((uint16_t)(tail >> 16));
        const float scale = (float)(sm & 15) * d * (1.0f / 45.0f);
        const float mn = fmaf((float)(sm >> 4), d, dmin);
        float sq = 0.f;
#pragma unroll
        for (int i = 0; i < 8; i++) sq = fmaf((float)((q >> (2 * i)) & 3u), xv[i], sq);
        return fmaf(scale, sq, mn * sx);
    }
}

// BitNet 1.58: the lane owns one 32-element block (10 bytes, 2-byte aligned); xv = its 32 activations, sx their sum
__device__ __forceinline__ float bitnet_lane_dot(const uint8_t *__restrict__ p, const float (&xv)[32], float sx) {
    uint32_t w[3];
    w[0] = ld_u16(p) | ((uint32_t)ld_u16(p + 2) << 16);
    w[1] = ld_u16(p + 4) | ((uint32_t)ld_u16(p + 6) << 16);
    w[2] = ld_u16(p + 8);
    const float d = h2f((uint16_t)w[0]);
    float st = 0.f;
    bool bad = false;
#pragma unroll
    for (int g = 0; g < 7; g++) {
        const int bi = 2 + g;
        uint32_t byte = (w[bi >> 2] >> (8 * (bi & 3))) & 0xFFu;
        bad |= byte >= 243u;           // leading digit 3: the reference throws IllegalStateException
#pragma unroll
        for (int pos = 0; pos < 5; pos++) {
            const int e = g * 5 + pos;
            if (e < 32) {
                const uint32_t qd = (byte * 171u) >> 9;   // byte / 3 for byte < 256
                st = fmaf((float)(byte - 3u * qd), xv[e], st);
                byte = qd;
            }
        }
    }
    const float r = d * (st - sx);      // d * sum((digit - 1) x)
    return bad ? __uint_as_float(0x7FC00000u) : r;
}

template <int TYPE>
__global__ void __launch_bounds__(256, 4) kq_gemv_kernel(const uint8_t *__restrict__ a, int64_t K, int64_t M,
                                                      const float *__restrict__ b, float *__restrict__ dst) {
    constexpr int BB = kq_block_bytes(TYPE), BE = kq_block_elems(TYPE), R = KQ_ROWS;
    __shared__ float red[8][R];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int64_t nb = K / BE;                    // blocks per row
    for (int64_t g = blockIdx.x; g * R < M; g += gridDim.x) {
        const uint8_t *row[R];
#pragma unroll
        for (int r = 0; r < R; r++) {
            const int64_t i = g * R + r;
            row[r] = a + (i < M ? i : M - 1) * nb * BB;     // rows past M repeat the last one; their sums are dropped
        }
        float acc[R];
#pragma unroll
        for (int r = 0; r < R; r++) acc[r] = 0.f;
        if (TYPE == T_BITNET_1_58) {
            for (int c = w * 32 + lane; c < (int)nb; c += 256) {
                float xv[32], sx = 0.f;
#pragma unroll
                for (int q = 0; q < 8; q++) {
                    const float4 f = __ldg((const float4 *)(b + c * 32) + q);
                    xv[4 * q] = f.x; xv[4 * q + 1] = f.y; xv[4 * q + 2] = f.z; xv[4 * q + 3] = f.w;
                    sx += (f.x + f.y) + (f.z + f.w);
                }
#pragma unroll
                for (int r = 0; r < R; r++) acc[r] += bitnet_lane_dot(row[r] + c * BB, xv, sx);
            }
        } else {
#pragma unroll 1      // 4 rows = 4 independent load streams per step already; unrolling further spills at 64 registers
            for (int c = w; c < (int)nb; c += 8) {
                const float4 x0 = __ldg((const float4 *)(b + c * 256 + lane * 8)), x1 = __ldg((const float4 *)(b + c * 256 + lane * 8 + 4));
                const float xv[8] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w};
                const float sx = ((x0.x + x0.y) + (x0.z + x0.w)) + ((x1.x + x1.y) + (x1.z + x1.w));
#pragma unroll
                for (int r = 0; r < R; r++) acc[r] += kq_lane_dot<TYPE>(row[r] + c * BB, xv, sx, lane);
            }
        }
#pragma unroll
        for (int r = 0; r < R; r++) {
            acc[r] = warp_sum(acc[r]);
            if (lane == 0) red[w][r] = acc[r];
        }
        __syncthreads();
        if (threadIdx.x < R && g * R + threadIdx.x < M) {
            float v = 0.f;
#pragma unroll
            for (int ww = 0; ww < 8; ww++) v += red[ww][threadIdx.x];
            dst[g * R + threadIdx.x] = v;
        }
        __syncthreads();
    }
}

// ---- every other shape: one warp per (row, tile of <= 8 columns), lanes stride over k ---------------------------------
template <int TYPE>
__global__ void __launch_bounds__(128) kq_matmul_generic_kernel(const uint8_t *__restrict__ a, int64_t K, int64_t M,
                                                                const float *__restrict__ b, int64_t N,
                                                                float *__restrict__ dst, int64_t ntiles_n) {
    constexpr int TN = 8, BE = kq_block_elems(TYPE);
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int64_t kwhole = (TYPE == T_BITNET_1_58) ? 0 : (K / 256) * 256;   // k below this lies in a whole 256-chunk
    for (int64_t w = warp; w < M * ntiles_n; w += nwarps) {
        const int64_t i = w / ntiles_n;
        const int64_t j0 = (w % ntiles_n) * TN;
        const int nj = (int)((N - j0) < TN ? (N - j0) : TN);
        float acc[TN];
#pragma unroll
        for (int j = 0; j < TN; j++) acc[j] = 0.f;
        for (int64_t k = lane; k < K; k += 32) {
            int64_t blk; int item; bool partial = false;
            if (k < kwhole) { const int64_t k0 = k & ~(int64_t)255; blk = (i * K + k0) >> 8; item = (int)(k - k0); }
            else { const int64_t flat = i * K + k; blk = flat / BE; item = (int)(flat % BE); partial = TYPE != T_BITNET_1_58; }
            const float wv = kq_value<TYPE>(a, blk, item, partial);
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

cudaError_t launch_kq_matmul(int type, const void *a, int64_t K, int64_t M, const float *b, int64_t N, float *dst,
                             int sm_count, cudaStream_t s) {
    if (M == 0 || N == 0) return cudaSuccess;
    const int be = kq_block_elems(type);
    if (be == 0) return cudaErrorNotSupported;
    const uint8_t *pa = (const uint8_t *)a;
    if (N == 1 && K > 0 && K % be == 0 && ((uintptr_t)a & 3) == 0 && ((uintptr_t)b & 15) == 0) {
        const int threads = 256;
        int64_t grid = (M + KQ_ROWS - 1) / KQ_ROWS;           // one CTA per 4 rows, 8 CTAs resident per SM
        if (grid > (int64_t)sm_count * 32) grid = (int64_t)sm_count * 32;
        switch (type) {
            case T_Q2_K: kq_gemv_kernel<T_Q2_K><<<(int)grid, threads, 0, s>>>(pa, K, M, b, dst); break;
            case T_Q4_K: kq_gemv_kernel<T_Q4_K><<<(int)grid, threads, 0, s>>>(pa, K, M, b, dst); break;
            case T_Q8_K: kq_gemv_kernel<T_Q8_K><<<(int)grid, threads, 0, s>>>(pa, K, M, b, dst); break;
            default: kq_gemv_kernel<T_BITNET_1_58><<<(int)grid, threads, 0, s>>>(pa, K, M, b, dst); break;
        }
        return cudaGetLastError();
    }
    const int64_t ntn = (N + 7) / 8;
    const int threads = 128;
    int64_t grid = (M * ntn * 32 + threads - 1) / threads;
    if (grid > (int64_t)sm_count * 32) grid = (int64_t)sm_count * 32;
    switch (type) {
        case T_Q2_K: kq_matmul_generic_kernel<T_Q2_K><<<(int)grid, threads, 0, s>>>(pa, K, M, b, N, dst, ntn); break;
        case T_Q4_K: kq_matmul_generic_kernel<T_Q4_K><<<(int)grid, threads, 0, s>>>(pa, K, M, b, N, dst, ntn); break;
        case T_Q8_K: kq_matmul_generic_kernel<T_Q8_K><<<(int)grid, threads, 0, s>>>(pa, K, M, b, N, dst, ntn); break;
        default: kq_matmul_generic_kernel<T_BITNET_1_58><<<(int)grid, threads, 0, s>>>(pa, K, M, b, N, dst, ntn); break;
    }
    return cudaGetLastError();
}

}  // namespace lk

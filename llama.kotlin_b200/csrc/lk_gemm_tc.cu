// lk_gemm_tc.cu — K2: prefill GEMM  dst[N,M] = dequant(a[K,M]) x b[N,K]  on the tcgen05 tensor cores.
//
// Replaces the N > 1 case of computeMatMul's Q x F32 branch (core/GGMLComputeOps.kt:1450-1528: for every
// (row i, column j) a dot of a's packed row i with column j of b, F32 accumulation) for the batched,
// dense-contraction shapes (BASELINE config 3: Q8_0 11008x4096 x 4096x512).
//
// Data flow per CTA (one 128-row x BN-column output tile, all of K):
//   * packed weight rows never leave their on-disk layout in HBM.  The producer warp copies, per 256-K
//     chunk, the 128 row pieces (8 blocks = 272 / 144 / 160 bytes each) into a 2-deep raw ring with one
//     2-D tiled TMA load (UTMALDG; 128 separate bulk copies were the bottleneck, profiles/r01c), completion
//     on an mbarrier;
//   * 8 dequant warps turn each raw piece into FP16 (one 32-element block per thread per 64-K step; the
//     value is round_fp16(d*q), one rounding, because d and q are both exact in FP16) and store it into
//     the A stage in the canonical K-major SWIZZLE_128B layout the tensor core reads;
//   * the activations were converted once per call by pack_b_kernel into FP16 tiles already in that
//     layout (workspace), so a B stage is a single bulk copy;
//   * one thread issues tcgen05.mma (cta_group::1, kind::f16, M=128, N=BN, K=16) with both operands in
//     shared memory and the F32 accumulator in tensor memory; tcgen05.commit releases the stage;
//   * after the last K step the dequant warps read the accumulator back (tcgen05.ld 32x32b) and store
//     dst rows directly (each thread owns one output row: 32 contiguous bytes per store pair).
//
// Numerics (precision modes, ctx option gemm_tc_precision; the reference is F32 throughout,
// core/GGMLComputeOps.kt:1516-1527):
//   2 (default) A = hi + lo and b = hi + lo, both halves FP16, three tcgen05.mma per K step into the same F32
//               accumulator (hi*hi + lo*hi + hi*lo).  d*q is EXACTLY hi + lo for Q8_0 / Q4_0 (d is a multiple of
//               2^-24 with 11 significant bits, q an integer of <= 8 bits: the residual of the FP16 product is
//               itself an FP16 number), b keeps 22 significant bits: F32-grade operands.
//   1           A = hi + lo, b single FP16 (two MMAs): weights exact, activations rounded to 11 bits.
//   0           single FP16 operands (one MMA): the round-1 kernel; ~3e-4 of max|dst| on random data, 1e-3 on
//               periodic data.
// Range: every column of b is scaled by an exact power of two (2^(13-e), e = exponent of the column's max |x|,
// undone in the epilogue), so activations of any F32 magnitude stay inside FP16; a weight block whose d*q could
// leave the FP16 range (|d| > 512 for Q8_0, > 4096 for Q4_0, ...; inf / NaN scales) makes its CTA recompute
// the tile in F32 with the reference's own operation order instead of storing the tensor-core result.
#include <cuda.h>   // CUtensorMap types only; the encoder is fetched through cudaGetDriverEntryPoint (no libcuda link)

#include "lk_common.cuh"
#include "lk_kernels.h"

namespace lk {
namespace {

constexpr int TC_BM = 128;            // rows of a per CTA = UMMA M
constexpr int TC_BK = 64;             // K per stage: 64 halfs = one 128-byte swizzle row
constexpr int TC_CHUNK_K = 256;       // K per raw chunk: 8 blocks, the smallest piece that is 16-byte aligned for all types
constexpr int TC_DQ_GROUP = 8;        // warps that fill one A stage (4 TMEM lane groups x 2 halves of the 64-K step)
constexpr int TC_DQ_GROUPS = 1;       // 2: two groups alternate 64-K steps (measured 4% slower on config 3: the tensor pipe, not the
                                      // convert chain, paces the loop; profiles/r01h, tools/gemm_tc_timeline.cu)
constexpr int TC_DQ_WARPS = TC_DQ_GROUP * TC_DQ_GROUPS;
constexpr int TC_CTRL_WARPS = 3;      // warp 0: activation-tile producer, warp 1: MMA issuer, warp 2: raw-weight producer
constexpr int TC_THREADS = (TC_CTRL_WARPS + TC_DQ_WARPS) * 32;
constexpr int TC_A_COL0 = 256;                        // TMEM: accumulator in columns [0, 256), A stages of 32 columns from 256
constexpr int TC_MAX_STAGES = 8;

// bytes of one row piece (256 K) and the row stride of the raw ring = inner extent of the TMA box.  The stride is
// an odd number of 16-byte units so that 8 consecutive rows hit 8 different bank groups with LDS.128; for Q4_1
// the box therefore reads 16 bytes past the piece (the next piece's first bytes, or zero fill at the row end).
template <int TYPE> struct Raw;
template <> struct Raw<T_Q8_0> { static constexpr int bytes = 272, stride = 272; };   // 17 x 16 B
template <> struct Raw<T_Q4_0> { static constexpr int bytes = 144, stride = 144; };   //  9 x 16 B
template <> struct Raw<T_Q4_1> { static constexpr int bytes = 160, stride = 176; };   // 10 -> 11 x 16 B

// 2-D tiled TMA load (SASS: UTMALDG): box {stride/2 uint16, 128 rows} at element coordinates (c0, c1);
// rows past the tensor's M are zero-filled and still counted in the transaction bytes.
__device__ __forceinline__ void tma_load_2d(void *smem_dst, const CUtensorMap *tm, int c0, int c1, uint64_t *bar, uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%2, %3}], [%4], %5;"
        ::"r"(smem_u32(smem_dst)), "l"(tm), "r"(c0), "r"(c1), "r"(smem_u32(bar)), "l"(policy)
        : "memory");
}
// 1-D bulk copy delivered to the same shared-memory offset of every CTA in `mask` of this cluster; each
// destination CTA's mbarrier (same offset) receives the transaction bytes.
__device__ __forceinline__ void tma_load_1d_multicast(void *smem_dst, const void *gmem_src, uint32_t bytes, uint64_t *bar, uint16_t mask,
                                                      uint64_t policy) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster.L2::cache_hint [%0], [%1], %2, [%3], %4, %5;"
        ::"r"(smem_u32(smem_dst)), "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar)), "h"(mask), "l"(policy)
        : "memory");
}
// One lane of the (converged) warp.  The single-thread instructions (TMA, tcgen05.mma/commit) take uniform-register
// operands: issued under elect.sync from a converged warp the compiler keeps addresses in uniform registers; issued
// under `if (lane == 0)` it wraps every one of them in an ELECT/R2UR/BRA.U.ANY waterfall loop (~100 clk each,
// profiles/r01f: the MMA thread spent more time issuing than the tensor pipe spent computing).
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "elect.sync _|p, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap *tm) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(tm) : "memory");
}

// ---- tcgen05 wrappers ---------------------------------------------------------------------------
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_alloc(uint32_t *slot, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T, one 128 x N x 16 step
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
// same with A read from tensor memory (lane = row, 32-bit column j = K elements 2j, 2j+1): no shared-memory traffic for A
__device__ __forceinline__ void umma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(tmem_d), "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
// 16 consecutive 32-bit columns of this thread's TMEM lane (lane = 32*(warp%4) + laneid)
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
        ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]),
          "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
        : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
// arrive on `bar` when every tcgen05.mma issued so far by this thread has completed
__device__ __forceinline__ void umma_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// the same arrive delivered to the barrier at this offset in every CTA of `mask`
__device__ __forceinline__ void umma_commit_multicast(uint64_t *bar, uint16_t mask) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"(mask) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
          "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// Shared-memory matrix descriptor, K-major, SWIZZLE_128B, dense 128-byte rows (cute/arch/mma_sm100_desc.hpp
// field layout): start>>4 [0,14), LBO>>4 [16,30) = 1 (unused for swizzled K-major), SBO>>4 [32,46) = 64
// (1024 B between 8-row groups), version [46,48) = 1, layout type [61,64) = 2.
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t saddr) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
// Instruction descriptor, kind::f16: D F32 (bit 4), A/B F16 (0), both K-major (0), N>>3 at [17,23), M>>4 at [24,29)
__device__ __forceinline__ uint32_t umma_idesc_f16(int n) { return (1u << 4) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24); }

// ---- FP16 pair arithmetic on raw 32-bit registers ---------------------------------------------------
__device__ __forceinline__ uint32_t h2_sub(uint32_t a, uint32_t b) {
    __half2 r = __hsub2(*reinterpret_cast<__half2 *>(&a), *reinterpret_cast<__half2 *>(&b));
    return *reinterpret_cast<uint32_t *>(&r);
}
__device__ __forceinline__ uint32_t h2_mul(uint32_t a, uint32_t b) {
    __half2 r = __hmul2(*reinterpret_cast<__half2 *>(&a), *reinterpret_cast<__half2 *>(&b));
    return *reinterpret_cast<uint32_t *>(&r);
}
__device__ __forceinline__ uint32_t h2_fma(uint32_t a, uint32_t b, uint32_t c) {
    __half2 r = __hfma2(*reinterpret_cast<__half2 *>(&a), *reinterpret_cast<__half2 *>(&b), *reinterpret_cast<__half2 *>(&c));
    return *reinterpret_cast<uint32_t *>(&r);
}
__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
    uint32_t r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
    return r;
}
__device__ __forceinline__ uint4 lds128(const uint8_t *p) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(smem_u32(p)));
    return v;
}
__device__ __forceinline__ void sts128(uint8_t *p, uint32_t x, uint32_t y, uint32_t z, uint32_t w) {
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(smem_u32(p)), "r"(x), "r"(y), "r"(z), "r"(w) : "memory");
}

// ---- raw block -> registers ---------------------------------------------------------------------------
// The row piece of one 256-K chunk starts 16-byte aligned in the raw ring; block BP (0..7) of it starts
// at byte BP*block_bytes, which is only 2-byte (Q8_0, Q4_0) or 4-byte (Q4_1) aligned.  Every block lies
// inside 3 (Q8_0) or 2 (Q4_x) aligned 16-byte windows, read with conflict-free LDS.128 (odd row stride in
// 16-byte units) and re-aligned with compile-time funnel shifts.
template <int NV>
__device__ __forceinline__ void load_window(const uint8_t *p, uint32_t *w) {
#pragma unroll
    for (int i = 0; i < NV; i++) {
        const uint4 v = lds128(p + 16 * i);
        w[4 * i + 0] = v.x; w[4 * i + 1] = v.y; w[4 * i + 2] = v.z; w[4 * i + 3] = v.w;
    }
}
// 32-bit word at byte offset OFF (even) of the window
template <int OFF>
__device__ __forceinline__ uint32_t word_at(const uint32_t *w) {
    if constexpr (OFF % 4 == 0) return w[OFF / 4];
    else return __funnelshift_r(w[OFF / 4], w[OFF / 4 + 1], 16);
}
template <int OFF>
__device__ __forceinline__ uint32_t half_at(const uint32_t *w) {   // 16-bit value at byte offset OFF, replicated into both halves
    if constexpr (OFF % 4 == 0) return prmt(w[OFF / 4], 0, 0x1010);
    else return prmt(w[OFF / 4], 0, 0x3232);
}

// FP16 bit patterns (positive halves order like integers; inf / NaN sort above every finite value)
constexpr uint32_t H_512 = 0x6000u, H_2048 = 0x6800u, H_4096 = 0x6C00u, H_16384 = 0x7400u;

// hi = round_fp16(q * d [+ m]) and, for AP == 2, lo = the residual.  For Q8_0 / Q4_0 the residual q*d - hi is an FP16
// number (see the header), so one HFMA2 delivers it exactly; for Q4_1 (q*d + m has up to 26 significant bits) it is
// taken in F32 and rounded to FP16: hi + lo then carries >= 21 bits.
template <int TYPE, int AP>
__device__ __forceinline__ void emit_pair(uint32_t q, uint32_t d2, uint32_t m2, uint32_t &hi, uint32_t &lo) {
    if constexpr (TYPE == T_Q4_1) {
        hi = h2_fma(q, d2, m2);
        if constexpr (AP == 2) {
            const float2 qf = __half22float2(*reinterpret_cast<const __half2 *>(&q));
            const float2 hf = __half22float2(*reinterpret_cast<const __half2 *>(&hi));
            const float df = __low2float(*reinterpret_cast<const __half2 *>(&d2)), mf = __low2float(*reinterpret_cast<const __half2 *>(&m2));
            const __half2 r = __floats2half2_rn(__fsub_rn(fmaf(qf.x, df, mf), hf.x), __fsub_rn(fmaf(qf.y, df, mf), hf.y));
            lo = *reinterpret_cast<const uint32_t *>(&r);
        }
    } else {
        hi = h2_mul(q, d2);
        if constexpr (AP == 2) lo = h2_fma(q, d2, hi ^ 0x80008000u);
    }
}

// One block (32 K elements) of a row -> 16 registers of FP16 pairs in K order (the A operand's TMEM row format),
// hi part and (AP == 2) lo part.  Returns true if the block's scale could take d*q outside the FP16 range.
// GG: canonical ggml / GGUF nibble order (byte j = elements j and j+16, ggml/src/ggml-quants.c:1524-1530) instead of the
// Kotlin accessors' (byte j = elements 2j and 2j+1); Q8_0 blocks are the same in both.
template <int TYPE, int BP, bool GG, int AP>
__device__ __forceinline__ bool dequant_block(const uint8_t *rowp, uint32_t (&out)[16], uint32_t (&out_lo)[16]) {
    constexpr int OFF = BP * block_bytes(TYPE);
    constexpr int S = OFF & 15;
    const uint8_t *win = rowp + (OFF & ~15);
    if constexpr (TYPE == T_Q8_0) {
        uint32_t w[12];
        load_window<3>(win, w);
        const uint32_t d2 = half_at<S>(w);
#pragma unroll
        for (int c = 0; c < 4; c++) {
#pragma unroll
            for (int h = 0; h < 2; h++) {
                const uint32_t x = word_at<S + 2>(w + 2 * c + h) ^ 0x80808080u;
                // bytes -> 0x6400 | (q + 128) = 1024 + q + 128, exact; minus 1152 -> q; times d: one rounding
                emit_pair<TYPE, AP>(h2_sub(prmt(x, 0x64646464u, 0x4140), 0x64806480u), d2, 0u, out[4 * c + 2 * h + 0], out_lo[4 * c + 2 * h + 0]);
                emit_pair<TYPE, AP>(h2_sub(prmt(x, 0x64646464u, 0x4342), 0x64806480u), d2, 0u, out[4 * c + 2 * h + 1], out_lo[4 * c + 2 * h + 1]);
            }
        }
        return (d2 & 0x7FFFu) > H_512;                 // |d| * 128 must stay below 65504
    } else {
        uint32_t w[8];
        load_window<2>(win, w);
        constexpr int QOFF = S + (TYPE == T_Q4_0 ? 2 : 4);
        const uint32_t d2 = half_at<S>(w);
        uint32_t m2 = 0;
        if constexpr (TYPE == T_Q4_1) m2 = half_at<S + 2>(w);
#pragma unroll
        for (int c = 0; c < 4; c++) {
            const uint32_t q = word_at<QOFF>(w + c);
            const uint32_t lo = q & 0x0F0F0F0Fu, hi = (q >> 4) & 0x0F0F0F0Fu;
            if constexpr (GG) {
                // quant word c: low nibbles = elements 4c..4c+3 -> out[2c], out[2c+1]; high nibbles = elements 16+4c.. -> out[8+2c], out[9+2c]
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const uint32_t src = (j & 2) ? hi : lo;
                    const uint32_t p = prmt(src, 0u, (j & 1) ? 0xB3A2u : 0x9180u) | 0x64006400u;   // (1024 + b0, 1024 + b1) / (b2, b3)
                    const int oi = ((j & 2) ? 8 : 0) + 2 * c + (j & 1);
                    emit_pair<TYPE, AP>(h2_sub(p, TYPE == T_Q4_0 ? 0x64086408u : 0x64006400u), d2, m2, out[oi], out_lo[oi]);
                }
                continue;
            }
#pragma unroll
            for (int j = 0; j < 4; j++) {
                // byte j -> elements (2j, 2j+1) = (low nibble, high nibble), core/GGMLTypes.kt nibble order;
                // selector nibbles with bit 3 set replicate the (zero) sign bit -> 0x00
                const uint32_t sel = (uint32_t)j | ((8u | j) << 4) | ((4u + j) << 8) | ((8u | j) << 12);
                const uint32_t p = prmt(lo, hi, sel) | 0x64006400u;   // (1024 + lo, 1024 + hi)
                // Q4_0: (nib - 8) * d;  Q4_1: nib * d + m, one rounding
                emit_pair<TYPE, AP>(h2_sub(p, TYPE == T_Q4_0 ? 0x64086408u : 0x64006400u), d2, m2, out[4 * c + j], out_lo[4 * c + j]);
            }
        }
        if constexpr (TYPE == T_Q4_0) return (d2 & 0x7FFFu) > H_4096;             // |d| * 8
        else return (d2 & 0x7FFFu) > H_2048 || (m2 & 0x7FFFu) > H_16384;          // |d| * 15 + |m|
    }
}

struct GemmTcParams {
    const uint8_t *a;      // packed weight rows
    const uint8_t *ws;     // FP16 activation tiles from pack_b_kernel: [K/64][BPARTS][n_pad][128 B]
    const float *cscale;   // [n_pad] per-column factor that undoes the power-of-two scaling of b (pack_b_kernel)
    const float *b;        // the F32 activations (read only by the F32 recomputation of an out-of-range tile)
    float *dst;
    int M, N, K;
    int row_bytes;         // K/32 * block_bytes
    int nt, bn;            // column tiles per row tile, columns per tile (multiple of 16, <= 256)
    int cs;                // cluster size: CTAs with consecutive row tiles and the same column tile share the B tile by multicast
    int n_pad;             // N rounded up to 32: rows of one K-step slab in ws
    int stages;
    int tmem_cols;
    int vec_ok;            // dst rows 16-byte aligned and N % 4 == 0
    int nib;
    unsigned long long *dbg;   // tools/gemm_tc_timeline.cu only (-DLK_TC_TIMELINE): 16 counters per CTA, else NULL
};

#ifdef LK_TC_TIMELINE
#define TC_T0() const long long t0_ = clock64()
#define TC_ACC(slot) do { if (p.dbg && lane == 0) p.dbg[(size_t)blockIdx.x * 16 + (slot)] += (unsigned long long)(clock64() - t0_); } while (0)
#define TC_MARK(slot) do { if (p.dbg && lane == 0) p.dbg[(size_t)blockIdx.x * 16 + (slot)] = (unsigned long long)(clock64() - t_begin); } while (0)
#else
#define TC_T0() do { } while (0)
#define TC_ACC(slot) do { } while (0)
#define TC_MARK(slot) do { } while (0)
#endif

#ifdef LK_TC_TIMELINE
#define g_tc_dbg_slot dbg_slot
#endif
// 64-K step J (0..3) of the current raw chunk: wait for the A stage (32 * AP TMEM columns) to be free, convert this
// thread's block, store it to the thread's TMEM lane, publish the stage to the MMA thread.
template <int TYPE, int J, bool GG, int AP>
__device__ __forceinline__ bool dequant_step(const uint8_t *rowp, uint32_t tmem_a_lane, uint64_t *full, uint64_t *a_empty, int qb,
                                             int lane, int stages, int &s, int &it, unsigned long long *dbg_slot = nullptr) {
    uint32_t o[16], ol[16];
    bool oor;
    if (qb) oor = dequant_block<TYPE, 2 * J + 1, GG, AP>(rowp, o, ol);
    else oor = dequant_block<TYPE, 2 * J, GG, AP>(rowp, o, ol);
#ifdef LK_TC_TIMELINE
    const long long tw_ = clock64();
#endif
    if (it > 0) mbar_wait(&a_empty[s], (it - 1) & 1);
#ifdef LK_TC_TIMELINE
    if (g_tc_dbg_slot && lane == 0) *g_tc_dbg_slot += (unsigned long long)(clock64() - tw_);
#endif
    tc_fence_after();
    const uint32_t col = tmem_a_lane + (uint32_t)(s * (TC_BK / 2) * AP + qb * 16);
    tmem_st16(col, o);
    if constexpr (AP == 2) tmem_st16(col + TC_BK / 2, ol);
    tmem_st_wait();
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(&full[s]);
    s += TC_DQ_GROUPS;                        // this warp's next step
    if (s >= stages) { s -= stages; it++; }
    return oor;
}

// The reference's own arithmetic for one weight element (core/GGMLComputeOps.kt:43-145 through the accessors of
// core/GGMLTypes.kt:543-732): used only by the F32 recomputation of a tile whose scales leave the FP16 range.
template <int TYPE>
__device__ __forceinline__ float tc_ref_weight(const uint8_t *a, int64_t flat, int nib) {
    constexpr int BB = block_bytes(TYPE);
    const uint8_t *p = a + (flat >> 5) * BB;
    const int item = (int)(flat & 31);
    const float d = h2f(*(const uint16_t *)p);
    if (TYPE == T_Q8_0) return __fmul_rn(d, (float)(int8_t)p[2 + item]);
    constexpr int QOFF = (TYPE == T_Q4_0) ? 2 : 4;
    const uint8_t byte = p[QOFF + (nib ? (item & 15) : (item >> 1))];
    const float q = (float)((nib ? (item >> 4) : (item & 1)) ? (byte >> 4) : (byte & 15));
    if (TYPE == T_Q4_0) return __fmul_rn(d, __fsub_rn(q, 8.0f));
    return __fadd_rn(__fmul_rn(d, q), h2f(*(const uint16_t *)(p + 2)));
}

// ---- the GEMM -------------------------------------------------------------------------------------
// Barriers: full[s] (1 producer arrive carrying the B bytes + 8 dequant warps), a_empty[s] (local tcgen05.commit:
// the A columns in TMEM are free), b_empty[s] (one commit from every CTA of the cluster: the B stage, which every
// CTA of the cluster writes a slice of, is free everywhere), raw_full[2] (producer arrive carrying 128 row pieces),
// raw_empty[2] (8 dequant warps), acc_full (commit).
// AP / BPARTS: FP16 parts of the A / B operand (1 or 2, see the header): AP * BPARTS - (AP == 2 && BPARTS == 2) MMAs per
// 16-K step into the same accumulator (the lo x lo product, 2^-22 of the term, is dropped).
template <int TYPE, bool GG, int AP, int BPARTS>
__global__ void __launch_bounds__(TC_THREADS, 1) gemm_tc_kernel(const GemmTcParams p, const __grid_constant__ CUtensorMap tm_a) {
    extern __shared__ uint8_t smem_dyn[];
    uint8_t *sm = smem_dyn + ((1024u - (smem_u32(smem_dyn) & 1023u)) & 1023u);   // SWIZZLE_128B atoms need 1024-byte alignment
    const int stages = p.stages;
    const int b_part = p.bn * 128;                 // one FP16 part of a B stage
    const int b_stride = b_part * BPARTS;
    uint8_t *sB = sm;
    uint8_t *sRaw = sB + stages * b_stride;
    uint64_t *full = reinterpret_cast<uint64_t *>(sRaw + 2 * TC_BM * Raw<TYPE>::stride);
    uint64_t *a_empty = full + TC_MAX_STAGES;
    uint64_t *b_empty = a_empty + TC_MAX_STAGES;
    uint64_t *raw_full = b_empty + TC_MAX_STAGES;
    uint64_t *raw_empty = raw_full + 2;
    uint64_t *acc_full = raw_empty + 2;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(acc_full + 1);
    volatile int *oor_flag = reinterpret_cast<volatile int *>(tmem_slot + 1);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
#ifdef LK_TC_TIMELINE
    const long long t_begin = clock64();
#endif
    const int cs = p.cs;
    const int crank = cs > 1 ? (int)cluster_ctarank() : 0;
    const int group = (int)blockIdx.x / cs;             // clusters are consecutive blockIdx.x
    const int m0 = ((group / p.nt) * cs + crank) * TC_BM;   // may lie past M (odd tile count): computed, never stored
    const int n0 = (group % p.nt) * p.bn;
    const int bn = min(p.bn, p.n_pad - n0);          // columns of this tile (multiple of 32)
    const int nkb = p.K / TC_BK, nchunks = p.K / TC_CHUNK_K;

    if (tid == 0) {
        for (int i = 0; i < stages; i++) { mbar_init(&full[i], 1 + TC_DQ_GROUP); mbar_init(&a_empty[i], 1); mbar_init(&b_empty[i], cs); }
        for (int i = 0; i < 2; i++) { mbar_init(&raw_full[i], 1); mbar_init(&raw_empty[i], TC_DQ_WARPS); }
        mbar_init(acc_full, 1);
        *oor_flag = 0;
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, (uint32_t)p.tmem_cols);
    tc_fence_before();
    if (cs > 1) cluster_sync_all();     // peers multicast into this CTA's stages and arrive on its barriers
    else __syncthreads();
    tc_fence_after();
    const uint32_t tmem_acc = *tmem_slot;

    if (warp == 0) {
        // ===== activation-tile producer (paced by the MMA through b_empty[]) =====
        pdl_wait();                                     // the tiles are pack_b_kernel's output (programmatic dependent launch)
        const uint64_t pol = l2_policy_evict_last();   // re-read by every row tile out of the 126 MB L2
        const uint32_t part_bytes = (uint32_t)bn * 128u;
        const uint32_t slice = part_bytes / (uint32_t)cs;       // rows [crank*bn/cs, +bn/cs): whole 8-row swizzle atoms
        const uint16_t mask = (uint16_t)((1u << cs) - 1u);
        int s = 0, it = 0;
        for (int kb = 0; kb < nkb; kb++) {
            { TC_T0(); if (it > 0) mbar_wait(&b_empty[s], (it - 1) & 1); TC_ACC(2); }
            if (elect_one()) {
                mbar_arrive_expect_tx(&full[s], part_bytes * BPARTS);   // this CTA's stage receives one slice per part from every CTA of the cluster
#pragma unroll
                for (int part = 0; part < BPARTS; part++) {
                    uint8_t *dstp = sB + (size_t)s * b_stride + (size_t)part * b_part + (size_t)crank * slice;
                    const uint8_t *srcp = p.ws + (((size_t)kb * BPARTS + part) * p.n_pad + n0) * 128 + (size_t)crank * slice;
                    if (cs > 1) tma_load_1d_multicast(dstp, srcp, slice, &full[s], mask, pol);
                    else tma_load_1d(dstp, srcp, slice, &full[s], pol);
                }
            }
            __syncwarp();
            if (++s == stages) { s = 0; it++; }
        }
    } else if (warp == 2) {
        // ===== raw-weight producer (paced only by the dequant warps through raw_empty[]: runs two chunks ahead) =====
        const uint64_t pol = l2_policy_evict_last();   // re-read by the other column tiles
        if (elect_one()) tma_prefetch_desc(&tm_a);
        for (int c = 0; c < nchunks; c++) {
            const int rs = c & 1;
            { TC_T0(); if (c >= 2) mbar_wait(&raw_empty[rs], ((c >> 1) - 1) & 1); TC_ACC(3); }
            if (elect_one()) {
                mbar_arrive_expect_tx(&raw_full[rs], (uint32_t)(TC_BM * Raw<TYPE>::stride));
                tma_load_2d(sRaw + (size_t)rs * TC_BM * Raw<TYPE>::stride, &tm_a, c * (Raw<TYPE>::bytes / 2), m0, &raw_full[rs], pol);
            }
            __syncwarp();
        }
    } else if (warp == 1) {
        // ===== MMA issuer: the warp stays converged, one elected lane issues =====
        const uint32_t idesc = umma_idesc_f16(bn);
        const uint16_t mask = (uint16_t)((1u << cs) - 1u);
        int s = 0, it = 0;
        for (int kb = 0; kb < nkb; kb++) {
            if (kb == 0) TC_MARK(7);
            { TC_T0(); mbar_wait(&full[s], it & 1); TC_ACC(1); }
            if (kb == 0) TC_MARK(8);
            tc_fence_after();
            if (elect_one()) {
                const uint32_t ta = tmem_acc + TC_A_COL0 + (uint32_t)(s * (TC_BK / 2) * AP);
                const uint64_t db = umma_desc_sw128(smem_u32(sB + (size_t)s * b_stride));
                const uint64_t db_lo = umma_desc_sw128(smem_u32(sB + (size_t)s * b_stride + b_part));
#pragma unroll
                for (int k = 0; k < TC_BK / 16; k++) {   // per 16-K step: +8 TMEM columns of A, +32 bytes inside B's swizzle row (+2 in the address field)
                    // small products first: they meet the accumulator before the large one of the same step
                    if constexpr (AP == 2) umma_f16_ts(tmem_acc, ta + TC_BK / 2 + 8 * k, db + 2 * k, idesc, (uint32_t)((kb | k) != 0));
                    if constexpr (BPARTS == 2) umma_f16_ts(tmem_acc, ta + 8 * k, db_lo + 2 * k, idesc, (uint32_t)(AP == 2 || (kb | k) != 0));
                    umma_f16_ts(tmem_acc, ta + 8 * k, db + 2 * k, idesc, (uint32_t)(AP == 2 || BPARTS == 2 || (kb | k) != 0));
                }
                umma_commit(&a_empty[s]);
                if (cs > 1) umma_commit_multicast(&b_empty[s], mask);
                else umma_commit(&b_empty[s]);
                if (kb == nkb - 1) umma_commit(acc_full);
            }
            __syncwarp();
            if (++s == stages) { s = 0; it++; }
        }
    } else {
        // ===== dequant warps: raw blocks -> FP16 A stage; then the epilogue =====
        // a warp may only touch TMEM lanes 32*(warp%4)..+31, and TMEM lane = tile row: that fixes the row of each
        // thread; the two warps sharing a lane group take the two 32-K halves (qb) of every 64-K step
        const int g = warp & 3, qb = ((warp - TC_CTRL_WARPS) >> 2) & 1, par = (warp - TC_CTRL_WARPS) >> 3;
        const int row = g * 32 + lane;
        const uint32_t tmem_a_lane = tmem_acc + ((uint32_t)(g * 32) << 16) + TC_A_COL0;
        unsigned long long *dq_dbg = nullptr;
#ifdef LK_TC_TIMELINE
        if (p.dbg && warp == TC_CTRL_WARPS) dq_dbg = p.dbg + (size_t)blockIdx.x * 16 + 5;
#endif
        int s = par, it = 0;          // first step of this group: 64-K step `par`
        bool oor = false;             // a scale of this thread's row leaves the FP16 range
        for (int c = 0; c < nchunks; c++) {
            const int rs = c & 1;
            { TC_T0(); mbar_wait(&raw_full[rs], (c >> 1) & 1); if (warp == TC_CTRL_WARPS) TC_ACC(4); }
            const uint8_t *rowp = sRaw + (size_t)(rs * TC_BM + row) * Raw<TYPE>::stride;
            if (TC_DQ_GROUPS == 1 || par == 0) {
                oor |= dequant_step<TYPE, 0, GG, AP>(rowp, tmem_a_lane, full, a_empty, qb, lane, stages, s, it, dq_dbg);
                if (TC_DQ_GROUPS == 1) oor |= dequant_step<TYPE, 1, GG, AP>(rowp, tmem_a_lane, full, a_empty, qb, lane, stages, s, it, dq_dbg);
                oor |= dequant_step<TYPE, 2, GG, AP>(rowp, tmem_a_lane, full, a_empty, qb, lane, stages, s, it, dq_dbg);
                if (TC_DQ_GROUPS == 1) oor |= dequant_step<TYPE, 3, GG, AP>(rowp, tmem_a_lane, full, a_empty, qb, lane, stages, s, it, dq_dbg);
            } else {
                oor |= dequant_step<TYPE, 1, GG, AP>(rowp, tmem_a_lane, full, a_empty, qb, lane, stages, s, it, dq_dbg);
                oor |= dequant_step<TYPE, 3, GG, AP>(rowp, tmem_a_lane, full, a_empty, qb, lane, stages, s, it, dq_dbg);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&raw_empty[rs]);
        }
        // epilogue: TMEM lane = output row; warp w may read lanes 32*(w%4)..+31
        if (warp == TC_CTRL_WARPS) TC_MARK(9);
        if (oor && m0 + row < p.M) *oor_flag = 1;              // rows past M are TMA zero fill
        pdl_wait();                                             // the column factors are pack_b_kernel's output
        mbar_wait(acc_full, 0);
        if (warp == TC_CTRL_WARPS) TC_MARK(10);
        tc_fence_after();
        asm volatile("bar.sync 2, %0;" ::"r"(TC_DQ_WARPS * 32) : "memory");   // every dequant thread has had its say on oor_flag
        const int half = qb + 2 * par;                         // the warps sharing a lane group split the columns
        const int ncol = bn / (2 * TC_DQ_GROUPS);              // multiple of 8
        const int grow = m0 + g * 32 + lane;
        float *drow = p.dst + (size_t)grow * p.N + n0;
        if (*oor_flag) {
            // F32 recomputation of this tile with the reference's operation order (sequential k, (d*q)*x rounded, then the
            // add): scales this large do not occur in model weights, so this is a correctness path, not a tuned one
            if (grow < p.M) {
                for (int cc = half * ncol; cc < (half + 1) * ncol; cc += 8) {
                    float acc[8];
#pragma unroll
                    for (int e = 0; e < 8; e++) acc[e] = 0.f;
                    for (int k = 0; k < p.K; k++) {
                        const float wv = tc_ref_weight<TYPE>(p.a, (int64_t)grow * p.K + k, p.nib);
                        const float *bp = p.b + (size_t)k * p.N + n0 + cc;
#pragma unroll
                        for (int e = 0; e < 8; e++)
                            if (n0 + cc + e < p.N) acc[e] = __fadd_rn(acc[e], __fmul_rn(wv, bp[e]));
                    }
#pragma unroll
                    for (int e = 0; e < 8; e++)
                        if (n0 + cc + e < p.N) drow[cc + e] = acc[e];
                }
            }
        } else {
        const float *csc = p.cscale + n0;
        // 16 columns per tcgen05.ld, two loads in flight before the wait: the TMEM read latency is paid once per 32 columns
        for (int cc = half * ncol; cc < (half + 1) * ncol; cc += 32) {
            uint32_t v[2][16];
            const bool second = cc + 16 < (half + 1) * ncol;       // ncol is a multiple of 16, not always of 32 (warp-uniform)
            tmem_ld16(tmem_acc + ((uint32_t)(g * 32) << 16) + (uint32_t)cc, v[0]);
            if (second) tmem_ld16(tmem_acc + ((uint32_t)(g * 32) << 16) + (uint32_t)(cc + 16), v[1]);
            tmem_ld_wait();
#pragma unroll
            for (int hh = 0; hh < 2; hh++) {
                if (hh == 1 && !second) break;
                const int c0 = cc + 16 * hh;
                float f[16];
#pragma unroll
                for (int q = 0; q < 4; q++) {                   // undo the columns' power-of-two scaling (exact)
                    const float4 sc = __ldg(reinterpret_cast<const float4 *>(csc + c0) + q);
                    f[4 * q] = __fmul_rn(__uint_as_float(v[hh][4 * q]), sc.x); f[4 * q + 1] = __fmul_rn(__uint_as_float(v[hh][4 * q + 1]), sc.y);
                    f[4 * q + 2] = __fmul_rn(__uint_as_float(v[hh][4 * q + 2]), sc.z); f[4 * q + 3] = __fmul_rn(__uint_as_float(v[hh][4 * q + 3]), sc.w);
                }
                if (grow < p.M) {
                    if (p.vec_ok && n0 + c0 + 16 <= p.N) {
#pragma unroll
                        for (int q = 0; q < 4; q++)
                            *reinterpret_cast<float4 *>(drow + c0 + 4 * q) = make_float4(f[4 * q], f[4 * q + 1], f[4 * q + 2], f[4 * q + 3]);
                    } else {
#pragma unroll
                        for (int e = 0; e < 16; e++)
                            if (n0 + c0 + e < p.N) drow[c0 + e] = f[e];
                    }
                }
            }
        }
        }
        tc_fence_before();
        if (warp == TC_CTRL_WARPS) TC_MARK(11);
    }
    if (cs > 1) cluster_sync_all();     // no CTA exits while a peer may still write its stages or barriers
    else __syncthreads();
    if (warp == 1) {
        __syncwarp();
        TC_MARK(12);
        tc_fence_after();
        tmem_dealloc(tmem_acc, (uint32_t)p.tmem_cols);
    }
}

// ---- activations: F32 [K][N] (N fastest, core/GGMLComputeOps.kt:1474 b index j + k*N) -> FP16 tiles ----
// Range: column n is multiplied by 2^(13 - e_n), e_n = exponent of max_k |b[k][n]| (exact; the largest element lands in
// [2^13, 2^14), the FP16 maximum is 65504), and the epilogue multiplies the accumulator by cscale[n] = 2^(e_n - 13).
// colmax_kernel takes the column maxima of |x| as unsigned bit patterns (finite < inf < NaN in that order; a
// non-finite column keeps factor 1 and poisons its outputs through the MMA like the reference's F32 arithmetic does).
__global__ void __launch_bounds__(256) colmax_kernel(const float *__restrict__ b, uint32_t *__restrict__ cmax, int K, int N) {
    __shared__ uint32_t part[8][32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n = (int)blockIdx.x * 32 + lane;
    const int kper = (K + (int)gridDim.y - 1) / (int)gridDim.y;
    const int k0 = (int)blockIdx.y * kper, k1 = min(K, k0 + kper);
    uint32_t m = 0;
    if (n < N)
        for (int k = k0 + warp; k < k1; k += 8) m = max(m, __float_as_uint(b[(size_t)k * N + n]) & 0x7FFFFFFFu);
    part[warp][lane] = m;
    __syncthreads();
    if (warp == 0) {
#pragma unroll
        for (int w = 1; w < 8; w++) m = max(m, part[w][lane]);
        if (n < N && m) atomicMax(cmax + n, m);
    }
}

// ws layout: for every 64-K step kb and FP16 part (hi, lo) one slab of n_pad rows x 128 bytes; row n holds the part of
// b[kb*64 .. +64][n] as 8 chunks of 8 halfs, chunk c stored at position c ^ (n & 7) (the 128-byte swizzle the UMMA
// descriptor names).  A column tile [n0, n0+bn) of a K step is then one contiguous, 1024-byte aligned byte range per part.
// Rows N..n_pad are zero.  One thread per 16-byte chunk; a warp reads 32 consecutive n per k (coalesced).
template <int BPARTS>
__global__ void __launch_bounds__(256) pack_b_kernel(const float *__restrict__ b, uint8_t *__restrict__ ws, const uint32_t *__restrict__ cmax,
                                                     float *__restrict__ cscale, int K, int N, int n_pad) {
    pdl_wait();                // cmax is colmax_kernel's output
    pdl_launch_dependents();   // the GEMM may start its prologue and its weight stream now
    const int64_t total = (int64_t)(K / TC_BK) * 8 * n_pad;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int n = (int)(t % n_pad);
        const int c = (int)((t / n_pad) & 7);
        const int64_t kb = t / ((int64_t)n_pad * 8);
        uint32_t o[4] = {0, 0, 0, 0}, ol[4] = {0, 0, 0, 0};
        if (n < N) {
            // e = biased exponent of the column maximum, clamped so that both factors are normal F32 numbers
            const uint32_t mb = cmax[n];
            int e = (int)(mb >> 23);
            const bool plain = e == 255 || mb == 0;          // non-finite or all-zero column: factor 1
            e = e < 14 ? 14 : e;
            const float up = plain ? 1.0f : __uint_as_float((uint32_t)(267 - e) << 23);     // 2^(13 - (e - 127))
            if (kb == 0 && c == 0) cscale[n] = plain ? 1.0f : __uint_as_float((uint32_t)(e - 13) << 23);
            const float *src = b + ((size_t)kb * TC_BK + (size_t)c * 8) * N + n;
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const float x0 = __fmul_rn(src[(size_t)(2 * j) * N], up), x1 = __fmul_rn(src[(size_t)(2 * j + 1) * N], up);
                const __half2 h = __floats2half2_rn(x0, x1);
                o[j] = *reinterpret_cast<const uint32_t *>(&h);
                if constexpr (BPARTS == 2) {
                    const float2 hf = __half22float2(h);
                    const __half2 l = __floats2half2_rn(__fsub_rn(x0, hf.x), __fsub_rn(x1, hf.y));   // exact residual, rounded to 11 bits
                    ol[j] = *reinterpret_cast<const uint32_t *>(&l);
                }
            }
        } else if (kb == 0 && c == 0) {
            cscale[n] = 1.0f;
        }
        uint8_t *dstp = ws + (((size_t)kb * BPARTS) * n_pad + n) * 128 + ((c ^ (n & 7)) << 4);
        *reinterpret_cast<uint4 *>(dstp) = make_uint4(o[0], o[1], o[2], o[3]);
        if constexpr (BPARTS == 2) *reinterpret_cast<uint4 *>(dstp + (size_t)n_pad * 128) = make_uint4(ol[0], ol[1], ol[2], ol[3]);
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn tensor_map_encoder() {
    static EncodeTiledFn fn = [] {
        void *ptr = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
            ptr = nullptr;
        return (EncodeTiledFn)ptr;
    }();
    return fn;
}

// The packed weight as a 2-D uint16 tensor {row_bytes/2, M}; one box = the 256-K pieces of 128 rows.
cudaError_t make_weight_map(CUtensorMap *tm, const void *a, int64_t M, int row_bytes, int box_bytes) {
    EncodeTiledFn enc = tensor_map_encoder();
    if (!enc) return cudaErrorNotSupported;
    const cuuint64_t gdim[2] = {(cuuint64_t)(row_bytes / 2), (cuuint64_t)M};
    const cuuint64_t gstride[1] = {(cuuint64_t)row_bytes};
    const cuuint32_t box[2] = {(cuuint32_t)(box_bytes / 2), (cuuint32_t)TC_BM};
    const cuuint32_t estride[2] = {1, 1};
    const CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, const_cast<void *>(a), gdim, gstride, box, estride,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? cudaSuccess : cudaErrorInvalidValue;
}

template <int TYPE, bool GG, int AP, int BPARTS>
cudaError_t launch_tc(const GemmTcParams &p, int grid, size_t smem, cudaStream_t s) {
    CUtensorMap tm;
    cudaError_t me = make_weight_map(&tm, p.a, p.M, p.row_bytes, Raw<TYPE>::stride);
    if (me != cudaSuccess) return me;
    // the attribute belongs to the (function, device) pair: set it on every launch (cheap), never cache it per process
    cudaError_t e = cudaFuncSetAttribute(gemm_tc_kernel<TYPE, GG, AP, BPARTS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return e;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(TC_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)p.cs;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;     // prologue + weight stream under pack_b_kernel's tail
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 2;
    return cudaLaunchKernelEx(&cfg, gemm_tc_kernel<TYPE, GG, AP, BPARTS>, p, tm);
}

template <int TYPE, bool GG>
cudaError_t launch_tc_prec(const GemmTcParams &p, int grid, size_t smem, cudaStream_t s, int prec) {
    if (prec == 0) return launch_tc<TYPE, GG, 1, 1>(p, grid, smem, s);
    if (prec == 1) return launch_tc<TYPE, GG, 2, 1>(p, grid, smem, s);
    return launch_tc<TYPE, GG, 2, 2>(p, grid, smem, s);
}

int raw_stride(int type) { return type == T_Q8_0 ? Raw<T_Q8_0>::stride : type == T_Q4_0 ? Raw<T_Q4_0>::stride : Raw<T_Q4_1>::stride; }
size_t tc_fixed_smem(int type) { return 1024 + 2 * (size_t)TC_BM * raw_stride(type) + (3 * TC_MAX_STAGES + 5) * 8 + 16; }

}  // namespace

void *g_gemm_tc_debug = nullptr;   // set by tools/gemm_tc_timeline.cu; NULL in the library

// workspace: [K/64][parts][n_pad][128 B] FP16 tiles (1024-byte aligned), then n_pad column maxima (u32) and n_pad
// column factors (f32).  Sized for two parts whatever the mode, so a mode switch never reallocates.
size_t gemm_tc_workspace_bytes(int64_t K, int64_t N) {
    const int64_t n_pad = (N + 31) / 32 * 32;
    return (size_t)(K / TC_BK) * 2 * (size_t)n_pad * 128 + (size_t)n_pad * 8;
}

// Tile plan: BN = columns per CTA (multiple of 32, <= 256), chosen so that the tile count fills whole waves of
// SMs.  Cost model per 64-K step: tensor pipe 2*BN clk per MMA part, L2->SM (raw piece + this CTA's
// 1/cs slice of the BN*128-byte activation tile per part) at ~43 B/clk per SM, shared memory (B write+read, raw
// write+read) at 128 B/clk.  A plan needs at least 3 pipeline stages in shared memory.
static int plan_bn(int type, int64_t M, int64_t N, int sm_count, int cs, int ap, int bparts, int *nt_out) {
    const int64_t n_pad = (N + 31) / 32 * 32, mt = ((M + TC_BM - 1) / TC_BM + cs - 1) / cs * cs;
    const double raw = 128.0 * block_bytes(type) * 2;
    const int nmma = ap * bparts - (ap == 2 && bparts == 2 ? 1 : 0);
    double best = 1e30;
    int best_bn = 0, best_nt = 0;
    for (int nt = (int)((n_pad + 255) / 256); nt <= (int)(n_pad / 32) && nt <= 64; nt++) {
        const int bn = (int)(((n_pad + nt - 1) / nt + 31) / 32 * 32);
        if ((int64_t)bn * (nt - 1) >= n_pad) continue;   // last tile would be empty
        const int stages = (int)((227 * 1024 - tc_fixed_smem(type)) / ((size_t)bn * 128 * bparts));
        if (stages < 3 && bn > 32) continue;
        const double mma = 2.0 * bn * nmma, l2 = (raw + 128.0 * bn * bparts / cs) / 43.0, sm = (2 * raw + 128.0 * bn * (bparts + nmma)) / 128.0;
        const double step = mma > l2 ? (mma > sm ? mma : sm) : (l2 > sm ? l2 : sm);
        const int64_t waves = (mt * nt + sm_count - 1) / sm_count;
        const double cost = (double)waves * (step + 40.0);
        if (cost < best) { best = cost; best_bn = bn; best_nt = nt; }
    }
    *nt_out = best_nt;
    return best_bn;
}

cudaError_t launch_gemm_tc(int type, const void *a, int64_t K, int64_t M, const float *b, int64_t N, float *dst,
                           void *workspace, int sm_count, int cluster, cudaStream_t s, int nib, int precision) {
    if (type != T_Q8_0 && type != T_Q4_0 && type != T_Q4_1) return cudaErrorNotSupported;
    if (K <= 0 || K % TC_CHUNK_K || N < 2 || M < 1 || M > (1 << 30) || N > (1 << 20) || K > (1 << 24)) return cudaErrorNotSupported;
    if (((uintptr_t)a & 15) || ((uintptr_t)workspace & 1023) || ((uintptr_t)b & 3)) return cudaErrorNotSupported;
    const int prec = precision < 0 ? 0 : precision > 2 ? 2 : precision;
    const int ap = prec >= 1 ? 2 : 1, bparts = prec >= 2 ? 2 : 1;
    GemmTcParams p;
    p.dbg = (unsigned long long *)g_gemm_tc_debug;
    p.a = (const uint8_t *)a; p.ws = (const uint8_t *)workspace; p.dst = dst; p.b = b;
    p.M = (int)M; p.N = (int)N; p.K = (int)K; p.nib = nib;
    p.row_bytes = (int)(K / LK_QK) * block_bytes(type);
    p.n_pad = (int)((N + 31) / 32 * 32);
    uint32_t *cmax = (uint32_t *)((uint8_t *)workspace + (size_t)(K / TC_BK) * 2 * (size_t)p.n_pad * 128);
    float *cscale = (float *)(cmax + p.n_pad);
    p.cscale = cscale;
    p.cs = cluster == 4 ? 4 : cluster == 1 ? 1 : 2;
    p.bn = plan_bn(type, M, N, sm_count, p.cs, ap, bparts, &p.nt);
    if (p.bn < 32) return cudaErrorNotSupported;
    p.tmem_cols = 512;     // accumulator [0, 256) + A stages of 32 * ap columns; one CTA per SM, so no contention
    p.vec_ok = (((uintptr_t)dst & 15) == 0 && (N % 4) == 0) ? 1 : 0;
    const size_t fixed = tc_fixed_smem(type);
    const size_t per_stage = (size_t)p.bn * 128 * bparts;
    int stages = (int)((227 * 1024 - fixed) / per_stage);
    const int tmem_stages = (512 - TC_A_COL0) / (TC_BK / 2 * ap);
    if (stages > tmem_stages) stages = tmem_stages;
    if (stages > TC_MAX_STAGES) stages = TC_MAX_STAGES;
    if (stages < 2) return cudaErrorNotSupported;
    p.stages = stages;
    const size_t smem = fixed + stages * per_stage;

    // column maxima -> scaled FP16 tiles (+ column factors) -> GEMM; the last two launches overlap their predecessor's
    // tail (programmatic dependent launch)
    cudaError_t e = cudaMemsetAsync(cmax, 0, (size_t)p.n_pad * 4, s);
    if (e != cudaSuccess) return e;
    {
        const int gx = (int)((N + 31) / 32);
        int gy = (sm_count * 4 + gx - 1) / gx;
        if (gy > (int)(K / 64)) gy = (int)(K / 64);
        if (gy < 1) gy = 1;
        colmax_kernel<<<dim3((unsigned)gx, (unsigned)gy), 256, 0, s>>>(b, cmax, (int)K, (int)N);
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
    }
    const int64_t chunks = (K / TC_BK) * 8 * (int64_t)p.n_pad;
    int pgrid = (int)((chunks + 255) / 256);
    if (pgrid > sm_count * 16) pgrid = sm_count * 16;
    if (bparts == 2) e = launch_kernel(pack_b_kernel<2>, dim3((unsigned)pgrid), dim3(256), 0, s, true, b, (uint8_t *)workspace, (const uint32_t *)cmax, cscale, (int)K, (int)N, p.n_pad);
    else e = launch_kernel(pack_b_kernel<1>, dim3((unsigned)pgrid), dim3(256), 0, s, true, b, (uint8_t *)workspace, (const uint32_t *)cmax, cscale, (int)K, (int)N, p.n_pad);
    if (e != cudaSuccess) return e;
    const int grid = (int)((((M + TC_BM - 1) / TC_BM + p.cs - 1) / p.cs) * p.nt * p.cs);
    switch (type) {
        case T_Q8_0: return launch_tc_prec<T_Q8_0, false>(p, grid, smem, s, prec);
        case T_Q4_0: return nib ? launch_tc_prec<T_Q4_0, true>(p, grid, smem, s, prec) : launch_tc_prec<T_Q4_0, false>(p, grid, smem, s, prec);
        default: return nib ? launch_tc_prec<T_Q4_1, true>(p, grid, smem, s, prec) : launch_tc_prec<T_Q4_1, false>(p, grid, smem, s, prec);
    }
}

}  // namespace lk

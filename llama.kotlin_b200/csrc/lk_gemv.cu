// lk_gemv.cu — K1: decode GEMV  dst[i] = sum_k deq(a)[i,k] * b[k]   (N == 1)
//
// Replaces the M x 1 case of computeMatMul's Q4_0 / Q4_1 / Q8_0 x F32 branches
// (core/GGMLComputeOps.kt:1448-1527 -> computeDotProductQ40F32/Q41F32/Q80F32,
// :120-145, :70-115, :43-68) on the reference's own block layouts (34/18/20 B,
// Kotlin nibble order core/GGMLTypes.kt:647-651).
//
// The op is HBM-bound on paper (3.5 flop/B) but at 40 weight elements per
// clock per SM the CUDA-core instruction budget is what decides it on B200:
// measured pipe rates (tools/ubench_pipes.cu) are 0.5 warp-instr/clk/SMSP for
// PRMT/LOP3/SHF (ALU pipe) and for DP4A/FFMA2 (FMA pipe), the two pipes issue
// in parallel.  A nibble -> F32 conversion costs >= 1 ALU op per element
// (>= 16 clk per 8-nibble word per warp, more than the ~25 clk HBM allows once
// latencies are added), so the inner product runs on DP4A instead:
//
//  * activations -> 23-bit block fixed point, once per CTA: for each 64-float
//    column, E = exponent of max|x|, T_k = rint(x_k * 2^(22-E)) (|T_k| <= 2^22),
//    stored as three byte planes (two's-complement bytes of T: planes 0/1
//    unsigned, plane 2 signed).  |x_k - T_k 2^(E-22)| <= 2^(E-23): one bit less
//    than F32 for the largest element of the column, i.e. far below the F32
//    accumulation error of the reference's sequential sum;
//  * per 8-nibble word: AND / SHR+AND isolate the low / high nibbles as bytes
//    (3 ALU ops), 2 x 3 DP4A accumulate exact int32 plane sums (6 FMA-pipe ops)
//    => 12 clk per word per warp on the FMA pipe, ALU pipe half idle;
//  * per block: I = (I2 << 16) + (I1 << 8) + I0 is exact in int32 for Q4
//    (|I| < 2^31 by construction), one I2F, then F32: d * (s*I - 8*s*sumT)
//    (Q4_0), d*s*I + m*s*sumT (Q4_1), d*s*I (Q8_0, planes combined in F32);
//  * persistent grid, one CTA per SM per launch; CTA c owns a contiguous,
//    balanced row range = one contiguous byte range of the packed matrix, which
//    a producer warp streams through an NSTAGES-deep shared-memory ring with
//    1-D TMA bulk copies (cp.async.bulk -> UBLKCP) completing on mbarriers;
//    weights use an L2 evict_first policy (read once);
//  * consumer thread (rg, c) owns block-pair column c (64 consecutive k) for the
//    rows of row group rg: its 192 activation bytes live in 48 registers for
//    the whole kernel, so shared memory only carries weight bytes (9 / 10 / 17
//    conflict-free word loads per 64 elements);
//  * per-row partial sums go to shared memory (one conflict-free STS per row per
//    thread); the cross-lane / cross-warp reduction happens once at the end;
//  * programmatic dependent launch: two CTAs (this launch's and the next one's)
//    fit on an SM; the next GEMV's producer prefetches its weights while this one
//    computes.  wait_mode says where griddepcontrol.wait sits: before the weight
//    prefetch (weights written by an earlier op), before reading x (default), or
//    at the very end (the launch touches nothing an in-flight launch writes or
//    reads-after-write; the end wait keeps completion order = launch order).
//
// Numerics: integer dot products are exact; the only roundings are the 23-bit
// activation rounding and ~4 F32 ops per block.  Measured error vs the oracle is
// ~1e-6 relative (bound documented in DESIGN.md), far inside the 1e-3 bar.
#include <algorithm>

#include "lk_common.cuh"
#include "lk_kernels.h"

namespace lk {

constexpr int GEMV_MAX_STAGES = 8;
// activation slot of one block column: NP planes x 8 words, then sumT, scale, pad: 7 (NP=3) or 5 (NP=2) x 16 B, an odd
// number of 16-byte units => conflict-free LDS.128
__host__ __device__ constexpr int gemv_xslot_words(int np) { return np * 8 + 4; }
constexpr int GEMV_RED_PAD = 8;        // floats of padding per partial-sum row (bank spread for finish_rows)

__device__ __forceinline__ int dp4a_uu(uint32_t a, uint32_t b, int c) {
    int d;
    asm("dp4a.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ int dp4a_us(uint32_t a, uint32_t b, int c) {  // a unsigned bytes, b signed bytes
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ int dp4a_su(uint32_t a, uint32_t b, int c) {  // a signed bytes, b unsigned bytes
    int d;
    asm("dp4a.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ int dp4a_ss(uint32_t a, uint32_t b, int c) {
    int d;
    asm("dp4a.s32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// Activation registers of one thread (one 32-element block column): q[plane][word].  For Q4 types
// word 2j holds elements 8j+{0,2,4,6} (the low nibbles of quant word j) and word 2j+1 elements
// 8j+{1,3,5,7}; for Q8_0 word j holds elements 4j..4j+3.
template <int NP>
struct XRegs {
    uint32_t q[NP][8];
    float c;        // Q4_0: -8 * s * sumT;  Q4_1 / Q4_K: s * sumT;  Q8_0 / Q8_K: unused;  Q2_K: s * sumT of elements 0..15
    float s;        // 2^(E-22) (NP = 3) or 2^(E-14) (NP = 2)
    float c2;       // Q2_K only: s * sumT of elements 16..31 (the block column spans two 16-element sub-blocks)
    uint32_t aux;   // BITNET_1_58 only: shared-space address of this lane's copy of the byte -> five-trit table
    uint32_t aux_sh;   // ... and log2 of the bytes between consecutive entries of one copy (3..7)
};

// BitNet 1.58 decode table (one per CTA, in shared memory): entry v = the five base-3 digits of byte v, each as a signed
// byte digit - 1, twice: .x = (t0, t1, t2, t3), .y = (t1, t2, t3, t4).  With the tail of a five-digit group always in .y
// and its head in .x, every word of four CONSECUTIVE elements is one PRMT of two table words (or a table word as is).
#ifndef LK_BITNET_UNR
#define LK_BITNET_UNR 4
#endif
struct BitnetLut { uint32_t w[512]; };
constexpr BitnetLut make_bitnet_lut() {
    BitnetLut t{};
    for (uint32_t e = 0; e < 256; e++) {
        uint32_t v = e, d[5] = {0, 0, 0, 0, 0};
        for (int i = 0; i < 5; i++) { const uint32_t q = v / 3u; d[i] = (v - 3u * q - 1u) & 0xFFu; v = q; }
        t.w[2 * e] = d[0] | (d[1] << 8) | (d[2] << 16) | (d[3] << 24);
        t.w[2 * e + 1] = d[1] | (d[2] << 8) | (d[3] << 16) | (d[4] << 24);
    }
    return t;
}
__device__ const BitnetLut g_bitnet_lut = make_bitnet_lut();     // 2 KB, built by the compiler; every CTA copies it into shared memory
__device__ __forceinline__ uint2 lds_u2(uint32_t addr) {
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
    return v;
}

// 16 quant bytes (4 aligned words) of 4-bit quants x the block's activations; exact int32
template <int NP>
__device__ __forceinline__ int q4_block_isum(const uint32_t (&q)[4], const XRegs<NP> &x) {
    int acc[NP];
#pragma unroll
    for (int pl = 0; pl < NP; pl++) acc[pl] = 0;
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const uint32_t lo = q[j] & 0x0F0F0F0Fu, hi = (q[j] >> 4) & 0x0F0F0F0Fu;
#pragma unroll
        for (int pl = 0; pl < NP; pl++) {       // the top plane holds signed bytes
            acc[pl] = (pl == NP - 1) ? dp4a_us(lo, x.q[pl][2 * j], acc[pl]) : dp4a_uu(lo, x.q[pl][2 * j], acc[pl]);
            acc[pl] = (pl == NP - 1) ? dp4a_us(hi, x.q[pl][2 * j + 1], acc[pl]) : dp4a_uu(hi, x.q[pl][2 * j + 1], acc[pl]);
        }
    }
    // exact: |.| < 2^31 (nibble <= 15, |T| <= 2^22 resp. 2^14, 32 terms)
    if (NP == 3) return (acc[NP - 1] << 16) + ((acc[1] << 8) + acc[0]);
    return (acc[1] << 8) + acc[0];
}

// dot of one block (32 consecutive k of one row) with the thread's activations.
// row: 4-byte aligned shared-memory address of the row; b: block index within the row.
// 18- and 34-byte blocks start on a 4-byte boundary only for even b: the thread loads the words
// that cover its block and one PRMT per word (selector chosen by parity) realigns them.
template <int TYPE, int NP>
__device__ __forceinline__ float block_dot(const uint32_t *row, int b, const XRegs<NP> &x) {
    const bool odd = b & 1;
    if (TYPE == T_Q4_0) {
        const uint32_t *wp = row + 4 * b + (b >> 1);
        uint32_t w[5];
#pragma unroll
        for (int i = 0; i < 5; i++) w[i] = wp[i];
        const uint32_t sel = odd ? 0x7654u : 0x5432u;
        uint32_t q[4];
#pragma unroll
        for (int j = 0; j < 4; j++) q[j] = __byte_perm(w[j], w[j + 1], sel);
        const float d = h2f((uint16_t)(odd ? (w[0] >> 16) : w[0]));
        const float f = (float)q4_block_isum(q, x);
        return d * fmaf(f, x.s, x.c);                       // d * (s*I - 8*s*sumT)
    } else if (TYPE == T_Q4_1) {
        const uint32_t *wp = row + 5 * b;
        uint32_t w[5];
#pragma unroll
        for (int i = 0; i < 5; i++) w[i] = wp[i];
        uint32_t q[4];
#pragma unroll
        for (int j = 0; j < 4; j++) q[j] = w[j + 1];
        const float d = h2f((uint16_t)(w[0] & 0xFFFFu)), m = h2f((uint16_t)(w[0] >> 16));
        const float f = (float)q4_block_isum(q, x);
        return fmaf(d, f * x.s, m * x.c);                   // d*s*I + m*s*sumT
    } else if (TYPE == T_Q4_K) {
        // the reference's own Q4_K (core/GGMLComputeOps.kt:262-306): 144-byte super-block = 36 words: d | dmin << 16,
        // scales[12], then 8 sub-blocks of 16 quant bytes.  Block column b = sub-block b & 7 of super-block b >> 3; its
        // quants are one aligned LDS.128.  w = (q/15) * ((sc & 63)/63 * d) + (qmin/63 * d + dmin).
        const uint32_t *wp = row + 36 * (b >> 3);
        const int sb = b & 7;
        const uint4 qv = *(const uint4 *)(wp + 4 + 4 * sb);
        const uint32_t q[4] = {qv.x, qv.y, qv.z, qv.w};
        const uint32_t hdr = wp[0];
        const uint32_t sbyte = (wp[1 + (sb >> 2)] >> (8 * (sb & 3))) & 0xFFu;
        const int hi_i = 2 * sb + 1;                                   // scales[2 sb + 1] & 15 for sb <= 5, else 0
        const uint32_t hbyte = sb <= 5 ? ((wp[1 + (hi_i >> 2)] >> (8 * (hi_i & 3))) & 0x0Fu) : 0u;
        const float d = h2f((uint16_t)(hdr & 0xFFFFu)), dmin = h2f((uint16_t)(hdr >> 16));
        const float scale = (float)(sbyte & 63u) * d * (1.0f / (63.0f * 15.0f));
        const float mn = fmaf((float)((sbyte >> 6) | (hbyte << 2)) * (1.0f / 63.0f), d, dmin);
        const float f = (float)q4_block_isum(q, x);
        return fmaf(scale, f * x.s, mn * x.c);                          // scale*s*I + min*s*sumT
    } else if (TYPE == T_Q2_K) {
        // the reference's own Q2_K (core/GGMLComputeOps.kt:152-236): 84-byte super-block = 21 words: scales[16] (lo nibble
        // scale, hi nibble min), 64 quant bytes (byte j = elements 4j..4j+3, 2 bits each), d | dmin << 16.  Block column
        // b = sub-blocks 2 (b & 7) and 2 (b & 7) + 1 (16 elements, one quant word, own scale / min each).
        // w = (q/3) * ((sc & 15)/15 * d) + ((sc >> 4) * d + dmin); plane word 4 h + j holds elements 16 h + {j, 4+j, 8+j, 12+j}.
        const uint32_t *wp = row + 21 * (b >> 3);
        const int sb = b & 7;
        const uint32_t hdr = wp[20];
        const uint32_t sc2 = (wp[sb >> 1] >> (16 * (sb & 1))) & 0xFFFFu;
        const float d = h2f((uint16_t)(hdr & 0xFFFFu)), dmin = h2f((uint16_t)(hdr >> 16));
        float r = 0.f;
#pragma unroll
        for (int h = 0; h < 2; h++) {
            const uint32_t w = wp[4 + 2 * sb + h];
            int acc[NP];
#pragma unroll
            for (int pl = 0; pl < NP; pl++) acc[pl] = 0;
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const uint32_t q = (w >> (2 * j)) & 0x03030303u;
#pragma unroll
                for (int pl = 0; pl < NP; pl++)
                    acc[pl] = (pl == NP - 1) ? dp4a_us(q, x.q[pl][4 * h + j], acc[pl]) : dp4a_uu(q, x.q[pl][4 * h + j], acc[pl]);
            }
            const int I = (NP == 3) ? (acc[NP - 1] << 16) + ((acc[1] << 8) + acc[0]) : (acc[1] << 8) + acc[0];
            const uint32_t sm = (sc2 >> (8 * h)) & 0xFFu;
            const float scale = (float)(sm & 15u) * d * (1.0f / 45.0f);
            const float mn = fmaf((float)(sm >> 4), d, dmin);
            r += fmaf(scale, (float)I * x.s, mn * (h ? x.c2 : x.c));
        }
        return r;
    } else if (TYPE == T_BITNET_1_58) {
        // the reference's BitNet 1.58 block (core/GGMLTypes.kt:920-1000, computeDotProductBitNet158F32
        // core/GGMLComputeOps.kt:635-660): 10 bytes = d (f16), 7 bytes of 5 base-3 digits (element 5g + p = digit p of byte
        // g, w = d * (digit - 1)), 1 unused.  Word-aligned for even b only.  Each digit byte goes through the CTA's table
        // (7 LDS.64), 5 PRMTs put the trits in element order, and the block is a signed-byte dot product against the
        // activation planes like Q8_0: exact int32, one I2F.  A byte >= 243 (leading digit 3: the reference throws
        // IllegalStateException, core/GGMLTypes.kt:975-981) makes the product NaN.
        const uint32_t *wp = row + 2 * b + (b >> 1);
        const uint32_t w0 = wp[0], w1 = wp[1], w2 = wp[2];
        const uint32_t q0 = odd ? __byte_perm(w0, w1, 0x5432) : w0;      // d | B0 << 16 | B1 << 24
        const uint32_t q1 = odd ? __byte_perm(w1, w2, 0x5432) : w1;      // B2 .. B5
        const uint32_t q2 = odd ? (w2 >> 16) : w2;                       // B6 in byte 0
        // x.aux = this lane's copy of the table (up to sixteen interleaved copies, entry v of copy r at byte
        // (v << aux_sh) + 8 r: with sixteen, the lanes of a half-warp always hit sixteen different bank pairs; with one
        // copy the look-ups run 3-4 deep in bank conflicts and the shared-memory pipe bounds the kernel — but a CTA
        // with few rows spends longer filling sixteen copies than it loses to the conflicts, so the host picks)
        const uint2 e0 = lds_u2(x.aux + (__byte_perm(q0, 0, 0x4442) << x.aux_sh)), e1 = lds_u2(x.aux + (__byte_perm(q0, 0, 0x4443) << x.aux_sh));
        const uint2 e2 = lds_u2(x.aux + (__byte_perm(q1, 0, 0x4440) << x.aux_sh)), e3 = lds_u2(x.aux + (__byte_perm(q1, 0, 0x4441) << x.aux_sh));
        const uint2 e4 = lds_u2(x.aux + (__byte_perm(q1, 0, 0x4442) << x.aux_sh)), e5 = lds_u2(x.aux + (__byte_perm(q1, 0, 0x4443) << x.aux_sh));
        const uint2 e6 = lds_u2(x.aux + (__byte_perm(q2, 0, 0x4440) << x.aux_sh));
        uint32_t q[8];
        q[0] = e0.x;                                   // elements 0..3   = byte 0 digits 0..3
        q[1] = __byte_perm(e0.y, e1.x, 0x6543);        // 4..7            = byte 0 digit 4, byte 1 digits 0..2
        q[2] = __byte_perm(e1.y, e2.x, 0x5432);        // 8..11           = byte 1 digits 3, 4, byte 2 digits 0, 1
        q[3] = __byte_perm(e2.y, e3.x, 0x4321);        // 12..15          = byte 2 digits 2..4, byte 3 digit 0
        q[4] = e3.y;                                   // 16..19          = byte 3 digits 1..4
        q[5] = e4.x;                                   // 20..23          = byte 4 digits 0..3
        q[6] = __byte_perm(e4.y, e5.x, 0x6543);        // 24..27
        q[7] = __byte_perm(e5.y, e6.x, 0x5432);        // 28..31          = byte 5 digits 3, 4, byte 6 digits 0, 1
        int acc[NP];
#pragma unroll
        for (int pl = 0; pl < NP; pl++) acc[pl] = 0;
#pragma unroll
        for (int j = 0; j < 8; j++)
#pragma unroll
            for (int pl = 0; pl < NP; pl++)
                acc[pl] = (pl == NP - 1) ? dp4a_ss(q[j], x.q[pl][j], acc[pl]) : dp4a_su(q[j], x.q[pl][j], acc[pl]);
        // a digit byte v >= 243  <=>  bit 7 of v and bit 7 of (v & 0x7F) + 13
        const uint32_t bad = (q0 & ((q0 & 0x7F7F7F7Fu) + 0x0D0D0D0Du) & 0x80800000u) | (q1 & ((q1 & 0x7F7F7F7Fu) + 0x0D0D0D0Du) & 0x80808080u) |
                             (q2 & ((q2 & 0x7F7F7F7Fu) + 0x0D0D0D0Du) & 0x00000080u);
        const float d = h2f((uint16_t)(q0 & 0xFFFFu));
        const int I = (NP == 3) ? (acc[NP - 1] << 16) + ((acc[1] << 8) + acc[0]) : (acc[1] << 8) + acc[0];   // |I| <= 32 * 2^22
        const float r = d * ((float)I * x.s);
        return bad ? __uint_as_float(0x7FC00000u) : r;
    } else if (TYPE == T_Q8_K) {
        // the reference's own Q8_K (core/GGMLComputeOps.kt:382-432): 292-byte super-block = 73 words: d (f32), 256 int8
        const uint32_t *wp = row + 73 * (b >> 3) + 1 + 8 * (b & 7);
        int acc[NP];
#pragma unroll
        for (int pl = 0; pl < NP; pl++) acc[pl] = 0;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const uint32_t q = wp[j];
#pragma unroll
            for (int pl = 0; pl < NP; pl++)
                acc[pl] = (pl == NP - 1) ? dp4a_ss(q, x.q[pl][j], acc[pl]) : dp4a_su(q, x.q[pl][j], acc[pl]);
        }
        const float d = __uint_as_float(row[73 * (b >> 3)]);
        float f;
        if (NP == 3) f = fmaf((float)acc[NP - 1], 65536.0f, (float)((acc[1] << 8) + acc[0]));
        else f = (float)((acc[1] << 8) + acc[0]);
        return d * (f * x.s);
    } else {
        const uint32_t *wp = row + 8 * b + (b >> 1);
        uint32_t w[9];
#pragma unroll
        for (int i = 0; i < 9; i++) w[i] = wp[i];
        const uint32_t sel = odd ? 0x7654u : 0x5432u;
        int acc[NP];
#pragma unroll
        for (int pl = 0; pl < NP; pl++) acc[pl] = 0;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const uint32_t q = __byte_perm(w[j], w[j + 1], sel);
#pragma unroll
            for (int pl = 0; pl < NP; pl++)
                acc[pl] = (pl == NP - 1) ? dp4a_ss(q, x.q[pl][j], acc[pl]) : dp4a_su(q, x.q[pl][j], acc[pl]);
        }
        const float d = h2f((uint16_t)(odd ? (w[0] >> 16) : w[0]));
        float f;
        if (NP == 3)   // planes combined in F32: the plane-2 term alone can exceed int32 for 8-bit quants
            f = fmaf((float)acc[NP - 1], 65536.0f, (float)((acc[1] << 8) + acc[0]));
        else           // |acc1| <= 32*127*128 < 2^19: the 16-bit combination is exact in int32
            f = (float)((acc[1] << 8) + acc[0]);
        return d * (f * x.s);
    }
}

// Kernel parameters.  Every launch gets a fresh parameter block, so the first touch of each of its
// cache lines is a constant-cache miss (~500 clk measured); the fields every thread needs sit in
// the first two 64-byte lines and the kernel touches both lines in its first instructions.
struct GemvParams {
    const uint8_t *a;     // packed rows, 16-byte aligned
    const float *b;       // K activations, 16-byte aligned
    float *dst;           // M_total outputs
    int K, M;             // this launch's rows
    int row_bytes;        // K/32 * blockBytes
    int nblk;             // K/32 block columns
    int ncw;              // column warps = ceil(nblk/32)
    int rg;               // row groups
    int cwarps;           // consumer warps = ncw * rg (producer is warp `cwarps`)
    int stage_rows;
    int nstages;
    int unit_rows;        // CTA row ranges are multiples of this (keeps TMA sizes 16-byte multiples)
    // ---- 64 bytes
    int units_per_cta, units_rem;   // M / unit_rows = units_per_cta * grid + units_rem
    int red_rows;         // rows the partial-sum buffer holds between flushes (multiple of stage_rows)
    int red_stride;       // floats per row in the partial-sum buffer (ncw * 8 + pad)
    int red_off, x_off, stage_off;  // dynamic smem offsets (bytes)
    int wait_mode;        // GEMV_WAIT_*
    short npeers;
    short nib;            // Q4 nibble order: 0 Kotlin accessors (byte j = elements 2j, 2j+1), 1 ggml / GGUF (byte j = elements j, j+16)
    int rg_shift;         // rg == 1 << rg_shift (row groups are a power of two: no integer division per stage)
    int ng;               // chained kernel only: warp groups of rg * ncw warps each; group g takes stages g, g + ng, ...
    int fix_gap;          // range fix-up: a block with a non-zero |x| below 2^-fix_gap of the block maximum is flagged (>= 255: never)
    int lut_off;          // BITNET_1_58: byte offset of the decode table in dynamic shared memory
    int lut_shift;        // ... log2 of its interleaved copies (0..4): entry v of copy r at byte (v << (3 + lut_shift)) + 8 r
    int64_t row_offset;   // dst index of row 0 (sharded weights)
    unsigned long long *tl;         // timeline buffer (tools/gemv_timeline.cu builds only), else NULL
    // ---- 120 bytes
    float *peer[8];       // optional peer-mapped copies of dst (fused all-gather)
    // ---- fused neighbours of the matmul in the layer graph (all optional, see GemvFusion in lk_kernels.h)
    const float *pro_w;   // RMSNorm gain (may be NULL with pro_mode 1)
    float *b_out;         // where CTA 0 also stores the activations it actually multiplies with (the unfused op's dst)
    const float *residual;   // dst[i] = residual[i] + (a x b)[i]
    float *dst_raw;       // optional: the bare product, when the residual sum goes to dst
    float pro_eps;
    int pair_off;         // epi_mode 1: first partial-sum row of segment 1
    int pro_mode;         // 0: b as is; 1: b = RMSNorm(b, pro_w, pro_eps)
    // ---- stacked segments: up to 3 weights of identical shape multiplied with the same activations in one launch
    // (Q/K/V, gate/up); every CTA runs its row range through each segment in turn (segment 0 is a / dst above)
    int nseg;
    int epi_mode;         // 1 (nseg == 2): epi_out[i] = silu(seg0[i]) * seg1[i], epi_act[i] = silu(seg0[i])
    const uint8_t *a_seg[2];
    float *dst_seg[2];
    float *epi_act, *epi_out;
    // ---- cross-GPU exchange folded into the launch (sharded matmuls, FUSED instantiation only; see GemvSync in lk_kernels.h)
    int out_off;                          // npeers > 0: byte offset of the output staging row (floats) the coalesced peer stores read
    int sig_n;                            // > 0: when the whole grid has stored, arrive on sig_peer[0..sig_n) (one per rank, own included)
    int wait_n;                           // > 0: before reading the activations wait for wait_n arrivals per epoch on wait_ctr
    unsigned long long *sig_peer[8];      // every rank's arrival counter of the signalled exchange point (peer-mapped)
    unsigned long long *sig_done;         // local: CTAs of this launch that have stored (the last one signals and re-arms it)
    unsigned long long *wait_ctr;         // local arrival counter of the awaited exchange point (monotonic)
    unsigned long long *wait_ticket;      // local: consumer CTAs that have taken a ticket (ticket / grid = epoch)
    unsigned *sync_timeout;               // local: bounded waits that gave up (never in a healthy run)
};

#ifdef LK_GEMV_TIMELINE
// slot 0 / 29: %globaltimer at entry / exit (aligns CTAs and launches); other slots: clock64 (cheap), slot 31 = clock64 at entry
#define LK_TL_SLOTS 32
__device__ __forceinline__ unsigned long long lk_gtime() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ unsigned lk_smid() { unsigned v; asm volatile("mov.u32 %0, %%smid;" : "=r"(v)); return v; }
#define LK_TL(slot) do { if (p.tl) { unsigned long long *q_ = p.tl + (size_t)blockIdx.x * LK_TL_SLOTS; \
    if ((slot) == 0) { q_[0] = lk_gtime(); q_[31] = (unsigned long long)clock64(); q_[30] = lk_smid(); } \
    else { q_[(slot)] = (unsigned long long)clock64(); if ((slot) == 7) q_[29] = lk_gtime(); } } } while (0)
// chained kernel: per CTA 8 header words (0: %globaltimer at entry, 1: clock64 at entry, 2: %smid, 3: %globaltimer at exit)
// + 8 clock64 stamps per phase (tools/chain_timeline.cu)
#define LK_CTL_STRIDE (8 + 8 * 8)
#define LK_CTL(ph, slot) do { if (cp.ph[0].tl) cp.ph[0].tl[(size_t)blockIdx.x * LK_CTL_STRIDE + 8 + (ph) * 8 + (slot)] = (unsigned long long)clock64(); } while (0)
#else
#define LK_TL(slot) do { } while (0)
#define LK_CTL(ph, slot) do { } while (0)
#endif

__device__ __forceinline__ float fmax3_nan_abs(float a, float b, float c) {   // max(|a|, |b|, |c|), NaN if any is NaN
    float d;
    asm("max.NaN.abs.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}

// ---- range fix-up -------------------------------------------------------------------------------------------------
// The planes hold T_k = rint(x_k / s) with ONE scale s = 2^(E-22) per 32-element block, so an element 2^g times smaller
// than its block's maximum keeps 23 - g bits and one more than 2^23 times smaller becomes 0, where the reference
// (F32 per element, core/GGMLComputeOps.kt:43-145) keeps 24 bits of every x_k.  That only shows when the large
// elements meet zero weights (Q4_0 nibble 8), but then it is a 100 % error.  prep_x therefore flags every block that
// holds a non-zero element below 2^-fix_gap of the block maximum (bit mask in the side words below), and before the
// partial sums of a row are reduced the CTA adds, for its rows and the flagged blocks only,
//     sum_k w(i, k) * r_k,   r_k = x_k - T_k * s   (exact in F32: both are multiples of ulp(x_k), |r_k| <= s / 2)
// with the reference's per-element weight formula, into the padding columns of the partial-sum buffer: x_k = T_k s + r_k
// exactly, so a flagged block is computed from the full F32 activations.  Cold path: natural activations flag a
// block in < 2 % of launches at the default fix_gap = 20 (P[|x| < 2^-20 max] ~ 2e-6 per element for N(0, 1)); the hot loop
// is untouched.  Elements between 2^-fix_gap and 1 of the maximum carry the documented absolute error <= s / 2.
constexpr int GEMV_SIDE_OFF = 128;        // bytes 128..255 of the barrier area: rinv, any-flag word, 28 mask words (896 block columns)
struct GemvSide {
    float rinv;
    uint32_t any;
    uint32_t mask[28];
};
static_assert(sizeof(GemvSide) <= 128, "side words share the first 256 bytes with the 16 mbarriers");
#ifdef LK_GEMV_NO_RANGE_FIX     // A/B builds only (tools/ab_build.sh): what the detection and the cold call sites cost
template <int TYPE> __host__ __device__ constexpr bool gemv_has_fixup() { return false; }
#else
template <int TYPE> __host__ __device__ constexpr bool gemv_has_fixup() { return TYPE == T_Q4_0 || TYPE == T_Q4_1 || TYPE == T_Q8_0 || TYPE == T_BITNET_1_58; }
#endif
// Where the detection sits and what the call sites cost was measured on one B200 box per run
// (profiles/r02_gemv_range_fix_ab.txt, headline step, GB/s): without detection and sites 4705; detection inside prep_x
// and light call sites (this code) 4626 (0.7 % each); the round-1 kernel 4613.  Rejected: detection by a dedicated 10th
// warp that scans x beside the consumers (-10.6 %; the extra warp alone, exiting at once, already costs 1.5 %), by the
// idle lanes of the producer warp (slower still: they share a warp with the lane that streams the weights), and a
// sixteen-argument call at the sites (spills; -7 % on the chained kernel although never taken; see gemv_fixup below).

__device__ __forceinline__ float fmin3_abs(float a, float b, float c) {   // min(|a|, |b|, |c|)
    float d;
    asm("min.abs.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}

// ---- activations -> byte planes, cooperatively by the consumer threads -------------------
// task = (block column, half): 16 consecutive floats; the two lanes of a pair share a block.
// Conversion without F2I (quarter-rate pipe): fmaf(v, c, 1.5*2^23) leaves rint(v*c) in the low
// mantissa bits; a non-finite element poisons the block scale with NaN instead.
// CG: read the activations / gain with ld.global.cg (L2 only).  The chained kernel below reads vectors that other CTAs
// of the SAME launch wrote a phase earlier; a non-coherent (ld.global.nc) or L1-cached load could return a stale line.
// CG = 0: ld.global.nc (read-only data path); 1: ld.global.cg (L2 only: the chained kernel, whose CTAs read vectors that
// other CTAs of the SAME launch wrote a phase earlier — an L1 line from the earlier phase would be stale); 2: plain
// ld.global (L1-allocating): the fused / sharded launches, whose activations may have been written over NVLink while
// the kernel was already running — but never read by it before its acquire, so no stale L1 line can exist, and the
// second and third pass of the RMSNorm prologue hit L1 instead of paying the L2 latency again.
template <int CG>
__device__ __forceinline__ float4 ld_vec4(const float4 *q) { return CG == 1 ? __ldcg(q) : CG == 2 ? *q : __ldg(q); }

// pre / use_pre: the first round's 16 activations of this thread, loaded by the caller at kernel entry (independent
// launches: x is final before the kernel starts, so its DRAM latency — every launch has its own cold x — hides under the
// barrier set-up instead of heading the critical path; ncu: 26 % of the stall samples sat on the first use of that load).
template <int TYPE, int NP, bool FUSED, int CG = 0, bool DETECT = true>
__device__ __forceinline__ void prep_x(const GemvParams &p, uint32_t *xs, int tid, int nthreads, GemvSide *side,
                                       const float4 *pre = nullptr, bool use_pre = false) {
    constexpr int TB = NP * 8 - 2;     // |T| <= 2^TB: 22 (three byte planes) or 14 (two)
    const int ntasks = p.nblk * 2;
    // Fused RMSNorm prologue (K3 folded in): every CTA needs all of x anyway, so it also takes the sum of squares
    // (F32 partials per thread, warp shuffles, one pass over <= 28 warp totals) and multiplies with 1/rms and the
    // gain while converting.  1/rms instead of the op's IEEE division: <= 2 ulp from the unfused RMSNorm.
    float rinv = 1.0f;
    // chained kernel and fused launches, K <= 16 * nthreads: every thread owns at most one task, so the values loaded for the sum of squares
    // (and the gain, fetched under the same latency) stay in registers for the conversion: one L2 round trip on the
    // hand-off's critical path instead of three.  Same per-thread order of operations as the two-pass code.
    const bool keep = CG != 0 && FUSED && p.pro_mode == 1 && ntasks <= nthreads;
    float v[16], gw[16];
    if (FUSED && p.pro_mode == 1) {
        float *part = (float *)(xs + (size_t)p.nblk * gemv_xslot_words(NP)) - 32;   // tail of the last slots: rewritten below
        float acc = 0.f;
        if (keep) {
            if (tid < ntasks) {
                const float4 *src = (const float4 *)(p.b + (size_t)tid * 16);
                const float4 *wsrc = (const float4 *)(p.pro_w + (size_t)tid * 16);
                float4 f[4], g[4];
#pragma unroll
                for (int i = 0; i < 4; i++) f[i] = ld_vec4<CG>(src + i);
                if (p.pro_w) {
#pragma unroll
                    for (int i = 0; i < 4; i++) g[i] = ld_vec4<CG>(wsrc + i);
                } else {
#pragma unroll
                    for (int i = 0; i < 4; i++) g[i] = make_float4(1.f, 1.f, 1.f, 1.f);
                }
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    v[4 * i] = f[i].x; v[4 * i + 1] = f[i].y; v[4 * i + 2] = f[i].z; v[4 * i + 3] = f[i].w;
                    gw[4 * i] = g[i].x; gw[4 * i + 1] = g[i].y; gw[4 * i + 2] = g[i].z; gw[4 * i + 3] = g[i].w;
                }
#pragma unroll
                for (int i = 0; i < 16; i++) acc = fmaf(v[i], v[i], acc);
            }
        } else {
#pragma unroll 1
        for (int task = tid; task < ntasks; task += nthreads) {
            const float4 *src = (const float4 *)(p.b + (size_t)task * 16);
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const float4 f = ld_vec4<CG>(src + i);
                acc = fmaf(f.x, f.x, acc); acc = fmaf(f.y, f.y, acc); acc = fmaf(f.z, f.z, acc); acc = fmaf(f.w, f.w, acc);
            }
        }
        }
        acc = warp_sum(acc);
        if ((tid & 31) == 0) part[tid >> 5] = acc;
        asm volatile("bar.sync 1, %0;" ::"r"(nthreads) : "memory");
        float tot = 0.f;
        for (int w = 0; w < (nthreads >> 5); w++) tot += part[w];
        rinv = 1.0f / sqrtf(tot / (float)p.K + p.pro_eps);
        if (tid == 0) side->rinv = rinv;
        asm volatile("bar.sync 1, %0;" ::"r"(nthreads) : "memory");     // `part` is about to be overwritten by the planes
    }
#pragma unroll 1
    for (int base = 0; base < ntasks; base += nthreads) {
        const int task = base + tid;
        const bool on = task < ntasks;
        if (on) {
            if (!keep) {
            const float4 *src = (const float4 *)(p.b + (size_t)task * 16);
            const bool from_pre = pre != nullptr && use_pre && base == 0;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const float4 f = from_pre ? pre[i] : ld_vec4<CG>(src + i);
                v[4 * i] = f.x; v[4 * i + 1] = f.y; v[4 * i + 2] = f.z; v[4 * i + 3] = f.w;
            }
            }
            if (FUSED && p.pro_mode == 1) {
#pragma unroll
                for (int i = 0; i < 16; i++) v[i] = __fmul_rn(v[i], rinv);
                if (keep) {
                    if (p.pro_w) {
#pragma unroll
                        for (int i = 0; i < 16; i++) v[i] = __fmul_rn(v[i], gw[i]);
                    }
                } else if (p.pro_w) {
                    const float4 *wsrc = (const float4 *)(p.pro_w + (size_t)task * 16);
#pragma unroll
                    for (int i = 0; i < 4; i++) {
                        const float4 f = ld_vec4<CG>(wsrc + i);
                        v[4 * i] = __fmul_rn(v[4 * i], f.x); v[4 * i + 1] = __fmul_rn(v[4 * i + 1], f.y);
                        v[4 * i + 2] = __fmul_rn(v[4 * i + 2], f.z); v[4 * i + 3] = __fmul_rn(v[4 * i + 3], f.w);
                    }
                }
                if (p.b_out && blockIdx.x == 0) {
                    float4 *o = (float4 *)(p.b_out + (size_t)task * 16);
#pragma unroll
                    for (int i = 0; i < 4; i++) o[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
                }
            }
        } else {
#pragma unroll
            for (int i = 0; i < 16; i++) v[i] = 0.f;
        }
#ifdef LK_GEMV_TIMELINE
        if (tid == 0 && base == 0) { if (v[0] + v[4] + v[8] + v[12] == 12345.f) xs[0] = 1; LK_TL(22); }
#endif
        // max |v| with NaN propagation (3-input max.NaN.abs): a NaN or inf element leaves exponent 255 in `ab`
        float amax = 0.f;
#pragma unroll
        for (int i = 0; i < 16; i += 2) amax = fmax3_nan_abs(amax, v[i], v[i + 1]);
        uint32_t ab = __float_as_uint(amax) & 0x7FFFFFFFu;
        ab = max(ab, __shfl_xor_sync(0xffffffffu, ab, 1));   // unsigned order: finite < inf < NaN
        int e = (int)(ab >> 23);
        const bool nonfinite = e == 255;
        if (DETECT && gemv_has_fixup<TYPE>()) {
            // Range check, cheap in the common case: min |x| over the block (3-input min.abs, 8 instructions).  Only a block
            // whose smallest |x| is zero or below 2^-fix_gap of its maximum takes the exact look (smallest NON-ZERO |x| as
            // 2 * bits - 1: the shift drops the sign, +-0 wraps to 0xFFFFFFFF); both lanes of a pair branch alike.
            float amin = __uint_as_float(0x7F800000u);
#pragma unroll
            for (int i = 0; i < 16; i += 2) amin = fmin3_abs(amin, v[i], v[i + 1]);
            uint32_t mb = __float_as_uint(amin);
            mb = min(mb, __shfl_xor_sync(0xffffffffu, mb, 1));
            if ((int)(mb >> 23) + p.fix_gap < e && !nonfinite) {
                uint32_t mn = 0xFFFFFFFFu;
#pragma unroll
                for (int i = 0; i < 16; i++) mn = min(mn, __float_as_uint(v[i]) * 2u - 1u);
                // only the two lanes of this pair are known to be here (the branch is data dependent): pair mask, not the full warp
                mn = min(mn, __shfl_xor_sync(3u << ((tid & 31) & ~1), mn, 1));
                const uint32_t t = mn + 1u;
                if (on && (task & 1) == 0 && t != 0u && (int)(t >> 24) + p.fix_gap < e) {
                    atomicOr(&side->mask[task >> 6], 1u << ((task >> 1) & 31));
                    side->any = 1u;
                }
            }
        }
        e = e < 24 ? 24 : e;
        // c = 2^(TB - (e - 126)) so that |x| < 2^(e-126) maps into |T| <= 2^TB;  s = 1/c.  A non-finite block gets
        // s = NaN, which poisons every product of the block whatever its T values are.
        const float c = nonfinite ? 0.f : __uint_as_float((uint32_t)(253 + TB - e) << 23);
        const float s = nonfinite ? __uint_as_float(0x7FC00000u) : __uint_as_float((uint32_t)(e + 1 - TB) << 23);
        int T[16];
        int sumT = 0;
#pragma unroll
        for (int i = 0; i < 16; i++) {
            T[i] = (int)__float_as_uint(fmaf(v[i], c, 12582912.0f)) - 0x4B400000;
            sumT += T[i];
        }
        const int sumT_half = sumT;                       // this lane's 16 elements (Q2_K: one sub-block)
        sumT += __shfl_xor_sync(0xffffffffu, sumT, 1);   // both halves of the block
#ifdef LK_GEMV_TIMELINE
        if (tid == 0 && base == 0) { if (sumT == 0x7FFFFFFF) xs[0] = 1; LK_TL(23); }
#endif
        if (on) {
            const int blk = task >> 1, h = task & 1;
            uint32_t *slot = xs + (size_t)blk * gemv_xslot_words(NP);
#pragma unroll
            for (int pl = 0; pl < NP; pl++) {
                const uint32_t s2 = 0x0040u + 0x0011u * pl;   // byte pl of a, byte pl of b
                uint32_t wds[4];
                if (TYPE == T_Q2_K) {
#pragma unroll
                    for (int j = 0; j < 4; j++)   // 2-bit field j of a quant word's four bytes = elements j, 4+j, 8+j, 12+j of my half
                        wds[j] = __byte_perm(__byte_perm(T[j], T[4 + j], s2), __byte_perm(T[8 + j], T[12 + j], s2), 0x5410);
                } else if (TYPE == T_Q8_0 || TYPE == T_Q8_K || TYPE == T_BITNET_1_58 || p.nib) {
#pragma unroll
                    for (int j = 0; j < 4; j++)
                        wds[j] = __byte_perm(__byte_perm(T[4 * j], T[4 * j + 1], s2), __byte_perm(T[4 * j + 2], T[4 * j + 3], s2), 0x5410);
                    if (TYPE != T_Q8_0 && TYPE != T_Q8_K && TYPE != T_BITNET_1_58) {
                        // ggml order: quant word j holds elements 4j..4j+3 in its low nibbles (plane word 2j) and elements
                        // 16+4j.. in its high nibbles (plane word 2j+1): my half h fills words 2j + h
#pragma unroll
                        for (int j = 0; j < 4; j++) slot[pl * 8 + 2 * j + h] = wds[j];
                        continue;
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < 2; j++) {   // nibble word j of my half: lo = even elements, hi = odd
                        wds[2 * j] = __byte_perm(__byte_perm(T[8 * j], T[8 * j + 2], s2), __byte_perm(T[8 * j + 4], T[8 * j + 6], s2), 0x5410);
                        wds[2 * j + 1] = __byte_perm(__byte_perm(T[8 * j + 1], T[8 * j + 3], s2), __byte_perm(T[8 * j + 5], T[8 * j + 7], s2), 0x5410);
                    }
                }
                *(uint4 *)(slot + pl * 8 + h * 4) = make_uint4(wds[0], wds[1], wds[2], wds[3]);
            }
            if (h == 0) { slot[NP * 8] = __float_as_uint((float)sumT); slot[NP * 8 + 1] = __float_as_uint(s); }
            if (TYPE == T_Q2_K) slot[NP * 8 + 2 + h] = __float_as_uint((float)sumT_half);
        }
    }
}

// this thread's block column of the activation planes: shared memory -> registers
template <int TYPE, int NP>
__device__ __forceinline__ void load_xregs(XRegs<NP> &x, const uint32_t *xs, int colc) {
    const uint4 *slot = (const uint4 *)(xs + (size_t)colc * gemv_xslot_words(NP));
#pragma unroll
    for (int pl = 0; pl < NP; pl++)
#pragma unroll
        for (int h = 0; h < 2; h++) {
            const uint4 v = slot[pl * 2 + h];
            x.q[pl][4 * h] = v.x; x.q[pl][4 * h + 1] = v.y; x.q[pl][4 * h + 2] = v.z; x.q[pl][4 * h + 3] = v.w;
        }
    const uint4 t = slot[NP * 2];
    x.s = __uint_as_float(t.y);
    x.c = ((TYPE == T_Q4_0) ? -8.0f : 1.0f) * x.s * __uint_as_float(TYPE == T_Q2_K ? t.z : t.x);
    x.c2 = x.s * __uint_as_float(t.w);
}

#ifndef LK_CHAIN_DETECT        // A/B builds only (tools/ab_build.sh)
#define LK_CHAIN_DETECT true
#endif
#ifndef LK_GEMV_DETECT
#define LK_GEMV_DETECT true
#endif
#ifdef LK_CHAIN_NO_SITES
#define LK_CHAIN_FIXUP(...)
#elif defined(LK_CHAIN_SITE_STUB)     // the flag test and a call of a trivial function
__device__ __noinline__ void gemv_fixup_stub(float *red, int tid) { red[tid] += 1.0f; }
#define LK_CHAIN_FIXUP(CGF, PRO_W, PRO_MODE, NTHR, AW, ROW_FIRST, NROWS, REDP) \
    if (*(volatile uint32_t *)(smem + GEMV_SIDE_OFF + 4) != 0u) gemv_fixup_stub(REDP, tid);
#elif defined(LK_CHAIN_SITE_TRAP)     // the flag test alone
#define LK_CHAIN_FIXUP(CGF, PRO_W, PRO_MODE, NTHR, AW, ROW_FIRST, NROWS, REDP) \
    if (*(volatile uint32_t *)(smem + GEMV_SIDE_OFF + 4) != 0u) __trap();
#else
#define LK_CHAIN_FIXUP(CGF, PRO_W, PRO_MODE, NTHR, AW, ROW_FIRST, NROWS, REDP)                                            \
    if (gemv_has_fixup<TYPE>() && __builtin_expect(*(volatile uint32_t *)(smem + GEMV_SIDE_OFF + 4) != 0u, 0)) {                                \
        gemv_fixup<TYPE, NP, 1, true, NTHR>(&p, AW, ROW_FIRST, NROWS, REDP, smem);                                        \
        asm volatile("bar.sync 1, %0;" ::"r"(NTHR) : "memory");                                                          \
        load_xregs<TYPE, NP>(x, xs, colc);      /* nothing of x has to survive the call */                                \
    }
#endif
enum { GEMV_WAIT_FULL = 0, GEMV_WAIT_X = 1, GEMV_WAIT_END = 2 };



// Cold path of the range fix-up (see GemvSide): for rows [row_first, row_first + n) of the packed matrix `a` and every
// flagged block, sum_k w(i, k) * r_k into padding column j = (rank of the block among the flagged ones) & 7 of the
// row's partial sums.  Scalars instead of the parameter struct: a reference to kernel parameters would force a local copy.
#ifndef LK_FIXUP_ATTR           // A/B builds only (tools/ab_build.sh)
#define LK_FIXUP_ATTR __noinline__
#endif
// First version of the work split, kept for the stand-alone launches: one thread per (row, column) walks the flagged blocks'
// 32 elements.  Slow when it runs (~9 us with one or two blocks flagged) but it never does on natural activations, and the
// stand-alone kernels are compiled around it: with the cooperative version below as their callee ptxas schedules the hot
// kernels differently (72 / 84 instead of 71 / 86 registers) and the headline step loses 2 % (4580-4590 -> 4487-4503 GB/s,
// same box, three runs each) although the path is never taken.  The chained kernel K1c takes the cooperative version.
template <int TYPE, int NP, bool CG>
__device__ __forceinline__ void gemv_fixup_body_rows(const uint8_t *a, const float *b, const float *pro_w, int pro_mode, int nib, int row_bytes,
                                                int nblk, int red_stride, int pad_col, int row_first, int n, float *red, const uint32_t *xs,
                                                const GemvSide *side, int tid, int nct) {
    const int nw = (nblk + 31) >> 5;
    const float rinv = side->rinv;
#pragma unroll 1
    for (int item = tid; item < n * GEMV_RED_PAD; item += nct) {
        const int r = item >> 3, j = item & 7;
        const uint8_t *rowp = a + (size_t)(row_first + r) * row_bytes;
        float sum = 0.f;
        int idx = 0;
#pragma unroll 1
        for (int w = 0; w < nw; w++) {
            uint32_t m = side->mask[w];
#pragma unroll 1
            while (m) {
                const int bit = __ffs((int)m) - 1;
                m &= m - 1;
                if ((idx++ & 7) != j) continue;
                const int blk = w * 32 + bit;
                const float s = __uint_as_float(xs[(size_t)blk * gemv_xslot_words(NP) + NP * 8 + 1]);
                const float c = __frcp_rn(s);                                 // a power of two: exact
                const uint8_t *bp = rowp + (size_t)blk * (TYPE == T_BITNET_1_58 ? 10 : block_bytes(TYPE));
                const float d = h2f((uint16_t)(bp[0] | (bp[1] << 8)));
                float mq = 0.f;
                if (TYPE == T_Q4_1) mq = h2f((uint16_t)(bp[2] | (bp[3] << 8)));
                constexpr int QOFF = TYPE == T_Q4_1 ? 4 : 2;
                float acc = 0.f;
#pragma unroll 1
                for (int k = 0; k < 32; k++) {
                    const int kk = blk * 32 + k;
                    float v = CG ? __ldcg(b + kk) : b[kk];
                    if (pro_mode == 1) {                                      // the fused RMSNorm prologue, same roundings as prep_x
                        v = __fmul_rn(v, rinv);
                        if (pro_w) v = __fmul_rn(v, CG ? __ldcg(pro_w + kk) : pro_w[kk]);
                    }
                    const int T = (int)__float_as_uint(fmaf(v, c, 12582912.0f)) - 0x4B400000;
                    const float rk = fmaf(-(float)T, s, v);                   // exact residual
                    float wv;                                                 // the reference's weight, core/GGMLComputeOps.kt:43-145
                    if (TYPE == T_Q8_0) {
                        wv = __fmul_rn(d, (float)(int8_t)bp[2 + k]);
                    } else if (TYPE == T_BITNET_1_58) {                       // digit k % 5 of byte k / 5, core/GGMLTypes.kt:975-990
                        const int g = (k * 13) >> 6;                          // k / 5 for k < 64
                        uint32_t dv = bp[2 + g];
                        for (int z = k - 5 * g; z > 0; z--) dv /= 3u;
                        wv = __fmul_rn(d, (float)((int)(dv % 3u) - 1));
                    } else {
                        const uint8_t byte = bp[QOFF + (nib ? (k & 15) : (k >> 1))];
                        const float q = (float)((nib ? (k >> 4) : (k & 1)) ? (byte >> 4) : (byte & 15));
                        wv = TYPE == T_Q4_0 ? __fmul_rn(d, __fsub_rn(q, 8.0f)) : __fadd_rn(__fmul_rn(d, q), mq);
                    }
                    acc = fmaf(wv, rk, acc);
                }
                sum += acc;
            }
        }
        red[(size_t)r * red_stride + pad_col + j] = sum;
    }
}
// Work split (round 2, second version): warp w owns rows w, w + nwarps, ... for EVERY flagged block and lane k owns element k
// of the block — one residual per lane and block, one quant byte and one scale per lane and row, the block's sum by a shuffle
// tree — so a single flagged block is spread over the whole CTA.  The first version gave each (row, column) to one thread
// that walked the 32 elements with a dependent L2 load each: ~9 us per launch with one or two blocks flagged
// (profiles/r02m_layer_types.txt).
template <int TYPE, int NP, bool CG>
__device__ __forceinline__ void gemv_fixup_body(const uint8_t *a, const float *b, const float *pro_w, int pro_mode, int nib, int row_bytes,
                                                int nblk, int red_stride, int pad_col, int row_first, int n, float *red, const uint32_t *xs,
                                                const GemvSide *side, int tid, int nct) {
    const int nw = (nblk + 31) >> 5;
    const float rinv = side->rinv;
    const int warp = tid >> 5, lane = tid & 31, nwarps = nct >> 5;
    constexpr int BBT = TYPE == T_BITNET_1_58 ? 10 : block_bytes(TYPE);
    constexpr int QOFF = TYPE == T_Q4_1 ? 4 : 2;
    // this lane's byte of a block and, for the 4-bit types, its nibble; BitNet: digit lane % 5 of byte lane / 5
    const int g5 = (lane * 13) >> 6;                              // lane / 5 for lane < 64
    const int qb = TYPE == T_Q8_0 ? lane : TYPE == T_BITNET_1_58 ? g5 : (nib ? (lane & 15) : (lane >> 1));
    const bool hi = (TYPE == T_Q4_0 || TYPE == T_Q4_1) && (nib ? (lane >> 4) : (lane & 1));
    for (int r = warp + nwarps * (lane >> 3); r < n; r += nwarps * 4) red[(size_t)r * red_stride + pad_col + (lane & 7)] = 0.f;
    __syncwarp();
    int idx = 0;
#pragma unroll 1
    for (int w = 0; w < nw; w++) {
        uint32_t m = side->mask[w];
#pragma unroll 1
        while (m) {
            const int bit = __ffs((int)m) - 1;
            m &= m - 1;
            const int j = idx++ & 7;
            const int blk = w * 32 + bit;
            const float s = __uint_as_float(xs[(size_t)blk * gemv_xslot_words(NP) + NP * 8 + 1]);
            const float c = __frcp_rn(s);                                 // a power of two: exact
            const int kk = blk * 32 + lane;
            float v = CG ? __ldcg(b + kk) : b[kk];
            if (pro_mode == 1) {                                          // the fused RMSNorm prologue, same roundings as prep_x
                v = __fmul_rn(v, rinv);
                if (pro_w) v = __fmul_rn(v, CG ? __ldcg(pro_w + kk) : pro_w[kk]);
            }
            const int T = (int)__float_as_uint(fmaf(v, c, 12582912.0f)) - 0x4B400000;
            const float rk = fmaf(-(float)T, s, v);                       // exact residual of this lane's element
#pragma unroll 2
            for (int r = warp; r < n; r += nwarps) {
                const uint8_t *bp = a + (size_t)(row_first + r) * row_bytes + (size_t)blk * BBT;
                const uint32_t hdr = __ldg((const unsigned short *)bp) | (TYPE == T_Q4_1 ? (uint32_t)__ldg((const unsigned short *)bp + 1) << 16 : 0u);
                const uint32_t byte = __ldg(bp + QOFF + qb);
                const float d = h2f((uint16_t)(hdr & 0xFFFFu));
                float wv;                                                 // the reference's weight, core/GGMLComputeOps.kt:43-145
                if (TYPE == T_Q8_0) {
                    wv = __fmul_rn(d, (float)(int8_t)byte);
                } else if (TYPE == T_BITNET_1_58) {                       // digit lane % 5 of byte lane / 5, core/GGMLTypes.kt:975-990
                    uint32_t dv = byte;
                    for (int z = lane - 5 * g5; z > 0; z--) dv /= 3u;
                    wv = __fmul_rn(d, (float)((int)(dv % 3u) - 1));
                } else {
                    const float q = (float)(hi ? (byte >> 4) : (byte & 15u));
                    wv = TYPE == T_Q4_0 ? __fmul_rn(d, __fsub_rn(q, 8.0f)) : __fadd_rn(__fmul_rn(d, q), h2f((uint16_t)(hdr >> 16)));
                }
                const float sum = warp_sum(__fmul_rn(wv, rk));
                if (lane == 0) red[(size_t)r * red_stride + pad_col + j] += sum;
            }
            __syncwarp();
        }
    }
}
// The call: six arguments, everything else read through the pointer to the launch's (or the phase's) parameters — the
// kernel parameter is __grid_constant__, so its address is a plain generic pointer into the constant bank, no local copy.
// A sixteen-argument call at the four sites made the kernels spill (ptxas: up to 200 bytes under the 56-register cap) and
// cost the chained kernel 7 % although the call is never taken on natural data (profiles/r02_gemv_range_fix_ab.txt);
// the callers also reload their activation registers after the call instead of keeping them alive across it.
// NCT > 0: compile-time thread count (chained kernel); else cwarps * 32.  FUSED: the launch may carry the RMSNorm prologue.
template <int TYPE, int NP, int CG, bool FUSED, int NCT>
__device__ LK_FIXUP_ATTR void gemv_fixup(const GemvParams *pp, const uint8_t *a, int row_first, int n, float *red, const uint8_t *smem) {
    if (NCT > 0)       // the chained kernel
        gemv_fixup_body<TYPE, NP, CG != 0>(a, pp->b, FUSED ? pp->pro_w : nullptr, FUSED ? pp->pro_mode : 0, pp->nib, pp->row_bytes, pp->nblk,
                                           pp->red_stride, pp->ncw * 8, row_first, n, red, (const uint32_t *)(smem + pp->x_off),
                                           (const GemvSide *)(smem + GEMV_SIDE_OFF), (int)threadIdx.x, NCT);
    else
        gemv_fixup_body_rows<TYPE, NP, CG != 0>(a, pp->b, FUSED ? pp->pro_w : nullptr, FUSED ? pp->pro_mode : 0, pp->nib, pp->row_bytes, pp->nblk,
                                                pp->red_stride, pp->ncw * 8, row_first, n, red, (const uint32_t *)(smem + pp->x_off),
                                                (const GemvSide *)(smem + GEMV_SIDE_OFF), (int)threadIdx.x, pp->cwarps * 32);
}

// cross-lane part of the final reduction: 8 lanes per row, 4 rows per warp per pass
__device__ __forceinline__ float silu_f(float x) { return __fmul_rn(x, __fdiv_rn(1.0f, __fadd_rn(1.0f, expf(-x)))); }   // = lk_elementwise.cu silu1


// ---- cross-GPU exchange, device side -----------------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_sys_add(unsigned long long *p, unsigned long long v) {
    asm volatile("red.release.sys.global.add.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// this CTA's n outputs (staged in shared memory by finish_rows / finish_pair) into every peer's copy of dst: one warp per
// peer, lanes on consecutive floats, so the NVLink writes leave as whole 128-byte lines instead of one 4-byte packet per row
__device__ __forceinline__ void peer_store_rows(const GemvParams &p, const float *outbuf, int n, int64_t out0, int warp, int lane) {
#pragma unroll 1
    for (int q = warp; q < p.npeers; q += p.cwarps) {
        float *dstq = p.peer[q] + out0;
#pragma unroll 1
        for (int i = lane; i < n; i += 32) dstq[i] = outbuf[i];
    }
}
// After the last peer store of the CTA (call with all consumer threads): EVERY CTA arrives on every rank's counter, with a
// weight — the CTAs of one launch add up to EXCHANGE_UNIT whatever the grid size — so the waiting side needs to know the
// number of ranks only, and nothing funnels through a last CTA (that cost a local atomic round trip and a second
// system-scope fence on the critical path).  fence.sys after the bar.sync: this CTA's peer stores (made by its other
// threads) are performed at system scope before the arrival is issued (release pattern: fence; relaxed red).
constexpr unsigned long long EXCHANGE_UNIT = 1ull << 20;
__device__ __forceinline__ void red_relaxed_sys_add(unsigned long long *p, unsigned long long v) {
    asm volatile("red.relaxed.sys.global.add.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void exchange_signal(const GemvParams &p, int tid, int nct) {
    asm volatile("bar.sync 1, %0;" ::"r"(nct) : "memory");
    if (tid == 0) {
        const unsigned long long g = gridDim.x, c = blockIdx.x;
        const unsigned long long w = (c + 1ull) * EXCHANGE_UNIT / g - c * EXCHANGE_UNIT / g;
        __threadfence_system();
#pragma unroll 1
        for (int q = 0; q < p.sig_n; q++) red_relaxed_sys_add(p.sig_peer[q], w);
    }
}
// Before the activations are read: wait until every rank's producer launch of this epoch has arrived.  The counter only
// grows (wait_n * EXCHANGE_UNIT per epoch); the epoch is this CTA's ticket / grid size.  Bounded: a rank that died cannot hang the device;
// a time-out is counted and reported by the next synchronising call on the context.
__device__ __forceinline__ void exchange_wait(const GemvParams &p, unsigned long long ticket) {
    const unsigned long long target = (ticket / gridDim.x + 1ull) * (unsigned long long)p.wait_n * EXCHANGE_UNIT;
    const long long t0 = clock64();
    while (ld_acquire_sys(p.wait_ctr) < target) {
        if (clock64() - t0 > 4000000000ll) { atomicAdd(p.sync_timeout, 1u); break; }
    }
}

template <bool FUSED, bool CG = false, bool PEER = FUSED>
__device__ __forceinline__ void finish_rows(const GemvParams &p, float *dst, const float *red, int nred, int64_t out0, int warp, int lane) {
    const int sub = lane & 7, rsel = lane >> 3;
    const int nc4 = p.red_stride >> 2;   // float4 chunks per row (columns beyond npairs hold zeros)
#pragma unroll 1
    for (int rb = warp * 4; rb < nred; rb += p.cwarps * 4) {   // warp-uniform trip count (shuffles inside)
        const int r = rb + rsel;
        float v = 0.f;
        if (r < nred) {
            const float4 *row = (const float4 *)(red + (size_t)r * p.red_stride);
#pragma unroll 2
            for (int c = sub; c < nc4; c += 8) {
                const float4 f = row[c];
                v += (f.x + f.y) + (f.z + f.w);
            }
        }
        v += __shfl_xor_sync(0xffffffffu, v, 1);
        v += __shfl_xor_sync(0xffffffffu, v, 2);
        v += __shfl_xor_sync(0xffffffffu, v, 4);
        if (sub == 0 && r < nred) {
            const int64_t o = out0 + r;
            if (FUSED && p.residual) {             // fused residual ADD (computeAdd's single F32 add, same rounding)
                if (p.dst_raw) p.dst_raw[o] = v;
                v = __fadd_rn(CG ? __ldcg(p.residual + o) : p.residual[o], v);
            }
            dst[o] = v;
            if (PEER && p.npeers) ((float *)((uint8_t *)red - p.red_off + p.out_off))[r] = v;   // staged for peer_store_rows
        }
    }
}

// gate/up pair (SwiGLU, LlamaMLP.forward model/LlamaModel.kt:159-176): rows [0, n) of the partial-sum buffer hold segment 0
// (gate), rows [off, off + n) segment 1 (up).  Stores g, u, silu(g) and silu(g)*u with the elementwise kernels' roundings.
__device__ __forceinline__ void finish_pair(const GemvParams &p, const float *red, int n, int off, int64_t out0, int warp, int lane) {
    const int sub = lane & 7, rsel = lane >> 3;
    const int nc4 = p.red_stride >> 2;
#pragma unroll 1
    for (int rb = warp * 4; rb < n; rb += p.cwarps * 4) {
        const int r = rb + rsel;
        float g = 0.f, u = 0.f;
        if (r < n) {
            const float4 *rg_ = (const float4 *)(red + (size_t)r * p.red_stride);
            const float4 *ru_ = (const float4 *)(red + (size_t)(off + r) * p.red_stride);
#pragma unroll 2
            for (int c = sub; c < nc4; c += 8) {
                const float4 a = rg_[c], b = ru_[c];
                g += (a.x + a.y) + (a.z + a.w);
                u += (b.x + b.y) + (b.z + b.w);
            }
        }
#pragma unroll
        for (int o = 1; o < 8; o <<= 1) { g += __shfl_xor_sync(0xffffffffu, g, o); u += __shfl_xor_sync(0xffffffffu, u, o); }
        if (sub == 0 && r < n) {
            const int64_t o = out0 + r;
            const float act = silu_f(g);
            p.dst[o] = g;
            p.dst_seg[0][o] = u;
            if (p.epi_act) p.epi_act[o] = act;
            const float hv = __fmul_rn(act, u);
            p.epi_out[o] = hv;
            if (p.npeers) ((float *)((uint8_t *)red - p.red_off + p.out_off))[r] = hv;   // sharded FFN: staged for peer_store_rows
        }
    }
}

// MINB = CTAs of consecutive launches that must fit on one SM together (registers are
// capped accordingly): with programmatic dependent launch the next GEMV's CTA sits beside
// this one and streams its weights while this one computes.
// FUSED = false is the plain computeMatMul launch: no prologue, one weight segment, plain store (the fused
// variants cost the plain one ~8 % when they share its code: measured, profiles/r01j_gemv_ab.txt).
// MODE 0: the plain launch.  3: the plain launch that is independent of everything in flight (WAIT_END; its activations
// are preloaded at kernel entry).  1: every fusion (prologue, segments, epilogues) + peers + exchange.  2: the plain launch with
// peer stores and the exchange signal / wait only (lk_compute_matmul_sharded: the 28-warp instantiation of mode 1 spills
// under its 64-register cap, this one does not).
template <int TYPE, int MAXW, int MINB, int NP, int MODE>
__global__ void __launch_bounds__((MAXW + 1) * 32, MINB) gemv_kernel(const __grid_constant__ GemvParams p) {
    constexpr bool FUSED = MODE == 1, PEER = MODE == 1 || MODE == 2;
    constexpr bool INDEP = MODE == 3;     // MODE 3: the plain launch in WAIT_END mode (independent of everything in flight), known at compile time
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t *full = (uint64_t *)smem;                   // [GEMV_MAX_STAGES]
    uint64_t *empty = full + GEMV_MAX_STAGES;            // [GEMV_MAX_STAGES]
    float *red = (float *)(smem + p.red_off);            // [red_rows][red_stride]
    uint32_t *xs = (uint32_t *)(smem + p.x_off);         // [nblk][gemv_xslot_words(NP)]
    uint8_t *stage0 = smem + p.stage_off;
    GemvSide *side = (GemvSide *)(smem + GEMV_SIDE_OFF);

    // The next kernel of the stream may be scheduled beside us as soon as every CTA has signalled.  An independent
    // launch (WAIT_END) signals at once; a launch that waits for its predecessor signals only after that wait, so
    // that whatever starts next finds everything before this launch complete (the host-side bookkeeping,
    // gemv_wait_mode in lk_shim.cu, closes its in-flight set at exactly that point).
    if (INDEP || p.wait_mode == GEMV_WAIT_END) pdl_launch_dependents();
    {   // touch both hot parameter lines now: the two constant-cache misses overlap instead of serialising later
        const int t0 = p.K, t1 = p.red_stride;
        asm volatile("" ::"r"(t0), "r"(t1));
    }
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int stage_bytes = p.stage_rows * p.row_bytes;
    const int nseg = FUSED ? p.nseg : 1;
    const int epi_mode = FUSED ? p.epi_mode : 0;

    // balanced contiguous row range of this CTA: the first `rem` CTAs own one unit more
    const int ub = p.units_per_cta * (int)blockIdx.x + min((int)blockIdx.x, p.units_rem);
    const int r_begin = ub * p.unit_rows;
    const int nrows = (p.units_per_cta + ((int)blockIdx.x < p.units_rem ? 1 : 0)) * p.unit_rows;

    if (warp == p.cwarps) {
        // ===== producer warp: lane 0 initialises the barriers and streams this CTA's byte range =====
        const int64_t seg_off = (int64_t)r_begin * p.row_bytes;
        const uint8_t *src = p.a + seg_off;
        int s = 0, left = nrows, seg = 0;
        uint8_t *dstp = stage0;
        uint64_t pol_w = 0;
        // next chunk of this CTA's byte stream: its row range of segment 0, then of segment 1, ...; a stage never
        // spans two segments (the consumers chunk the same way)
        auto advance = [&]() -> bool {
            if (left > 0) return true;
            if (!FUSED || ++seg >= nseg) return false;
            src = p.a_seg[seg - 1] + seg_off;
            left = nrows;
            return true;
        };
        if (lane == 0) {
            LK_TL(0);
#pragma unroll 1
            for (int i = 0; i < p.nstages; i++) {
                mbar_init(&full[i], 1);
                mbar_init(&empty[i], p.cwarps);
            }
            fence_mbar_init();
            if (!INDEP && p.wait_mode == GEMV_WAIT_FULL) pdl_wait();
            pol_w = l2_policy_evict_first();
#pragma unroll 1
            for (; s < p.nstages && advance(); s++) {   // first pass over the ring: no waiting
                const int rows = min(p.stage_rows, left);
                const uint32_t bytes = (uint32_t)(rows * p.row_bytes);
                mbar_arrive_expect_tx(&full[s], bytes);
                tma_load_1d(dstp, src, bytes, &full[s], pol_w);
                src += bytes; left -= rows; dstp += stage_bytes;
            }
            LK_TL(1);
        }
        __syncwarp();
        __syncthreads();   // consumers may now wait on the barriers
        if (lane == 0 && s == p.nstages) {
            int round = 1;
            s = 0; dstp = stage0;
#pragma unroll 1
            while (advance()) {
                mbar_wait(&empty[s], (round - 1) & 1);
                const int rows = min(p.stage_rows, left);
                const uint32_t bytes = (uint32_t)(rows * p.row_bytes);
                mbar_arrive_expect_tx(&full[s], bytes);
                tma_load_1d(dstp, src, bytes, &full[s], pol_w);
                src += bytes; left -= rows; dstp += stage_bytes;
                if (++s == p.nstages) { s = 0; round++; dstp = stage0; }
            }
            LK_TL(2);
        }
        return;
    }

    // ===== consumer warps =====
    // Everything that does not depend on the activations happens before griddepcontrol.wait.
    const int nct = p.cwarps * 32;
    // an independent launch (WAIT_END) may read its activations at once: issue the first round's loads now
    float4 pre[4];
    constexpr bool use_pre = INDEP;
    if (use_pre && tid < p.nblk * 2) {
        const float4 *src = (const float4 *)(p.b + (size_t)tid * 16);
#pragma unroll
        for (int i = 0; i < 4; i++) pre[i] = __ldg(src + i);
    }
    const int cw = p.rg == 1 ? warp : warp % p.ncw;   // column warp
    const int rgi = p.rg == 1 ? 0 : warp / p.ncw;     // row group
    const int col = cw * 32 + lane;                   // block column
    const bool active = col < p.nblk;
    const int colc = active ? col : 0;                // idle lanes redo column 0 and contribute 0
    // zero the partial-sum buffer's padding columns once (finish_rows reads whole rows)
#pragma unroll 1
    for (int r = warp; r < p.red_rows; r += p.cwarps)
        if (lane < GEMV_RED_PAD) red[r * p.red_stride + p.ncw * 8 + lane] = 0.f;
    if (warp == 0) { if (lane < 28) side->mask[lane] = 0u; if (lane == 28) side->any = 0u; }
    uint32_t lut_addr = 0;
    if (TYPE == T_BITNET_1_58) {            // the CTA's byte -> five-trit table (read after the bar.sync that follows prep_x)
        lut_addr = smem_u32(smem + p.lut_off);
        const int lsh = p.lut_shift;
#pragma unroll 4
        for (int i = tid; i < (256 << lsh); i += nct) {        // item i = entry i >> lsh, copy i & (copies - 1), at byte 8 i: conflict-free stores
            const uint2 e = __ldg((const uint2 *)g_bitnet_lut.w + (i >> lsh));
            asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(lut_addr + 8u * i), "r"(e.x), "r"(e.y) : "memory");
        }
        lut_addr += 8u * (lane & ((1 << lsh) - 1));
    }
    unsigned long long xticket = 0;
    if (PEER && p.wait_n > 0 && tid == 0) xticket = atomicAdd(p.wait_ticket, 1ull);   // issued early: its latency hides under the barrier below
    __syncthreads();   // barriers initialised by the producer
    if (!INDEP && p.wait_mode != GEMV_WAIT_END) {
        pdl_wait();                // activations may be the previous kernel's output
        pdl_launch_dependents();   // the next kernel can start its prologue / weight prefetch under our main loop
    }
    if (PEER && p.wait_n > 0) {   // sharded matmul: the activations are gathered from every rank (peer stores + arrival counters)
        if (tid == 0) exchange_wait(p, xticket);
        asm volatile("bar.sync 1, %0;" ::"r"(nct) : "memory");
    }
    if (tid == 0) LK_TL(3);
    prep_x<TYPE, NP, FUSED, PEER ? 2 : 0, LK_GEMV_DETECT>(p, xs, tid, nct, side, INDEP ? pre : nullptr, use_pre);   // fused launches: plain L1-allocating loads (see ld_vec4)
    if (tid == 0) LK_TL(24);
    asm volatile("bar.sync 1, %0;" ::"r"(nct) : "memory");   // activation planes written
    if (tid == 0) LK_TL(4);
    // range fix-up (cold): some block holds elements its block scale cannot represent; see GemvSide.  The flag is re-read
    // from shared memory at each site so that nothing extra stays live across the stage loop.
#ifdef LK_GEMV_NO_SITES
#define LK_GEMV_FIXUP(...)
#else
#define LK_GEMV_FIXUP(AW, ROW_FIRST, NROWS, REDP)                                                                          \
    if (gemv_has_fixup<TYPE>() && __builtin_expect(*(volatile uint32_t *)(smem + GEMV_SIDE_OFF + 4) != 0u, 0)) {                                \
        gemv_fixup<TYPE, NP, PEER ? 2 : 0, FUSED, 0>(&p, AW, ROW_FIRST, NROWS, REDP, smem);                              \
        asm volatile("bar.sync 1, %0;" ::"r"(nct) : "memory");                                                           \
        load_xregs<TYPE, NP>(x, xs, colc);      /* nothing of x has to survive the call */                                \
    }
#endif

    XRegs<NP> x;
    x.aux = lut_addr;
    x.aux_sh = 3u + (uint32_t)p.lut_shift;
    load_xregs<TYPE, NP>(x, xs, colc);
    if (tid == 0) LK_TL(5);

    // Stage loop: one barrier wait / release per stage; inside, this warp's rows (rgi, rgi+rg, ...)
    // are processed UNR at a time so that UNR independent dot-product chains are in flight.
    constexpr int UNR = TYPE == T_BITNET_1_58 ? LK_BITNET_UNR : 4;    // (BitNet: 6 and 8 rows per step measured no faster, profiles/r02g_bitnet_gemv.txt)
    const int rg_words = p.rg * (p.row_bytes >> 2);
    int red_base = 0, s = 0, par = 0, it = 0;
    const uint32_t *sp = (const uint32_t *)stage0 + (size_t)rgi * (p.row_bytes >> 2);   // my first row of the stage
    float *const rp0 = red + (size_t)rgi * p.red_stride + (col >> 2);
    float *rp = rp0;
    const int rstep = p.rg * p.red_stride;
#pragma unroll 1
    for (int seg = 0; seg < nseg; seg++) {
    float *const seg_dst = (!FUSED || seg == 0) ? p.dst : p.dst_seg[seg - 1];
    if (epi_mode == 1 && seg == 1) rp = rp0 + (size_t)p.pair_off * p.red_stride;   // the host sized the buffer for both
#pragma unroll 1
    for (int row0 = 0; row0 < nrows; row0 += p.stage_rows, it++) {
        if (epi_mode != 1 && row0 + p.stage_rows - red_base > p.red_rows) {
            // partial-sum buffer full (a CTA that owns more rows than it holds)
            asm volatile("bar.sync 1, %0;" ::"r"(nct) : "memory");
            LK_GEMV_FIXUP((!FUSED || seg == 0) ? p.a : p.a_seg[seg - 1], r_begin + red_base, row0 - red_base, red)
            finish_rows<FUSED, false, PEER>(p, seg_dst, red, row0 - red_base, p.row_offset + r_begin + red_base, warp, lane);
            if (PEER && p.npeers) {
                asm volatile("bar.sync 1, %0;" ::"r"(nct) : "memory");
                peer_store_rows(p, (const float *)(smem + p.out_off), row0 - red_base, p.row_offset + r_begin + red_base, warp, lane);
            }
            asm volatile("bar.sync 1, %0;" ::"r"(nct) : "memory");
            rp -= (size_t)(row0 - red_base) * p.red_stride;
            red_base = row0;
        }
        const int rows_here = min(p.stage_rows, nrows - row0);
        int nr = (rows_here - rgi + p.rg - 1) >> p.rg_shift;   // my rows in this stage
        mbar_wait(&full[s], par);
        if (tid == 0 && it < 6) LK_TL(8 + 2 * it);
        const uint32_t *wp = sp;
        float *rq = rp;
#pragma unroll 1
        for (; nr >= UNR; nr -= UNR) {
            float v[UNR];
#pragma unroll
            for (int u = 0; u < UNR; u++) v[u] = block_dot<TYPE, NP>(wp + u * rg_words, colc, x);
#pragma unroll
            for (int u = 0; u < UNR; u++) v[u] = active ? v[u] : 0.f;
#pragma unroll
            for (int u = 0; u < UNR; u++) v[u] += __shfl_xor_sync(0xffffffffu, v[u], 1);
#pragma unroll
            for (int u = 0; u < UNR; u++) v[u] += __shfl_xor_sync(0xffffffffu, v[u], 2);
            if ((lane & 3) == 0) {
#pragma unroll
                for (int u = 0; u < UNR; u++) rq[u * rstep] = v[u];     // one partial per 4 block columns
            }
            wp += UNR * rg_words;
            rq += UNR * rstep;
        }
        // the rows left over: two at a time, then one
#ifdef LK_GEMV_TAIL2
        if (nr >= 2) {
            float v[2];
#pragma unroll
            for (int u = 0; u < 2; u++) v[u] = block_dot<TYPE, NP>(wp + u * rg_words, colc, x);
#pragma unroll
            for (int u = 0; u < 2; u++) v[u] = active ? v[u] : 0.f;
#pragma unroll
            for (int u = 0; u < 2; u++) v[u] += __shfl_xor_sync(0xffffffffu, v[u], 1);
#pragma unroll
            for (int u = 0; u < 2; u++) v[u] += __shfl_xor_sync(0xffffffffu, v[u], 2);
            if ((lane & 3) == 0) { rq[0] = v[0]; rq[rstep] = v[1]; }
            wp += 2 * rg_words;
            rq += 2 * rstep;
            nr -= 2;
        }
#endif
#pragma unroll 1
        for (; nr > 0; nr--) {
            float v = block_dot<TYPE, NP>(wp, colc, x);
            v = active ? v : 0.f;
            v += __shfl_xor_sync(0xffffffffu, v, 1);
            v += __shfl_xor_sync(0xffffffffu, v, 2);
            if ((lane & 3) == 0) *rq = v;
            wp += rg_words;
            rq += rstep;
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);  // this warp is done reading the stage
        if (tid == 0 && it < 6) LK_TL(9 + 2 * it);
        sp += (size_t)p.stage_rows * (p.row_bytes >> 2);
        rp += (size_t)p.stage_rows * p.red_stride;
        if (++s == p.nstages) { s = 0; par ^= 1; sp -= (size_t)p.nstages * p.stage_rows * (p.row_bytes >> 2); }
    }
    if (epi_mode == 1) continue;                // the pair is finished together below
    if (tid == 0) LK_TL(6);
    asm volatile("bar.sync 1, %0;" ::"r"(nct) : "memory");
    if (tid == 0) LK_TL(20);
    LK_GEMV_FIXUP((!FUSED || seg == 0) ? p.a : p.a_seg[seg - 1], r_begin + red_base, nrows - red_base, red)
    finish_rows<FUSED, false, PEER>(p, seg_dst, red, nrows - red_base, p.row_offset + r_begin + red_base, warp, lane);
    if (PEER && p.npeers) {
        asm volatile("bar.sync 1, %0;" ::"r"(nct) : "memory");
        peer_store_rows(p, (const float *)(smem + p.out_off), nrows - red_base, p.row_offset + r_begin + red_base, warp, lane);
    }
    if (seg + 1 < nseg) {                        // the buffer is reused by the next segment
        asm volatile("bar.sync 1, %0;" ::"r"(nct) : "memory");
        rp = rp0;
        red_base = 0;
    }
    }
    if (epi_mode == 1) {
        asm volatile("bar.sync 1, %0;" ::"r"(nct) : "memory");
        LK_GEMV_FIXUP(p.a, r_begin, nrows, red)
        LK_GEMV_FIXUP(p.a_seg[0], r_begin, nrows, red + (size_t)p.pair_off * p.red_stride)
        finish_pair(p, red, nrows, p.pair_off, p.row_offset + r_begin, warp, lane);
        if (p.npeers) {
            asm volatile("bar.sync 1, %0;" ::"r"(nct) : "memory");
            peer_store_rows(p, (const float *)(smem + p.out_off), nrows, p.row_offset + r_begin, warp, lane);
        }
    }
    if (PEER && p.sig_n > 0) exchange_signal(p, tid, nct);   // every CTA arrives on every rank's counter
    if (tid == 0) {
        LK_TL(7);
        // keep completion order = launch order so a later launch's wait covers every earlier one
        if (INDEP || p.wait_mode == GEMV_WAIT_END) pdl_wait();
    }
}

// ---------------------------------------------------------------------------
// K1c: a CHAIN of dependent GEMVs as one persistent launch.
//
// A decoder layer at batch 1 is four GEMV launches, each of which needs the whole output of the one before
// ([RMSNorm -> Q/K/V], [O + residual], [RMSNorm -> gate/up + SwiGLU], [down + residual], model/LlamaModel.kt:253-285).
// As separate launches every hand-off costs a grid completion + launch + parameter fetch + barrier set-up, and the
// weight stream stops meanwhile.  Here one CTA per SM runs the phases back to back:
//   * the producer warp streams the CTA's row range of phase 0, then of phase 1, ... through ONE ring of equally
//     sized slots and never waits for anything but a free slot: weights do not depend on activations, so while the
//     consumers sit in the hand-off the ring fills with the next phase's rows;
//   * the consumers separate phases with a grid-wide barrier (arrive counter + generation word in global memory, the
//     cooperative-groups pattern: bar.sync, one thread fences, arrives and polls) and read every vector another CTA
//     may have written with ld.global.cg;
//   * all CHAIN_W consumer warps take part in every phase's prologue, barriers and epilogue.  In the stage loop they
//     work as `ng` independent groups of rg * ncw warps (K = 4096: two groups of 2 x 4; K = 11008: one group of 11,
//     five warps without rows): group g computes stages g, g + ng, ... so that, as with two CTAs per SM, one group's
//     barrier round or slow warp does not stall the other; every working warp waits for and releases every slot, in
//     order, and warp 0 also releases it for the warps without rows.  Measured (profiles/r01t_gemv_chain_ab.txt): 12
//     consumer warps with 128 registers beat 15 and 16 (96 registers: spills in the dot-product loop).
// Results are bit-identical to the same launches issued one by one (same block_dot, same reduction order).
constexpr int CHAIN_MAX = 8;      // phases per launch (2 decoder layers without attention)
constexpr int CHAIN_W = 12;       // consumer warps (K up to 12288; with the producer 13 warps: 128 registers each)
struct GemvChainParams {
    GemvParams ph[CHAIN_MAX];     // cwarps = CHAIN_W, red_off / x_off / stage_off / nstages identical in every phase
    int nphase;
    int slot_bytes;               // ring slot size (>= every phase's stage_rows * row_bytes)
    int wait_mode;                // GEMV_WAIT_FULL: a weight was written by an op that may still run -> the producer waits too
    unsigned *bar;                // words 0-1: 64-bit arrival counter (8-byte aligned), word 2: time-outs (never in a healthy run)
};
static_assert(sizeof(GemvChainParams) <= 4000, "kernel parameter space");
static_assert(sizeof(GemvParams) >= 260, "the parameter prefetch below touches five 64-byte lines");

__device__ __forceinline__ unsigned long long ld_acquire_gpu(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_gpu_add(unsigned long long *p, unsigned long long v) {
    asm volatile("red.release.gpu.global.add.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

template <int TYPE, int NP>
__global__ void __launch_bounds__((CHAIN_W + 1) * 32, 1) gemv_chain_kernel(const __grid_constant__ GemvChainParams cp) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t *full = (uint64_t *)smem;                   // [GEMV_MAX_STAGES]
    uint64_t *empty = full + GEMV_MAX_STAGES;            // [GEMV_MAX_STAGES]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int bid = (int)blockIdx.x;
    const int nslots = cp.ph[0].nstages;
    float *red = (float *)(smem + cp.ph[0].red_off);
    uint32_t *xs = (uint32_t *)(smem + cp.ph[0].x_off);
    uint8_t *ring = smem + cp.ph[0].stage_off;
    GemvSide *side = (GemvSide *)(smem + GEMV_SIDE_OFF);
    constexpr int NCT = CHAIN_W * 32;

    if (warp == CHAIN_W) {
        // ===== producer warp: one thread streams this CTA's rows of every phase, in phase order =====
        if (lane == 0) {
#pragma unroll 1
            for (int i = 0; i < nslots; i++) {
                mbar_init(&full[i], 1);
                mbar_init(&empty[i], CHAIN_W);
            }
            fence_mbar_init();
        }
        __syncwarp();
        __syncthreads();   // consumers may now wait on the barriers
        if (lane == 0) {
            if (cp.wait_mode == GEMV_WAIT_FULL) pdl_wait();
            const uint64_t pol_w = l2_policy_evict_first();
            int s = 0, round = 0;
#pragma unroll 1
            for (int ph = 0; ph < cp.nphase; ph++) {
                const GemvParams &p = cp.ph[ph];
                const int ub = p.units_per_cta * bid + min(bid, p.units_rem);
                const int nrows = (p.units_per_cta + (bid < p.units_rem ? 1 : 0)) * p.unit_rows;
                const int64_t seg_off = (int64_t)ub * p.unit_rows * p.row_bytes;
#pragma unroll 1
                for (int seg = 0; seg < p.nseg; seg++) {
                    const uint8_t *src = (seg == 0 ? p.a : p.a_seg[seg - 1]) + seg_off;
#pragma unroll 1
                    for (int left = nrows; left > 0;) {
                        if (round > 0) mbar_wait(&empty[s], (round - 1) & 1);
                        const int rows = min(p.stage_rows, left);
                        const uint32_t bytes = (uint32_t)(rows * p.row_bytes);
                        mbar_arrive_expect_tx(&full[s], bytes);
                        tma_load_1d(ring + (size_t)s * cp.slot_bytes, src, bytes, &full[s], pol_w);
                        src += bytes; left -= rows;
                        if (++s == nslots) { s = 0; round++; }
                    }
                }
                LK_CTL(ph, 7);     // the phase's last copy is issued
            }
        }
        return;
    }

    // ===== consumer warps =====
    if (warp == 0) { if (lane < 28) side->mask[lane] = 0u; if (lane == 28) side->any = 0u; }
    __syncthreads();   // barriers initialised by the producer
#ifdef LK_GEMV_TIMELINE
    if (tid == 0 && cp.ph[0].tl) { unsigned long long *q_ = cp.ph[0].tl + (size_t)blockIdx.x * LK_CTL_STRIDE; q_[0] = lk_gtime(); q_[1] = (unsigned long long)clock64(); q_[2] = lk_smid(); }
#endif
    pdl_wait();                // phase 0's activations may be the previous kernel's output
    pdl_launch_dependents();   // the successor waits for this whole grid anyway; let it be scheduled early
    // Grid barrier = one 64-bit arrival counter that only ever grows: every launch adds (nphase - 1) * gridDim.x, so between
    // launches it is a multiple of gridDim.x, and a CTA that reads it late (others have already arrived at barrier 0, which
    // cannot complete without this CTA) still finds the same base by rounding down.  Barrier k is complete when the counter
    // reaches base + (k + 1) * gridDim.x: one release-add and one acquire-poll per CTA, no second hop through a last arriver.
    unsigned long long *const bar_cnt = (unsigned long long *)cp.bar;
    unsigned long long bar_base = 0;
    if (tid == 0) {
        bar_base = ld_acquire_gpu(bar_cnt);
        bar_base -= bar_base % gridDim.x;
    }
    int s = 0, par = 0;
    constexpr int UNR = 4;
#pragma unroll 1
    for (int ph = 0; ph < cp.nphase; ph++) {
        const GemvParams &p = cp.ph[ph];
        if (tid == 0) LK_CTL(ph, 0);
        if (ph > 0) {
            // ---- grid barrier: every CTA has stored the previous phase's outputs ----
            asm volatile("bar.sync 1, %0;" ::"r"(NCT) : "memory");
            if (tid == 0) {
                LK_CTL(ph, 1);
                red_release_gpu_add(bar_cnt, 1ull);
                const unsigned long long target = bar_base + (unsigned long long)ph * gridDim.x;
                unsigned polls = 0;
                while ((long long)(ld_acquire_gpu(bar_cnt) - target) < 0) {
                    if (++polls > (1u << 22)) { atomicAdd(cp.bar + 2, 1u); break; }   // never hang the device
                }
                LK_CTL(ph, 2);
            }
            asm volatile("bar.sync 1, %0;" ::"r"(NCT) : "memory");
        }
        if (tid == 32 && ph + 1 < cp.nphase) {   // warm the constant cache with the next phase's parameter lines
            const int *q = (const int *)&cp.ph[ph + 1];
            const int t0 = q[0], t1 = q[16], t2 = q[32], t3 = q[48], t4 = q[64];
            asm volatile("" ::"r"(t0), "r"(t1), "r"(t2), "r"(t3), "r"(t4));
        }
        const int ub = p.units_per_cta * bid + min(bid, p.units_rem);
        const int r_begin = ub * p.unit_rows;
        const int nrows = (p.units_per_cta + (bid < p.units_rem ? 1 : 0)) * p.unit_rows;
        const int gwarps = p.ncw << p.rg_shift;           // warps per group
        const int grp = warp / gwarps;                    // my group
        const int wig = warp - grp * gwarps;
        const int cw = wig % p.ncw;                       // column warp
        const int rgi = wig / p.ncw;                      // row group within my group
        const bool working = grp < p.ng;                  // warps beyond ng groups have no rows in this phase: they skip the
                                                          // ring altogether (no polling beside the working warps) and warp 0
                                                          // releases every slot on their behalf
        const uint32_t my_arrivals = warp == 0 ? (uint32_t)(1 + CHAIN_W - p.ng * gwarps) : 1u;
        int turn = 0;                                     // whose stage comes next
        const int col = cw * 32 + lane;                   // block column
        const bool active = working && col < p.nblk;
        const int colc = col < p.nblk ? col : 0;
        const int row_words = p.row_bytes >> 2;
        const int nseg = p.nseg, epi_mode = p.epi_mode;
#pragma unroll 1
        for (int r = warp; r < p.red_rows; r += CHAIN_W)
            if (lane < GEMV_RED_PAD) red[r * p.red_stride + p.ncw * 8 + lane] = 0.f;
        prep_x<TYPE, NP, true, 1, LK_CHAIN_DETECT>(p, xs, tid, NCT, side);
        asm volatile("bar.sync 1, %0;" ::"r"(NCT) : "memory");   // activation planes (and the padding zeros) written
        if (tid == 0) LK_CTL(ph, 3);

        XRegs<NP> x;
        load_xregs<TYPE, NP>(x, xs, colc);
#ifdef LK_GEMV_TIMELINE
        bool first_stage = true;
#endif

        const int rg_words = p.rg * row_words;
        const int rstep = p.rg * p.red_stride;
        int red_base = 0;
        float *const rp0 = red + (size_t)rgi * p.red_stride + (col >> 2);
        float *rp = rp0;
#pragma unroll 1
        for (int seg = 0; seg < nseg; seg++) {
            float *const seg_dst = seg == 0 ? p.dst : p.dst_seg[seg - 1];
            if (epi_mode == 1 && seg == 1) rp = rp0 + (size_t)p.pair_off * p.red_stride;
#pragma unroll 1
            for (int row0 = 0; row0 < nrows; row0 += p.stage_rows) {
                if (epi_mode != 1 && row0 + p.stage_rows - red_base > p.red_rows) {
                    // partial-sum buffer full (a CTA that owns more rows than it holds)
                    asm volatile("bar.sync 1, %0;" ::"r"(NCT) : "memory");
                    LK_CHAIN_FIXUP(true, p.pro_w, p.pro_mode, NCT, seg == 0 ? p.a : p.a_seg[seg - 1], r_begin + red_base, row0 - red_base, red)
                    finish_rows<true, true>(p, seg_dst, red, row0 - red_base, p.row_offset + r_begin + red_base, warp, lane);
                    asm volatile("bar.sync 1, %0;" ::"r"(NCT) : "memory");
                    rp -= (size_t)(row0 - red_base) * p.red_stride;
                    red_base = row0;
                }
                const int rows_here = min(p.stage_rows, nrows - row0);
                int nr = (working && turn == grp) ? (rows_here - rgi + p.rg - 1) >> p.rg_shift : 0;   // my rows in this stage
                if (++turn == p.ng) turn = 0;
                if (working) mbar_wait(&full[s], par);
#ifdef LK_GEMV_TIMELINE
                if (first_stage && tid == 0) { LK_CTL(ph, 4); first_stage = false; }
#endif
                const uint32_t *wp = (const uint32_t *)(ring + (size_t)s * cp.slot_bytes) + (size_t)rgi * row_words;
                float *rq = rp;
#pragma unroll 1
                for (; nr >= UNR; nr -= UNR) {
                    float v[UNR];
#pragma unroll
                    for (int u = 0; u < UNR; u++) v[u] = block_dot<TYPE, NP>(wp + u * rg_words, colc, x);
#pragma unroll
                    for (int u = 0; u < UNR; u++) v[u] = active ? v[u] : 0.f;
#pragma unroll
                    for (int u = 0; u < UNR; u++) v[u] += __shfl_xor_sync(0xffffffffu, v[u], 1);
#pragma unroll
                    for (int u = 0; u < UNR; u++) v[u] += __shfl_xor_sync(0xffffffffu, v[u], 2);
                    if ((lane & 3) == 0) {
#pragma unroll
                        for (int u = 0; u < UNR; u++) rq[u * rstep] = v[u];     // one partial per 4 block columns
                    }
                    wp += UNR * rg_words;
                    rq += UNR * rstep;
                }
#pragma unroll 1
                for (; nr > 0; nr--) {
                    float v = block_dot<TYPE, NP>(wp, colc, x);
                    v = active ? v : 0.f;
                    v += __shfl_xor_sync(0xffffffffu, v, 1);
                    v += __shfl_xor_sync(0xffffffffu, v, 2);
                    if ((lane & 3) == 0) *rq = v;
                    wp += rg_words;
                    rq += rstep;
                }
                __syncwarp();
                if (working && lane == 0) mbar_arrive_n(&empty[s], my_arrivals);  // this warp is done reading the slot
                rp += (size_t)p.stage_rows * p.red_stride;
                if (++s == nslots) { s = 0; par ^= 1; }
            }
            if (tid == 0 && seg + 1 == nseg) LK_CTL(ph, 5);
            if (epi_mode == 1) continue;                // the pair is finished together below
            asm volatile("bar.sync 1, %0;" ::"r"(NCT) : "memory");
            LK_CHAIN_FIXUP(true, p.pro_w, p.pro_mode, NCT, seg == 0 ? p.a : p.a_seg[seg - 1], r_begin + red_base, nrows - red_base, red)
            finish_rows<true, true>(p, seg_dst, red, nrows - red_base, p.row_offset + r_begin + red_base, warp, lane);
            if (seg + 1 < nseg) {                        // the buffer is reused by the next segment
                asm volatile("bar.sync 1, %0;" ::"r"(NCT) : "memory");
                rp = rp0;
                red_base = 0;
            }
        }
        if (epi_mode == 1) {
            asm volatile("bar.sync 1, %0;" ::"r"(NCT) : "memory");
            LK_CHAIN_FIXUP(true, p.pro_w, p.pro_mode, NCT, p.a, r_begin, nrows, red)
            LK_CHAIN_FIXUP(true, p.pro_w, p.pro_mode, NCT, p.a_seg[0], r_begin, nrows, red + (size_t)p.pair_off * p.red_stride)
            finish_pair(p, red, nrows, p.pair_off, p.row_offset + r_begin, warp, lane);
        }
        // re-arm the range flags for the next phase (its prep_x runs after the grid barrier's bar.sync)
        if (warp == 0 && side->any != 0u) { if (lane < 28) side->mask[lane] = 0u; if (lane == 28) side->any = 0u; }
        if (tid == 0) LK_CTL(ph, 6);
    }
#ifdef LK_GEMV_TIMELINE
    if (tid == 0 && cp.ph[0].tl) cp.ph[0].tl[(size_t)blockIdx.x * LK_CTL_STRIDE + 3] = lk_gtime();
#endif
}

// ---------------------------------------------------------------------------
// host side: shape analysis + launch
// ---------------------------------------------------------------------------
template <int TYPE, int MAXW, int MINB, int NP, int MODE>
static cudaError_t launch_one(const GemvParams &p, int grid, int smem, int pdl, cudaStream_t s) {
    auto kern = gemv_kernel<TYPE, MAXW, MINB, NP, MODE>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    if (e != cudaSuccess) return e;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3((unsigned)((p.cwarps + 1) * 32));
    cfg.dynamicSmemBytes = (size_t)smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, p);
}

template <int TYPE, int NP, int MODE>
static cudaError_t launch_fused(const GemvParams &p, int grid, int smem, int pdl, cudaStream_t s) {
    if (p.cwarps <= 8) return launch_one<TYPE, 8, 2, NP, MODE>(p, grid, smem, pdl, s);
    if (p.cwarps <= 12) return launch_one<TYPE, 12, 2, NP, MODE>(p, grid, smem, pdl, s);
    if (p.cwarps <= 16) return launch_one<TYPE, 16, 2, NP, MODE>(p, grid, smem, pdl, s);
    return launch_one<TYPE, 28, 1, NP, MODE>(p, grid, smem, pdl, s);   // K up to 28672 (Llama-2-70B down projection): one CTA per SM
}

template <int TYPE, int NP>
static cudaError_t launch_type(const GemvParams &p, int mode, int grid, int smem, int pdl, cudaStream_t s) {
    return mode == 1 ? launch_fused<TYPE, NP, 1>(p, grid, smem, pdl, s) : mode == 2 ? launch_fused<TYPE, NP, 2>(p, grid, smem, pdl, s)
         : mode == 3 ? launch_fused<TYPE, NP, 3>(p, grid, smem, pdl, s) : launch_fused<TYPE, NP, 0>(p, grid, smem, pdl, s);
}
template <int TYPE, int NP>
static cudaError_t launch_plain(const GemvParams &p, int grid, int smem, int pdl, cudaStream_t s) {   // K-quants: no fused variants
    return launch_fused<TYPE, NP, 0>(p, grid, smem, pdl, s);
}

cudaError_t launch_gemv(int type, const void *a, int64_t K, int64_t M, const float *b, float *dst,
                        int64_t row_offset, const GemvPeers &peers, const GemvTuning &tune, int sm_count,
                        cudaStream_t s, const GemvFusion *fusion, const GemvSync *sync) {
    const bool bitnet = type == T_BITNET_1_58;                              // 10-byte blocks of 32 ternary weights
    const bool kq = type == T_Q4_K || type == T_Q8_K || type == T_Q2_K || bitnet;     // the reference's K-quant super-blocks (256 elements) / BitNet: plain launches only
    if (kq && ((!bitnet && (K % 256) != 0) || fusion)) return cudaErrorNotSupported;
    const int bb = bitnet ? 10 : block_bytes(type);
    if ((bb == 0 && !kq) || K <= 0 || M <= 0 || (K % 64) != 0 || K > 32 * 32 * 28 || M > (1 << 30))
        return cudaErrorNotSupported;
    if (((uintptr_t)a & 15) || ((uintptr_t)b & 15)) return cudaErrorNotSupported;

    GemvParams p = {};
    p.a = (const uint8_t *)a; p.b = b; p.dst = dst;
    p.npeers = (short)peers.n;
    p.nib = (short)((tune.nibble_order && !kq) ? 1 : 0);
    for (int i = 0; i < peers.n && i < 8; i++) p.peer[i] = peers.ptr[i];
    p.K = (int)K; p.M = (int)M; p.row_offset = row_offset;
    p.nseg = 1;
    if (fusion) {
        if (((uintptr_t)fusion->pro_w & 15) || ((uintptr_t)fusion->b_out & 15)) return cudaErrorNotSupported;
        p.pro_mode = fusion->pro_mode; p.pro_w = fusion->pro_w; p.pro_eps = fusion->pro_eps; p.b_out = fusion->b_out;
        p.residual = fusion->residual; p.dst_raw = fusion->dst_raw;
        if (fusion->nseg_extra < 0 || fusion->nseg_extra > 2) return cudaErrorNotSupported;
        // stacked segments take peers / a row offset only as the SwiGLU pair (sharded FFN: the peers then are copies of epi_out)
        if (fusion->nseg_extra > 0 && (fusion->residual || ((peers.n > 0 || row_offset != 0) && fusion->epi_mode != 1))) return cudaErrorNotSupported;
        for (int i = 0; i < fusion->nseg_extra; i++) {
            if (!fusion->a_extra[i] || !fusion->dst_extra[i] || ((uintptr_t)fusion->a_extra[i] & 15)) return cudaErrorNotSupported;
            p.a_seg[i] = (const uint8_t *)fusion->a_extra[i];
            p.dst_seg[i] = fusion->dst_extra[i];
        }
        p.nseg = 1 + fusion->nseg_extra;
        if (fusion->epi_mode == 1) {
            if (fusion->nseg_extra != 1 || !fusion->epi_out) return cudaErrorNotSupported;
            p.epi_mode = 1; p.epi_act = fusion->epi_act; p.epi_out = fusion->epi_out;
        }
    }
    if (sync) {
        if (kq || sync->sig_n < 0 || sync->sig_n > 8 || sync->wait_n < 0) return cudaErrorNotSupported;
        p.sig_n = sync->sig_n; p.sig_done = sync->sig_done;
        for (int i = 0; i < sync->sig_n; i++) p.sig_peer[i] = sync->sig_peer[i];
        p.wait_n = sync->wait_n; p.wait_ctr = sync->wait_ctr; p.wait_ticket = sync->wait_ticket;
        p.sync_timeout = sync->timeout;
        if ((p.sig_n && !p.sig_done) || (p.wait_n && (!p.wait_ctr || !p.wait_ticket)) || !p.sync_timeout) return cudaErrorNotSupported;
    }
    p.tl = (unsigned long long *)tune.timeline;
    p.wait_mode = tune.wait_mode;
    p.fix_gap = (tune.fix_gap >= 255 || (tune.act_planes == 2 && !tune.fix_gap_set)) ? 255 : std::max(2, tune.fix_gap - (tune.act_planes == 2 ? 8 : 0));
    p.row_bytes = (kq && !bitnet) ? (int)(K / 256) * (type == T_Q4_K ? 144 : type == T_Q8_K ? 292 : 84) : (int)(K / 32) * bb;
    p.nblk = (int)(K / 32);
    p.ncw = (p.nblk + 31) / 32;
    const int target_warps = tune.variant > 0 ? tune.variant : 8;   // consumer warps to aim for
    p.rg = target_warps / p.ncw;
    if (p.rg < 1) p.rg = 1;
    while (p.rg > 1 && p.rg * p.ncw > 16) p.rg--;
    if (p.ncw > 28) return cudaErrorNotSupported;
    p.rg_shift = 0;
    while ((2 << p.rg_shift) <= p.rg) p.rg_shift++;
    p.rg = 1 << p.rg_shift;                        // round down to a power of two
    p.cwarps = p.rg * p.ncw;
    p.unit_rows = 1;
    while ((p.unit_rows * p.row_bytes) % 16) p.unit_rows *= 2;   // <= 8 because row_bytes is even
    if (M % p.unit_rows) return cudaErrorNotSupported;

    // two CTAs (this launch and the next) share one SM's 227 KB when cwarps <= 12
    const int budget = tune.smem_budget > 0 ? tune.smem_budget : (p.ncw > 16 ? 216 : 112) * 1024;   // > 16 column warps: the
                                                                                                     // x planes alone need up to 98 KB
    // Stage sizing, measured on B200 (profiles/r01k_gemv_stage_ab.txt): a launch that waits for its predecessor runs
    // best with stages of rg * 4 rows per 40 KB target (every warp gets whole 4-row steps, fewer barrier rounds on the
    // critical path); independent launches, which overlap pairwise on an SM, with 32 KB stages of any row count.
    const bool dependent = tune.wait_mode != GEMV_WAIT_END;
    const int stage_target = tune.stage_bytes > 0 ? tune.stage_bytes : bitnet ? 16 * 1024 : (dependent ? 40 : 32) * 1024;   // BitNet: 32 KB of the budget hold the table
    int quant = p.unit_rows;                       // stage_rows: multiple of unit_rows (TMA size) and of rg
    while (quant % p.rg) quant += p.unit_rows;
    const bool row_quant = tune.row_quant < 0 ? (dependent || bitnet) : tune.row_quant != 0;
    if (row_quant) {
        int q4 = p.unit_rows;
        while (q4 % (p.rg * (bitnet ? LK_BITNET_UNR : 4))) q4 += p.unit_rows;
        if ((int64_t)q4 * p.row_bytes <= stage_target + stage_target / 4) quant = q4;
    }
    int sr = (stage_target / p.row_bytes / quant) * quant;
    if (sr < quant) sr = quant;
    p.stage_rows = sr;
    const int stage_bytes = p.stage_rows * p.row_bytes;

    int grid = sm_count * (tune.ctas_per_sm > 0 ? tune.ctas_per_sm : 1);
    const int units = (int)(M / p.unit_rows);
    if (grid > units) grid = units;
    const int max_cta_rows = (int)(((int64_t)units + grid - 1) / grid) * p.unit_rows;

    p.units_per_cta = units / grid; p.units_rem = units % grid;
    p.red_stride = p.ncw * 8 + GEMV_RED_PAD;   // one partial per 4 block columns; pad spreads rows over banks
    int red_rows = ((max_cta_rows + p.stage_rows - 1) / p.stage_rows) * p.stage_rows;
    const int red_cap = (24 * 1024 / (p.red_stride * 4) / p.stage_rows) * p.stage_rows;
    if (red_rows > red_cap) red_rows = red_cap;
    if (red_rows < p.stage_rows) red_rows = p.stage_rows;
    if (p.epi_mode == 1) {
        // a pair keeps both segments' partial sums until the end (no flush): rows [0, n) gate, [pair_off, pair_off + n) up;
        // 11008 rows over 148 CTAs at K = 4096 need 24 KB
        p.pair_off = (max_cta_rows + 3) & ~3;
        red_rows = 2 * p.pair_off;
        if ((size_t)red_rows * p.red_stride * 4 > 40 * 1024) return cudaErrorNotSupported;
    }
    p.red_rows = red_rows;
    p.red_off = 256;
    p.out_off = p.red_off + p.red_rows * p.red_stride * 4;            // output staging row for the coalesced peer stores
    p.x_off = (p.out_off + (p.npeers > 0 ? p.red_rows * 4 : 0) + 15) & ~15;
    const int np = tune.act_planes == 2 ? 2 : 3;
    p.stage_off = (p.x_off + p.nblk * gemv_xslot_words(np) * 4 + 127) & ~127;
    if (bitnet) {       // interleaved copies of the 2 KB trit table: sixteen where a CTA has enough rows to pay for filling them
        int copies = tune.bitnet_copies > 0 ? tune.bitnet_copies : ((int64_t)max_cta_rows * p.nblk >= 8192 ? 16 : 1);
        p.lut_shift = 0;
        while ((2 << p.lut_shift) <= copies) p.lut_shift++;
        p.lut_off = p.stage_off;
        p.stage_off += 2048 << p.lut_shift;
    }
    int nst = (budget - p.stage_off) / stage_bytes;
    const int max_st = tune.max_stages > 0 ? tune.max_stages : GEMV_MAX_STAGES;
    if (nst > max_st) nst = max_st;
    if (nst > GEMV_MAX_STAGES) nst = GEMV_MAX_STAGES;
    if (nst < 2) {
        nst = 2;  // very long rows: take what two stages need, up to the 227 KB limit
        if (p.stage_off + 2 * stage_bytes > 227 * 1024) return cudaErrorNotSupported;
    }
    p.nstages = nst;
    const int smem = p.stage_off + nst * stage_bytes;
    // peer stores and the exchange signal / wait live in the fused instantiation only (the plain kernel keeps its size)
    const bool full = p.pro_mode != 0 || p.residual != nullptr || p.nseg > 1;
    const int fused = full ? 1 : (p.npeers > 0 || p.sig_n > 0 || p.wait_n > 0) ? 2 : (!kq && p.wait_mode == GEMV_WAIT_END) ? 3 : 0;     // kernel MODE
    if (kq && fused) return cudaErrorNotSupported;     // (the K-quant instantiations are mode 0 only)

    switch (type * 4 + np) {
        case T_Q4_0 * 4 + 2: return launch_type<T_Q4_0, 2>(p, fused, grid, smem, tune.pdl, s);
        case T_Q4_1 * 4 + 2: return launch_type<T_Q4_1, 2>(p, fused, grid, smem, tune.pdl, s);
        case T_Q8_0 * 4 + 2: return launch_type<T_Q8_0, 2>(p, fused, grid, smem, tune.pdl, s);
        case T_Q4_0 * 4 + 3: return launch_type<T_Q4_0, 3>(p, fused, grid, smem, tune.pdl, s);
        case T_Q4_1 * 4 + 3: return launch_type<T_Q4_1, 3>(p, fused, grid, smem, tune.pdl, s);
        case T_Q8_0 * 4 + 3: return launch_type<T_Q8_0, 3>(p, fused, grid, smem, tune.pdl, s);
        case T_Q4_K * 4 + 2: return launch_plain<T_Q4_K, 2>(p, grid, smem, tune.pdl, s);
        case T_Q4_K * 4 + 3: return launch_plain<T_Q4_K, 3>(p, grid, smem, tune.pdl, s);
        case T_Q8_K * 4 + 2: return launch_plain<T_Q8_K, 2>(p, grid, smem, tune.pdl, s);
        case T_Q8_K * 4 + 3: return launch_plain<T_Q8_K, 3>(p, grid, smem, tune.pdl, s);
        case T_Q2_K * 4 + 2: return launch_plain<T_Q2_K, 2>(p, grid, smem, tune.pdl, s);
        case T_Q2_K * 4 + 3: return launch_plain<T_Q2_K, 3>(p, grid, smem, tune.pdl, s);
        case T_BITNET_1_58 * 4 + 2: return launch_plain<T_BITNET_1_58, 2>(p, grid, smem, tune.pdl, s);
        case T_BITNET_1_58 * 4 + 3: return launch_plain<T_BITNET_1_58, 3>(p, grid, smem, tune.pdl, s);
        default: return cudaErrorNotSupported;
    }
}

// ---- K1c host side: one parameter block per phase, common shared-memory layout --------------
cudaError_t launch_gemv_chain(int type, int n, const GemvChainLink *links, const GemvTuning &tune, int sm_count,
                              cudaStream_t s, unsigned *bar) {
    if (n < 1 || n > CHAIN_MAX || !links || !bar || block_bytes(type) == 0) return cudaErrorNotSupported;
    const int bb = block_bytes(type);
    const int np = tune.act_planes == 2 ? 2 : 3;
    const int grid = sm_count;
    const int slot_target = (tune.stage_bytes > 0 ? tune.stage_bytes : 40 * 1024);
    static_assert(CHAIN_MAX == GEMV_CHAIN_MAX, "lk_kernels.h");
    GemvChainParams cp = {};
    int red_bytes = 0, x_bytes = 0, slot_bytes = 0;
    for (int i = 0; i < n; i++) {
        const GemvChainLink &L = links[i];
        const GemvFusion &f = L.fu;
        GemvParams &p = cp.ph[i];
        const int64_t K = L.K, M = L.M;
        if (K <= 0 || M <= 0 || (K % 64) != 0 || K > 32 * 32 * CHAIN_W || M > (1 << 30)) return cudaErrorNotSupported;
        if (((uintptr_t)L.a & 15) || ((uintptr_t)L.b & 15) || ((uintptr_t)f.pro_w & 15) || ((uintptr_t)f.b_out & 15)) return cudaErrorNotSupported;
        p.a = (const uint8_t *)L.a; p.b = L.b; p.dst = L.dst;
        p.nib = (short)(tune.nibble_order ? 1 : 0);
        p.K = (int)K; p.M = (int)M;
        p.pro_mode = f.pro_mode; p.pro_w = f.pro_w; p.pro_eps = f.pro_eps; p.b_out = f.b_out;
        p.residual = f.residual; p.dst_raw = f.dst_raw;
        if (f.nseg_extra < 0 || f.nseg_extra > 2 || (f.nseg_extra > 0 && f.residual)) return cudaErrorNotSupported;
        for (int j = 0; j < f.nseg_extra; j++) {
            if (!f.a_extra[j] || !f.dst_extra[j] || ((uintptr_t)f.a_extra[j] & 15)) return cudaErrorNotSupported;
            p.a_seg[j] = (const uint8_t *)f.a_extra[j];
            p.dst_seg[j] = f.dst_extra[j];
        }
        p.nseg = 1 + f.nseg_extra;
        if (f.epi_mode == 1) {
            if (f.nseg_extra != 1 || !f.epi_out) return cudaErrorNotSupported;
            p.epi_mode = 1; p.epi_act = f.epi_act; p.epi_out = f.epi_out;
        }
        p.wait_mode = GEMV_WAIT_X;
        p.fix_gap = (tune.fix_gap >= 255 || (np == 2 && !tune.fix_gap_set)) ? 255 : std::max(2, tune.fix_gap - (np == 2 ? 8 : 0));
        p.row_bytes = (int)(K / 32) * bb;
        p.nblk = (int)(K / 32);
        p.ncw = (p.nblk + 31) / 32;
        if (p.ncw > CHAIN_W) return cudaErrorNotSupported;
        p.rg_shift = 0;
        // warps per group to aim for (8: the single launch's shape), never more than the kernel has (gemv_warps = 16 is a
        // setting the sweep tools use: unclamped it gave ng = 0, no warp arriving on empty[] and a producer that spins forever)
        const int gtarget = std::min(tune.variant > 0 ? tune.variant : 8, CHAIN_W);
        while ((2 << p.rg_shift) * p.ncw <= gtarget) p.rg_shift++;
        p.rg = 1 << p.rg_shift;                        // row groups per warp group
        p.ng = CHAIN_W / (p.ncw * p.rg);               // warp groups taking turns at the stages
        if (p.ng < 1) return cudaErrorNotSupported;
        p.cwarps = CHAIN_W;                            // every consumer warp takes part in prologue / epilogue
        p.unit_rows = 1;
        while ((p.unit_rows * p.row_bytes) % 16) p.unit_rows *= 2;
        if (M % p.unit_rows) return cudaErrorNotSupported;
        int quant = p.unit_rows;                       // stage_rows: multiple of unit_rows (TMA size) and of rg
        while (quant % p.rg) quant += p.unit_rows;
        {
            int q4 = p.unit_rows;                      // whole 4-row steps per warp where a slot has room for them
            while (q4 % (p.rg * 4)) q4 += p.unit_rows;
            if ((int64_t)q4 * p.row_bytes <= slot_target + slot_target / 4) quant = q4;
        }
        int sr = (slot_target / p.row_bytes / quant) * quant;
        if (sr < quant) sr = quant;
        p.stage_rows = sr;
        if ((int64_t)sr * p.row_bytes > 72 * 1024) return cudaErrorNotSupported;
        slot_bytes = std::max(slot_bytes, sr * p.row_bytes);
        const int units = (int)(M / p.unit_rows);
        const int max_cta_rows = (int)(((int64_t)units + grid - 1) / grid) * p.unit_rows;
        p.units_per_cta = units / grid; p.units_rem = units % grid;
        p.red_stride = p.ncw * 8 + GEMV_RED_PAD;
        int red_rows = ((max_cta_rows + p.stage_rows - 1) / p.stage_rows) * p.stage_rows;
        const int red_cap = (24 * 1024 / (p.red_stride * 4) / p.stage_rows) * p.stage_rows;
        if (red_rows > red_cap) red_rows = red_cap;
        if (red_rows < p.stage_rows) red_rows = p.stage_rows;
        if (p.epi_mode == 1) {
            p.pair_off = (max_cta_rows + 3) & ~3;
            red_rows = 2 * p.pair_off;
            if ((size_t)red_rows * p.red_stride * 4 > 40 * 1024) return cudaErrorNotSupported;
        }
        p.red_rows = red_rows;
        red_bytes = std::max(red_bytes, red_rows * p.red_stride * 4);
        x_bytes = std::max(x_bytes, p.nblk * gemv_xslot_words(np) * 4);
    }
    const int red_off = 256;
    const int x_off = (red_off + red_bytes + 15) & ~15;
    const int stage_off = (x_off + x_bytes + 127) & ~127;
    slot_bytes = (slot_bytes + 127) & ~127;
    int nslots = (227 * 1024 - stage_off) / slot_bytes;
    if (tune.max_stages > 0 && nslots > tune.max_stages) nslots = tune.max_stages;
    if (nslots > GEMV_MAX_STAGES) nslots = GEMV_MAX_STAGES;
    if (nslots < 2) return cudaErrorNotSupported;
    for (int i = 0; i < n; i++) {
        cp.ph[i].red_off = red_off; cp.ph[i].x_off = x_off; cp.ph[i].stage_off = stage_off; cp.ph[i].nstages = nslots;
    }
    cp.nphase = n;
    cp.slot_bytes = slot_bytes;
    cp.wait_mode = tune.wait_mode == GEMV_WAIT_FULL ? GEMV_WAIT_FULL : GEMV_WAIT_X;
    cp.bar = bar;
    cp.ph[0].tl = (unsigned long long *)tune.timeline;
    const int smem = stage_off + nslots * slot_bytes;

    void (*kern)(const GemvChainParams) = nullptr;
    switch (type * 4 + np) {
        case T_Q4_0 * 4 + 2: kern = gemv_chain_kernel<T_Q4_0, 2>; break;
        case T_Q4_1 * 4 + 2: kern = gemv_chain_kernel<T_Q4_1, 2>; break;
        case T_Q8_0 * 4 + 2: kern = gemv_chain_kernel<T_Q8_0, 2>; break;
        case T_Q4_0 * 4 + 3: kern = gemv_chain_kernel<T_Q4_0, 3>; break;
        case T_Q4_1 * 4 + 3: kern = gemv_chain_kernel<T_Q4_1, 3>; break;
        case T_Q8_0 * 4 + 3: kern = gemv_chain_kernel<T_Q8_0, 3>; break;
        default: return cudaErrorNotSupported;
    }
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    // The grid barrier needs every CTA resident at once.  What can be checked is checked: the device must be able to
    // hold the grid (one CTA of this size per SM on an unpartitioned device); a launch that would not fit is declined
    // and the caller issues the links one by one.  What cannot be checked from here (another stream or process holding
    // SMs for longer than the barrier's bounded spin) ends as a counted time-out that lk_sync / lk_buf_download report.
    int per_sm = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, (CHAIN_W + 1) * 32, (size_t)smem);
    if (e != cudaSuccess) return e;
    if (per_sm < 1 || (int64_t)per_sm * sm_count < grid) return cudaErrorNotSupported;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3((unsigned)((CHAIN_W + 1) * 32));
    cfg.dynamicSmemBytes = (size_t)smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = tune.pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, cp);
}

}  // namespace lk

/*
 * lk_oracle.h — CPU restatement of llama.kotlin's quantized-matmul hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load this library, and only as the checker or as the
 * timed CPU baseline.  The product path (llama.kotlin_b200/) never links,
 * imports or falls back to it.
 *
 * Parity status: PINNED BY THE REFERENCE'S OWN GOLDEN VECTORS (tests/golden,
 * tests/test_oracle_golden.py: every matmul / dot / round-trip / SiLU /
 * RMSNorm vector the reference's nativeTest files hold for this path) and by
 * a second, independent numpy restatement (oracle/np_oracle.py).  The Kotlin
 * reference itself cannot be compiled or run here (no JVM / Konan / Gradle,
 * no network), so there is no oracle/_ref for the Kotlin path; block-level
 * byte layouts are pinned by source citation only.
 *
 * All paths below are relative to
 *   /root/reference/src/nativeMain/kotlin/ai/solace/llamakotlin/
 * and cite the function each routine restates.
 *
 * Build: gcc -O2 -ffp-contract=off -fno-fast-math (see oracle/Makefile) so
 * that every F32 operation is a single IEEE-754 binary32 operation in source
 * order, as Kotlin/Native emits for linuxX64 (no FMA contraction).
 */
#ifndef LK_ORACLE_H
#define LK_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* GGMLType ordinals — core/GGMLTypes.kt:145-153 */
enum {
    LKO_TYPE_F32  = 0,
    LKO_TYPE_F16  = 1,
    LKO_TYPE_Q4_0 = 2,
    LKO_TYPE_Q4_1 = 3,
    LKO_TYPE_Q8_0 = 6,
    LKO_TYPE_Q2_K = 8,
    LKO_TYPE_Q4_K = 10,
    LKO_TYPE_Q8_K = 13,
    LKO_TYPE_BITNET_1_58 = 40   /* no GGMLType.fromValue id in the reference; same id as include/llamakotlin_cuda.h */
};

/* status codes mirror include/llamakotlin_cuda.h so tests can compare 1:1 */
enum {
    LKO_OK          = 0,
    LKO_EINVAL_DIM  = 1,
    LKO_EINVAL_TYPE = 2,
    LKO_EBOUNDS     = 3,
    LKO_ENOTIMPL    = 4
};

/* core/NumericConversions.kt:9-54 (halfToFloat) and :61-124 (floatToHalf,
 * including the half-denormal off-by-one at :83). */
float    lko_half_to_float(uint16_t h);
uint16_t lko_float_to_half(float f);

/* bytes per 32-element block; 0 for non-block types. core/GGMLTypes.kt:99-133 */
int lko_block_bytes(int type);

/* core/GGMLComputeOps.kt:1040-1087 (Q8_0, Q4_0) and :1162-1203 (Q4_1).
 * n must be a multiple of 32; returns LKO_EINVAL_DIM otherwise. */
int lko_quantize(int type, const float *x, int64_t n, uint8_t *out);

/* core/GGMLComputeOps.kt:918-964. n elements, flat order. */
int lko_dequantize(int type, const uint8_t *blocks, int64_t n, float *out);

/* SURVEY §8f N1 — canonical ggml / GGUF nibble order (byte j = elements j and j+16,
 * /root/reference/ggml/src/ggml-quants.c:1515-1555) next to the Kotlin accessors' order (byte j = elements 2j and
 * 2j+1, core/GGMLTypes.kt:647-651).  lko_repack_nibbles permutes the 32 nibbles of every Q4 block between the two
 * (to_ggml = 1: Kotlin -> ggml, 0: ggml -> Kotlin; Q8_0 is a copy); lko_dequantize_ggml restates dequantize_row_q*.
 * Both are checked bit for bit against the reference's own compiled C (oracle/_ref, tests/test_ref_ggml.py). */
int lko_repack_nibbles(int type, const uint8_t *src, int64_t nblocks, uint8_t *dst, int to_ggml);
int lko_dequantize_ggml(int type, const uint8_t *blocks, int64_t n, float *out);

/* core/GGMLComputeOps.kt:43-68 / :120-145 / :70-115: one dot product,
 * row i of `a` (K/32 blocks per row) with column j of b (ne=[N,K]). */
float lko_dot(int a_type, const uint8_t *a, int64_t K, int64_t i,
              const float *b, int64_t N, int64_t j);

/* core/GGMLComputeOps.kt:1435-1565 computeMatMul for a.type in
 * {F32, Q4_0, Q4_1, Q8_0} x F32 -> F32.  a.ne=[K,M], b.ne=[N,K] (N fastest),
 * dst.ne=[N,M] (N fastest).  Single thread, reference loop order. */
int lko_matmul(int a_type, const void *a, int64_t K, int64_t M,
               const float *b, int64_t N, float *dst);

/* Same arithmetic, same per-output k order (bit-identical results), output
 * rows split over `nthreads` pthreads.  Used as the all-cores CPU baseline. */
int lko_matmul_mt(int a_type, const void *a, int64_t K, int64_t M,
                  const float *b, int64_t N, float *dst, int nthreads);

/* Rows [row0,row1) only — lets bench.py time a bounded sample of a big
 * matrix.  dst is the full [N,M] buffer; only the selected rows are written. */
int lko_matmul_rows(int a_type, const void *a, int64_t K, int64_t M,
                    const float *b, int64_t N, float *dst,
                    int64_t row0, int64_t row1, int nthreads);

/* R1: nativeTest core/GGMLComputeOpsTest.kt:502-521 (manualRMSNorm: f64
 * accumulate, rms = (float)sqrt(mean + eps), x / rms) and the weighted form
 * of model/LlamaModel.kt:48-63 (… * weight).  weight may be NULL. */
void lko_rms_norm(const float *x, const float *weight, float eps,
                  int64_t ne0, int64_t nrows, float *dst);

/* R2: nativeTest core/GGMLComputeOpsTest.kt:491-499: x * (1/(1+exp(-x))). */
void lko_silu(const float *x, int64_t n, float *dst);
/* SwiGLU composition, model/LlamaModel.kt:165-175: silu(gate) * up. */
void lko_silu_mul(const float *gate, const float *up, int64_t n, float *dst);

/* R3: model/LlamaAttention.kt:242-288: per row max, exp(x-max), sequential
 * F32 sum, divide. */
void lko_softmax(const float *x, int64_t ne0, int64_t nrows, float *dst);

/* R4: core/GGMLComputeOps.kt:735-759 / :811-835 F32 same-shape add / mul. */
void lko_add(const float *a, const float *b, int64_t n, float *dst);
void lko_mul(const float *a, const float *b, int64_t n, float *dst);

/* N4 attention core — model/LlamaAttention.kt:29-72 (applyRoPE), :166-195 (scale), :104-123 (computeAttention, one
 * query row, heads per the KVCache layout of model/KVCache.kt:19-34). */
void lko_rope(const float *x, int head_dim, int nheads, int ntok, int position, float theta, float *dst);
void lko_scale(const float *x, int64_t n, float factor, float *dst);
void lko_attention_decode(const float *q, const float *kc, const float *vc, int64_t len, int head_dim, int nheads,
                          float *dst);

/* N4 K-quants and BitNet 1.58 (lk_oracle_kq.c): the reference's own Q2_K / Q4_K / Q8_K / BITNET_1_58 formats —
 * core/GGMLComputeOps.kt:152-432, :635-660 (dot products), :966-1036 + :2125-2368 (dequantize), :1365-1426 +
 * :1741-2120 (quantizers), :1481-1513 (computeMatMul branches). */
int   lko_kq_block_bytes(int type);
int   lko_kq_block_elems(int type);
int   lko_kq_quantize(int type, const float *x, int64_t n, uint8_t *out);
int   lko_kq_dequantize(int type, const uint8_t *blocks, int64_t n, float *out);
float lko_kq_dot(int type, const uint8_t *a, int64_t K, int64_t i, const float *b, int64_t N, int64_t j);
int   lko_kq_matmul(int type, const uint8_t *a, int64_t K, int64_t M, const float *b, int64_t N, float *dst);

#ifdef __cplusplus
}
#endif
#endif

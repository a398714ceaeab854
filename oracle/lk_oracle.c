/*
 * lk_oracle.c — CPU restatement of llama.kotlin's quantized-matmul hot path.
 * TEST INFRASTRUCTURE ONLY (see lk_oracle.h for the rules and parity status).
 *
 * Reference paths are relative to
 *   /root/reference/src/nativeMain/kotlin/ai/solace/llamakotlin/
 */
#include "lk_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

#define QK 32 /* QK8_0 = QK4_0 = QK4_1 = 32, core/GGMLTypes.kt:84-86 */

static inline float bits_to_float(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static inline uint32_t float_to_bits(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }

/* ---- core/NumericConversions.kt:9-54 halfToFloat ------------------------ */
float lko_half_to_float(uint16_t h_bits) {
    uint32_t h = h_bits;
    uint32_t sign = h >> 15, exp = (h >> 10) & 0x1F, mant = h & 0x3FF;
    if (exp == 0) {
        if (mant == 0) return bits_to_float(sign << 31);          /* :20-21 */
        int e = 0;                                                /* :23-29 */
        while ((mant & 0x400) == 0) { mant <<= 1; e--; }
        mant &= 0x3FF;
        uint32_t f32exp = (uint32_t)((e + 1) + (127 - 15));       /* :34 */
        return bits_to_float((sign << 31) | (f32exp << 23) | (mant << 13));
    } else if (exp == 0x1F) {
        if (mant == 0) return bits_to_float((sign << 31) | 0x7F800000u);
        return bits_to_float((sign << 31) | 0x7F800000u | (mant << 13) | (1u << 22)); /* :44 */
    }
    return bits_to_float((sign << 31) | ((exp - 15 + 127) << 23) | (mant << 13));     /* :46-52 */
}

/* ---- core/NumericConversions.kt:61-124 floatToHalf ---------------------- *
 * Kotlin Int shifts mask the count to 5 bits and Int arithmetic wraps; both
 * are restated with uint32_t arithmetic and signed compares.  The denormal
 * branch keeps the reference's `shift = 127 - exp` (:83).       */
uint16_t lko_float_to_half(float f_val) {
    uint32_t f32bits = float_to_bits(f_val);
    uint32_t fSign = (f32bits >> 16) & 0x8000u;
    int32_t absF = (int32_t)(f32bits & 0x7FFFFFFFu);

    if (absF > 0x47FFEFFF) {                                      /* :67-73 */
        int nz = (absF & 0x007FFFFF) != 0;
        return (uint16_t)(fSign | 0x7C00u | (nz ? 0x0200u : 0u));
    }
    if (absF < 0x38800000) {                                      /* :75-100 */
        uint32_t fMant = ((uint32_t)absF & 0x007FFFFFu) | 0x00800000u;
        int32_t shift = 127 - (absF >> 23);                       /* :83 (sic) */
        uint32_t hMant = (shift < 24) ? (fMant >> (shift & 31)) : 0u;
        int32_t roundBits = (int32_t)(fMant & ((1u << (shift & 31)) - 1u));
        int32_t half = (int32_t)(1u << ((shift - 1) & 31));
        if (roundBits > half || (roundBits == half && (hMant & 1u) != 0)) {
            uint32_t t = hMant + 1;
            if (t == 0x0400u) return (uint16_t)(fSign | (1u << 10));
            return (uint16_t)(fSign | t);
        }
        return (uint16_t)(fSign | hMant);
    }
    uint32_t hExp = (((uint32_t)absF >> 23) - 112u) << 10;        /* :107 */
    uint32_t hMant = ((uint32_t)absF & 0x007FFFFFu) >> 13;
    if ((absF & 0x00001000) != 0) {                               /* :112-122 */
        if ((absF & 0x00000FFF) != 0 || (hMant & 1u) != 0) {
            hMant++;
            if (hMant == 0x0400u) {
                hMant = 0;
                return (uint16_t)(fSign | (hExp + (1u << 10)) | hMant);
            }
        }
    }
    return (uint16_t)(fSign | hExp | hMant);
}

int lko_block_bytes(int type) {
    switch (type) {                       /* core/GGMLTypes.kt:99-133 */
        case LKO_TYPE_Q4_0: return 18;
        case LKO_TYPE_Q4_1: return 20;
        case LKO_TYPE_Q8_0: return 34;
        default: return 0;
    }
}

static inline void put_u16le(uint8_t *p, uint16_t v) { p[0] = (uint8_t)(v & 0xFF); p[1] = (uint8_t)(v >> 8); }
static inline uint16_t get_u16le(const uint8_t *p) { return (uint16_t)(p[0] | ((uint16_t)p[1] << 8)); }

/* kotlin.math.max / maxOf(Float, Float): NaN if either is NaN. */
static inline float kt_max(float a, float b) { if (isnan(a) || isnan(b)) return NAN; return a > b ? a : b; }
static inline float kt_min(float a, float b) { if (isnan(a) || isnan(b)) return NAN; return a < b ? a : b; }

/* kotlin.math.round(Float) = ties-to-even; Float.toInt() saturates, NaN -> 0;
 * then coerceIn(lo, hi). */
static inline int kt_round_to_int_clamped(float v, int lo, int hi) {
    float r = nearbyintf(v); /* FE_TONEAREST is the process default */
    if (isnan(r)) return (0 < lo) ? lo : (0 > hi ? hi : 0);
    if (r <= (float)lo) return lo;
    if (r >= (float)hi) return hi;
    return (int)r;
}

/* ---- core/GGMLComputeOps.kt:1058-1071 Q8_0 ------------------------------ */
static void quantize_q8_0(const float *x, int64_t n, uint8_t *out) {
    for (int64_t b = 0; b < n / QK; b++) {
        const float *xb = x + b * QK;
        uint8_t *ob = out + b * 34;
        float amax = 0.0f;
        for (int k = 0; k < QK; k++) amax = kt_max(amax, fabsf(xb[k]));
        float scale = (amax == 0.0f) ? 1.0f : amax / 127.0f;
        float invS = 1.0f / scale;
        put_u16le(ob, lko_float_to_half(scale));
        for (int k = 0; k < QK; k++)
            ob[2 + k] = (uint8_t)(int8_t)kt_round_to_int_clamped(xb[k] * invS, -128, 127);
    }
}

/* ---- core/GGMLComputeOps.kt:1072-1087 Q4_0 (Kotlin nibble order, F2/F3) - */
static void quantize_q4_0(const float *x, int64_t n, uint8_t *out) {
    for (int64_t b = 0; b < n / QK; b++) {
        const float *xb = x + b * QK;
        uint8_t *ob = out + b * 18;
        float amax = 0.0f;
        for (int k = 0; k < QK; k++) amax = kt_max(amax, fabsf(xb[k]));
        float scale = (amax == 0.0f) ? 1.0f : amax / 8.0f;
        float invS = (scale == 0.0f) ? 0.0f : 1.0f / scale;
        put_u16le(ob, lko_float_to_half(scale));
        for (int j = 0; j < QK / 2; j++) {
            int q1 = kt_round_to_int_clamped(xb[2 * j] * invS + 8.0f, 0, 15);
            int q2 = kt_round_to_int_clamped(xb[2 * j + 1] * invS + 8.0f, 0, 15);
            ob[2 + j] = (uint8_t)((q1 & 0x0F) | ((q2 & 0x0F) << 4));
        }
    }
}

/* ---- core/GGMLComputeOps.kt:1162-1203 Q4_1 ------------------------------ */
static void quantize_q4_1(const float *x, int64_t n, uint8_t *out) {
    for (int64_t b = 0; b < n / QK; b++) {
        const float *xb = x + b * QK;
        uint8_t *ob = out + b * 20;
        float fmin = xb[0], fmax = xb[0];
        for (int k = 1; k < QK; k++) { fmin = kt_min(fmin, xb[k]); fmax = kt_max(fmax, xb[k]); }
        float d = (fmax - fmin) / 15.0f;
        if (d == 0.0f) d = 1.0f;
        float m = fmin;
        float invD = 1.0f / d;
        put_u16le(ob, lko_float_to_half(d));
        put_u16le(ob + 2, lko_float_to_half(m));
        for (int j = 0; j < QK / 2; j++) {
            int q1 = kt_round_to_int_clamped((xb[2 * j] - m) * invD, 0, 15);
            int q2 = kt_round_to_int_clamped((xb[2 * j + 1] - m) * invD, 0, 15);
            ob[4 + j] = (uint8_t)((q1 & 0x0F) | ((q2 & 0x0F) << 4));
        }
    }
}

int lko_quantize(int type, const float *x, int64_t n, uint8_t *out) {
    if (n % QK != 0) return LKO_EINVAL_DIM;   /* require(numElements % QK == 0) */
    switch (type) {
        case LKO_TYPE_Q8_0: quantize_q8_0(x, n, out); return LKO_OK;
        case LKO_TYPE_Q4_0: quantize_q4_0(x, n, out); return LKO_OK;
        case LKO_TYPE_Q4_1: quantize_q4_1(x, n, out); return LKO_OK;
        default: return LKO_ENOTIMPL;
    }
}

/* ---- block field accessors: core/GGMLTypes.kt:543-732 -------------------- *
 * Called once PER ELEMENT, exactly like the reference (scale re-decoded for
 * every one of the 32 items of a block, §3.1 of SURVEY.md).                  */
static inline float q8_0_scale(const uint8_t *a, int64_t blk) { return lko_half_to_float(get_u16le(a + blk * 34)); }
static inline int8_t q8_0_weight(const uint8_t *a, int64_t blk, int item) { return (int8_t)a[blk * 34 + 2 + item]; }
static inline float q4_0_scale(const uint8_t *a, int64_t blk) { return lko_half_to_float(get_u16le(a + blk * 18)); }
static inline int q4_0_nibble(const uint8_t *a, int64_t blk, int item) {
    uint8_t p = a[blk * 18 + 2 + item / 2];
    return (item % 2 == 0) ? (p & 0x0F) : ((p >> 4) & 0x0F);     /* :647-651 */
}
static inline float q4_1_scale(const uint8_t *a, int64_t blk) { return lko_half_to_float(get_u16le(a + blk * 20)); }
static inline float q4_1_min(const uint8_t *a, int64_t blk) { return lko_half_to_float(get_u16le(a + blk * 20 + 2)); }
static inline int q4_1_nibble(const uint8_t *a, int64_t blk, int item) {
    uint8_t p = a[blk * 20 + 4 + item / 2];
    return (item % 2 == 0) ? (p & 0x0F) : ((p >> 4) & 0x0F);     /* :726-730 */
}

/* ---- core/GGMLComputeOps.kt:918-964 dequantizeTensor --------------------- */
int lko_dequantize(int type, const uint8_t *a, int64_t n, float *out) {
    if (n % QK != 0) return LKO_EINVAL_DIM;
    int64_t nb = n / QK, f = 0;
    switch (type) {
        case LKO_TYPE_Q8_0:
            for (int64_t b = 0; b < nb; b++) {
                float s = q8_0_scale(a, b);
                for (int k = 0; k < QK; k++) out[f++] = s * (float)q8_0_weight(a, b, k);          /* :936 */
            }
            return LKO_OK;
        case LKO_TYPE_Q4_0:
            for (int64_t b = 0; b < nb; b++) {
                float s = q4_0_scale(a, b);
                for (int k = 0; k < QK; k++) out[f++] = s * ((float)q4_0_nibble(a, b, k) - 8.0f); /* :945 */
            }
            return LKO_OK;
        case LKO_TYPE_Q4_1:
            for (int64_t b = 0; b < nb; b++) {
                float d = q4_1_scale(a, b), m = q4_1_min(a, b);
                for (int k = 0; k < QK; k++) out[f++] = d * (float)q4_1_nibble(a, b, k) + m;      /* :958 */
            }
            return LKO_OK;
        default:
            return LKO_ENOTIMPL;
    }
}

/* ---- canonical ggml nibble order -------------------------- *
 * The Kotlin accessors put element e of a Q4 block in byte e/2, nibble e&1 (core/GGMLTypes.kt:647-651);
 * ggml and every GGUF file put it in byte e%16, nibble e/16 (/root/reference/ggml/src/ggml-quants.c:1524-1530,
 * :1545-1551).  Q8_0 blocks are identical in both.  to_ggml = 0: ggml order in, Kotlin order out; 1: the inverse.
 * Scales / mins are copied; the map is a bijection on the 32 nibbles, so repack(repack(x, 1), 0) == x.      */
int lko_repack_nibbles(int type, const uint8_t *src, int64_t nblocks, uint8_t *dst, int to_ggml) {
    const int bb = lko_block_bytes(type);
    if (bb == 0) return LKO_ENOTIMPL;
    if (type == LKO_TYPE_Q8_0) { memcpy(dst, src, (size_t)nblocks * bb); return LKO_OK; }
    const int qoff = type == LKO_TYPE_Q4_0 ? 2 : 4;
    for (int64_t b = 0; b < nblocks; b++) {
        const uint8_t *s = src + b * bb;
        uint8_t *d = dst + b * bb, nib[32];
        memcpy(d, s, qoff);
        for (int e = 0; e < QK; e++)       /* element e of the source block */
            nib[e] = to_ggml ? (uint8_t)((s[qoff + e / 2] >> (4 * (e & 1))) & 15) : (uint8_t)((s[qoff + e % 16] >> (4 * (e / 16))) & 15);
        memset(d + qoff, 0, 16);
        for (int e = 0; e < QK; e++) {
            if (to_ggml) d[qoff + e % 16] |= (uint8_t)(nib[e] << (4 * (e / 16)));
            else d[qoff + e / 2] |= (uint8_t)(nib[e] << (4 * (e & 1)));
        }
    }
    return LKO_OK;
}

/* Direct restatement of dequantize_row_q4_0 / q4_1 / q8_0 (/root/reference/ggml/src/ggml-quants.c:1515-1555,
 * :1609-1625): y[j] = (lo - 8) * d, y[j + 16] = (hi - 8) * d; lo * d + m, hi * d + m; q * d.               */
int lko_dequantize_ggml(int type, const uint8_t *a, int64_t n, float *out) {
    if (n % QK != 0) return LKO_EINVAL_DIM;
    const int bb = lko_block_bytes(type);
    if (bb == 0) return LKO_ENOTIMPL;
    if (type == LKO_TYPE_Q8_0) return lko_dequantize(type, a, n, out);
    const int qoff = type == LKO_TYPE_Q4_0 ? 2 : 4;
    for (int64_t b = 0; b < n / QK; b++) {
        const uint8_t *s = a + b * bb;
        const float d = lko_half_to_float(get_u16le(s));
        const float m = type == LKO_TYPE_Q4_1 ? lko_half_to_float(get_u16le(s + 2)) : 0.0f;
        for (int j = 0; j < 16; j++) {
            const int lo = s[qoff + j] & 0x0F, hi = s[qoff + j] >> 4;
            if (type == LKO_TYPE_Q4_0) {
                out[b * QK + j] = (float)(lo - 8) * d;
                out[b * QK + j + 16] = (float)(hi - 8) * d;
            } else {
                out[b * QK + j] = (float)lo * d + m;
                out[b * QK + j + 16] = (float)hi * d + m;
            }
        }
    }
    return LKO_OK;
}

/* ---- dot products: core/GGMLComputeOps.kt:43-68, :120-145, :70-115 ------- *
 * sum += (dequantized) * f32, F32, strictly sequential in k.                 */
static float dot_q8_0(const uint8_t *a, int64_t K, int64_t i, const float *b, int64_t N, int64_t j) {
    float sum = 0.0f;
    for (int64_t k = 0; k < K; k++) {
        int64_t flat = i * K + k, blk = flat / QK; int item = (int)(flat % QK);
        float dq = q8_0_scale(a, blk) * (float)q8_0_weight(a, blk, item);
        sum += dq * b[j + k * N];
    }
    return sum;
}
static float dot_q4_0(const uint8_t *a, int64_t K, int64_t i, const float *b, int64_t N, int64_t j) {
    float sum = 0.0f;
    for (int64_t k = 0; k < K; k++) {
        int64_t flat = i * K + k, blk = flat / QK; int item = (int)(flat % QK);
        float dq = q4_0_scale(a, blk) * ((float)q4_0_nibble(a, blk, item) - 8.0f);
        sum += dq * b[j + k * N];
    }
    return sum;
}
static float dot_q4_1(const uint8_t *a, int64_t K, int64_t i, const float *b, int64_t N, int64_t j) {
    float sum = 0.0f;
    for (int64_t k = 0; k < K; k++) {
        int64_t flat = i * K + k, blk = flat / QK; int item = (int)(flat % QK);
        float dq = q4_1_scale(a, blk) * (float)q4_1_nibble(a, blk, item) + q4_1_min(a, blk);
        sum += dq * b[j + k * N];
    }
    return sum;
}
static float dot_f32(const float *a, int64_t K, int64_t i, const float *b, int64_t N, int64_t j) {
    float sum = 0.0f;                                  /* :1532-1543 */
    for (int64_t l = 0; l < K; l++) sum += a[l + i * K] * b[j + l * N];
    return sum;
}

float lko_dot(int a_type, const uint8_t *a, int64_t K, int64_t i, const float *b, int64_t N, int64_t j) {
    switch (a_type) {
        case LKO_TYPE_Q8_0: return dot_q8_0(a, K, i, b, N, j);
        case LKO_TYPE_Q4_0: return dot_q4_0(a, K, i, b, N, j);
        case LKO_TYPE_Q4_1: return dot_q4_1(a, K, i, b, N, j);
        case LKO_TYPE_F32:  return dot_f32((const float *)a, K, i, b, N, j);
        default: return NAN;
    }
}

/* ---- core/GGMLComputeOps.kt:1435-1565 computeMatMul ----------------------- */
static int matmul_check(int a_type, int64_t K, int64_t M, int64_t N) {
    if (K < 0 || M < 0 || N < 0) return LKO_EINVAL_DIM;
    switch (a_type) {
        case LKO_TYPE_F32: return LKO_OK;
        case LKO_TYPE_Q4_0: case LKO_TYPE_Q4_1: case LKO_TYPE_Q8_0:
            /* blocks are addressed as flatIndex/32 over a contiguous tensor
             * (F9): M*K must be whole blocks or the accessor bounds `require`
             * trips on the last partial block. */
            return ((M * K) % QK == 0) ? LKO_OK : LKO_EINVAL_DIM;
        default: return LKO_ENOTIMPL;
    }
}

static void matmul_range(int a_type, const void *a, int64_t K, const float *b, int64_t N,
                         float *dst, int64_t r0, int64_t r1) {
    for (int64_t i = r0; i < r1; i++)          /* HOT LOOP 1 (:1520) */
        for (int64_t j = 0; j < N; j++)        /* HOT LOOP 2 (:1521) */
            dst[j + i * N] = lko_dot(a_type, (const uint8_t *)a, K, i, b, N, j);
}

int lko_matmul(int a_type, const void *a, int64_t K, int64_t M, const float *b, int64_t N, float *dst) {
    int rc = matmul_check(a_type, K, M, N);
    if (rc != LKO_OK) return rc;
    matmul_range(a_type, a, K, b, N, dst, 0, M);
    return LKO_OK;
}

typedef struct {
    int a_type; const void *a; int64_t K; const float *b; int64_t N; float *dst; int64_t r0, r1;
} mm_job;

static void *mm_worker(void *p) {
    mm_job *j = (mm_job *)p;
    matmul_range(j->a_type, j->a, j->K, j->b, j->N, j->dst, j->r0, j->r1);
    return NULL;
}

int lko_matmul_rows(int a_type, const void *a, int64_t K, int64_t M, const float *b, int64_t N,
                    float *dst, int64_t row0, int64_t row1, int nthreads) {
    int rc = matmul_check(a_type, K, M, N);
    if (rc != LKO_OK) return rc;
    if (row0 < 0 || row1 > M || row0 > row1) return LKO_EBOUNDS;
    int64_t rows = row1 - row0;
    if (nthreads < 1) nthreads = 1;
    if (nthreads > rows) nthreads = rows > 0 ? (int)rows : 1;
    if (nthreads == 1) { matmul_range(a_type, a, K, b, N, dst, row0, row1); return LKO_OK; }
    pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * (size_t)nthreads);
    mm_job *jobs = (mm_job *)malloc(sizeof(mm_job) * (size_t)nthreads);
    for (int t = 0; t < nthreads; t++) {
        jobs[t] = (mm_job){a_type, a, K, b, N, dst, row0 + rows * t / nthreads, row0 + rows * (t + 1) / nthreads};
        pthread_create(&th[t], NULL, mm_worker, &jobs[t]);
    }
    for (int t = 0; t < nthreads; t++) pthread_join(th[t], NULL);
    free(th); free(jobs);
    return LKO_OK;
}

int lko_matmul_mt(int a_type, const void *a, int64_t K, int64_t M, const float *b, int64_t N,
                  float *dst, int nthreads) {
    return lko_matmul_rows(a_type, a, K, M, b, N, dst, 0, M, nthreads);
}

/* ---- R1 RMSNorm: nativeTest core/GGMLComputeOpsTest.kt:502-521 ----------- */
void lko_rms_norm(const float *x, const float *weight, float eps, int64_t ne0, int64_t nrows, float *dst) {
    for (int64_t r = 0; r < nrows; r++) {
        const float *xr = x + r * ne0; float *dr = dst + r * ne0;
        if (ne0 == 0) continue;
        double sumSq = 0.0;
        for (int64_t i = 0; i < ne0; i++) { double v = (double)xr[i]; sumSq += v * v; }
        double meanSq = sumSq / (double)ne0;
        float rms = (float)sqrt(meanSq + (double)eps);
        if (rms == 0.0f || isnan(rms)) { for (int64_t i = 0; i < ne0; i++) dr[i] = 0.0f; continue; }
        for (int64_t i = 0; i < ne0; i++) {
            float v = xr[i] / rms;
            dr[i] = weight ? v * weight[i] : v;      /* model/LlamaModel.kt:59-62 */
        }
    }
}

/* ---- R2 SiLU: nativeTest core/GGMLComputeOpsTest.kt:491-499 -------------- */
static inline float silu1(float x) { return x * (1.0f / (1.0f + expf(-x))); }
void lko_silu(const float *x, int64_t n, float *dst) { for (int64_t i = 0; i < n; i++) dst[i] = silu1(x[i]); }
void lko_silu_mul(const float *g, const float *u, int64_t n, float *dst) {
    for (int64_t i = 0; i < n; i++) dst[i] = silu1(g[i]) * u[i];
}

/* ---- R3 softmax: model/LlamaAttention.kt:242-288 -------------------------- */
void lko_softmax(const float *x, int64_t ne0, int64_t nrows, float *dst) {
    for (int64_t r = 0; r < nrows; r++) {
        const float *xr = x + r * ne0; float *dr = dst + r * ne0;
        float maxVal = -INFINITY;
        for (int64_t d = 0; d < ne0; d++) if (xr[d] > maxVal) maxVal = xr[d];
        float sum = 0.0f;
        for (int64_t d = 0; d < ne0; d++) { dr[d] = expf(xr[d] - maxVal); sum += dr[d]; }
        for (int64_t d = 0; d < ne0; d++) dr[d] = dr[d] / sum;
    }
}

/* ---- R4 add / mul: core/GGMLComputeOps.kt:735-759, :811-835 --------------- */
void lko_add(const float *a, const float *b, int64_t n, float *dst) { for (int64_t i = 0; i < n; i++) dst[i] = a[i] + b[i]; }
void lko_mul(const float *a, const float *b, int64_t n, float *dst) { for (int64_t i = 0; i < n; i++) dst[i] = a[i] * b[i]; }

/* ---- attention core -------------------------------------------------------------------------- */

/* LlamaAttention.applyRoPE, model/LlamaAttention.kt:29-72.  x: ntok rows of nheads*head_dim floats (ne[0] = hidden,
 * ne[1] = sequence); pairs (i, i+1) inside each head; angle = (position + s) / theta^((i/2)/(headDim/2)) in F64,
 * cos / sin rounded to F32, then x*cos - y*sin and x*sin + y*cos in F32 (two products rounded, then one add).  */
void lko_rope(const float *x, int head_dim, int nheads, int ntok, int position, float theta, float *dst) {
    const int64_t row = (int64_t)head_dim * nheads;
    for (int s = 0; s < ntok; s++)
        for (int h = 0; h < nheads; h++)
            for (int i = 0; i < head_dim; i += 2) {
                const int pos = position + s;                                              /* :52 */
                const double exponent = ((double)i / 2.0) / ((double)head_dim / 2.0);      /* :53 */
                const double inv_freq = 1.0 / pow((double)theta, exponent);                /* :54 */
                const double angle = (double)pos * inv_freq;                               /* :55 */
                const float c = (float)cos(angle), sn = (float)sin(angle);                 /* :56-57 */
                const float xv = x[s * row + h * head_dim + i], yv = x[s * row + h * head_dim + i + 1];
                dst[s * row + h * head_dim + i] = xv * c - yv * sn;                        /* :62 */
                dst[s * row + h * head_dim + i + 1] = xv * sn + yv * c;                    /* :63 */
            }
}

/* LlamaAttention.scale, model/LlamaAttention.kt:166-195: value * factor. */
void lko_scale(const float *x, int64_t n, float factor, float *dst) {
    for (int64_t i = 0; i < n; i++) dst[i] = x[i] * factor;
}

/* LlamaAttention.computeAttention for one query row, model/LlamaAttention.kt:104-123, per head of the KVCache layout
 * [headDim, numHeads, seq] (model/KVCache.kt:19-34): scores = q.K^T as an F32 matmul (sequential k, the F32 x F32
 * loop of core/GGMLComputeOps.kt:1532-1543), * 1/sqrt(headDim) (:108-109), softmax (:242-288, lko_softmax),
 * weights.V with the same sequential F32 matmul.  len = rows of the cache in use.                            */
void lko_attention_decode(const float *q, const float *kc, const float *vc, int64_t len, int head_dim, int nheads,
                          float *dst) {
    const int64_t row = (int64_t)head_dim * nheads;
    const float scale = 1.0f / sqrtf((float)head_dim);
    float *sc = (float *)malloc((size_t)len * 2 * sizeof(float)), *w = sc + len;
    for (int h = 0; h < nheads; h++) {
        for (int64_t s = 0; s < len; s++) {
            float sum = 0.0f;
            for (int d = 0; d < head_dim; d++) sum += q[h * head_dim + d] * kc[s * row + h * head_dim + d];
            sc[s] = sum * scale;
        }
        lko_softmax(sc, len, 1, w);
        for (int d = 0; d < head_dim; d++) {
            float sum = 0.0f;
            for (int64_t s = 0; s < len; s++) sum += w[s] * vc[s * row + h * head_dim + d];
            dst[h * head_dim + d] = sum;
        }
    }
    free(sc);
}

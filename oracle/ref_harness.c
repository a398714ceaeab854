/*
 * ref_harness.c — thin C entry points over the reference tree's OWN ggml C sources.
 *
 * TEST INFRASTRUCTURE ONLY (see lk_oracle.h).  This file is compiled TOGETHER WITH the unmodified
 * sources under /root/reference/ggml/src (ggml.c, ggml-quants.c, ggml-alloc.c, ggml-backend.c,
 * ggml-aarch64.c), where they lie, into oracle/_ref/libggml_ref.so by oracle/Makefile (target `ref`).
 * No reference source is copied into this repository; only this harness (ours) is.
 *
 * What it pins:
 *   - ggml_fp16_to_fp32 / ggml_fp32_to_fp16  (the conversions core/NumericConversions.kt ports)
 *   - dequantize_row_q4_0 / q4_1 / q8_0      (ggml/src/ggml-quants.c:1515-1555, :1609) — the CANONICAL ggml
 *     nibble order (byte j = elements j and j+16), i.e. what real GGUF files hold
 *   - quantize_row_q8_0_ref / q4_0_ref / q4_1_ref (:661-740, :841-862)
 *   - ggml_rms_norm / ggml_silu / ggml_soft_max / ggml_mul_mat / ggml_rope through ggml_graph_compute: the
 *     C implementation the Kotlin ops are a port of (R1-R3 have no executable Kotlin code)
 * The ggml mul_mat quantizes the activations to Q8_0 first (vec_dot_q4_0_q8_0), so it agrees with the
 * Kotlin F32 dot only to ~1e-2; it is used as the timed all-cores CPU reference, not as the parity bar.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "ggml.h"

/* from ggml-quants.h (not a public header; declared here to avoid including internal headers) */
void dequantize_row_q4_0(const void *x, float *y, int64_t k);
void dequantize_row_q4_1(const void *x, float *y, int64_t k);
void dequantize_row_q8_0(const void *x, float *y, int64_t k);
void quantize_row_q4_0_ref(const float *x, void *y, int64_t k);
void quantize_row_q4_1_ref(const float *x, void *y, int64_t k);
void quantize_row_q8_0_ref(const float *x, void *y, int64_t k);

#define EXPORT __attribute__((visibility("default")))

static int g_inited = 0;
EXPORT void ref_init(void) {
    if (g_inited) return;
    struct ggml_init_params ip = {1 << 20, NULL, 0};
    struct ggml_context *c = ggml_init(ip); /* fills the f16 tables on first call */
    ggml_free(c);
    g_inited = 1;
}

EXPORT float ref_fp16_to_fp32(uint16_t h) { ref_init(); return ggml_fp16_to_fp32((ggml_fp16_t)h); }
EXPORT uint16_t ref_fp32_to_fp16(float f) { ref_init(); return (uint16_t)ggml_fp32_to_fp16(f); }

/* type ids are the GGMLType / ggml_type ordinals: Q4_0 = 2, Q4_1 = 3, Q8_0 = 8 in ggml (6 in the Kotlin enum) */
static enum ggml_type map_type(int t) {
    switch (t) {
        case 0: return GGML_TYPE_F32;
        case 2: return GGML_TYPE_Q4_0;
        case 3: return GGML_TYPE_Q4_1;
        case 6: return GGML_TYPE_Q8_0; /* Kotlin ordinal */
        default: return GGML_TYPE_COUNT;
    }
}

EXPORT int ref_dequantize(int type, const void *src, int64_t n, float *dst) {
    ref_init();
    switch (type) {
        case 2: dequantize_row_q4_0(src, dst, n); return 0;
        case 3: dequantize_row_q4_1(src, dst, n); return 0;
        case 6: dequantize_row_q8_0(src, dst, n); return 0;
    }
    return 2;
}

EXPORT int ref_quantize(int type, const float *x, int64_t n, void *out) {
    ref_init();
    switch (type) {
        case 2: quantize_row_q4_0_ref(x, out, n); return 0;
        case 3: quantize_row_q4_1_ref(x, out, n); return 0;
        case 6: quantize_row_q8_0_ref(x, out, n); return 0;
    }
    return 2;
}

/* op: 0 rms_norm(eps), 1 silu, 2 soft_max, 3 rope (NORM mode 0: pairs (2i, 2i+1); p0 = position, p1 = freq_base);
 * x: nrows rows of ne0 floats.  For rope, x is [ne0 = head_dim, nrows = heads] at one position. */
EXPORT int ref_unary(int op, const float *x, int64_t ne0, int64_t nrows, float p0, float p1, float *out, int nthreads) {
    ref_init();
    const size_t bytes = (size_t)ne0 * nrows * 4;
    struct ggml_init_params ip = {bytes * 4 + (64u << 20), NULL, 0};
    struct ggml_context *c = ggml_init(ip);
    if (!c) return 1;
    struct ggml_tensor *t, *r = NULL;
    if (op == 3) {
        t = ggml_new_tensor_3d(c, GGML_TYPE_F32, ne0, nrows, 1);
        memcpy(t->data, x, bytes);
        struct ggml_tensor *pos = ggml_new_tensor_1d(c, GGML_TYPE_I32, 1);
        ((int32_t *)pos->data)[0] = (int32_t)p0;
        r = ggml_rope_ext(c, t, pos, NULL, (int)ne0, 0, 0, p1, 1.0f, 0.0f, 1.0f, 0.0f, 0.0f);
    } else {
        t = ggml_new_tensor_2d(c, GGML_TYPE_F32, ne0, nrows);
        memcpy(t->data, x, bytes);
        if (op == 0) r = ggml_rms_norm(c, t, p0);
        else if (op == 1) r = ggml_silu(c, t);
        else if (op == 2) r = ggml_soft_max(c, t);
    }
    if (!r) { ggml_free(c); return 2; }
    struct ggml_cgraph *g = ggml_new_graph(c);
    ggml_build_forward_expand(g, r);
    const int st = (int)ggml_graph_compute_with_ctx(c, g, nthreads > 0 ? nthreads : 1);
    memcpy(out, r->data, bytes);
    ggml_free(c);
    return st;
}

/* dst[M x N] (M fastest, ggml layout) = a[K x M rows] . b[K x N] (K fastest).  a holds ggml-order blocks.
 * reps > 1 recomputes the graph (timing); returns seconds per compute in *sec if sec != NULL. */
#include <time.h>
EXPORT int ref_mul_mat(int a_type, const void *a, int64_t K, int64_t M, const float *b, int64_t N, float *dst,
                       int nthreads, int reps, double *sec) {
    ref_init();
    const enum ggml_type gt = map_type(a_type);
    if (gt == GGML_TYPE_COUNT) return 2;
    const size_t abytes = ggml_row_size(gt, K) * (size_t)M;
    const size_t need = abytes + (size_t)K * N * 4 + (size_t)M * N * 4 + (size_t)K * N * 2 + (256u << 20);
    struct ggml_init_params ip = {need, NULL, 0};
    struct ggml_context *c = ggml_init(ip);
    if (!c) return 1;
    struct ggml_tensor *ta = ggml_new_tensor_2d(c, gt, K, M);
    struct ggml_tensor *tb = ggml_new_tensor_2d(c, GGML_TYPE_F32, K, N);
    memcpy(ta->data, a, abytes);
    memcpy(tb->data, b, (size_t)K * N * 4);
    struct ggml_tensor *r = ggml_mul_mat(c, ta, tb);
    struct ggml_cgraph *g = ggml_new_graph(c);
    ggml_build_forward_expand(g, r);
    int st = 0;
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (int i = 0; i < (reps > 0 ? reps : 1) && st == 0; i++) st = (int)ggml_graph_compute_with_ctx(c, g, nthreads > 0 ? nthreads : 1);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    if (sec) *sec = ((t1.tv_sec - t0.tv_sec) + 1e-9 * (t1.tv_nsec - t0.tv_nsec)) / (reps > 0 ? reps : 1);
    memcpy(dst, r->data, (size_t)M * N * 4);
    ggml_free(c);
    return st;
}

/* A GGUF file written by the reference tree's own writer (gguf_* in ggml.c): KV pairs general.architecture = "llama",
 * general.name, general.alignment (u32), llama.block_count / embedding_length / rope.freq_base, a string array, plus
 * `nt` 2-D tensors (type ids as in map_type, payload bytes given).  Fixture generator for the GGUF reader tests. */
EXPORT int ref_gguf_write(const char *path, int alignment, int nt, const char **names, const int *types,
                          const int64_t *ne0, const int64_t *ne1, const void **payload) {
    ref_init();
    size_t total = 1 << 20;
    for (int i = 0; i < nt; i++) total += ggml_row_size(map_type(types[i]), ne0[i]) * (size_t)ne1[i] + 4096;
    struct ggml_init_params ip = {total, NULL, 0};
    struct ggml_context *c = ggml_init(ip);
    if (!c) return 1;
    struct gguf_context *g = gguf_init_empty();
    gguf_set_val_str(g, "general.architecture", "llama");
    gguf_set_val_str(g, "general.name", "lk-b200-fixture");
    gguf_set_val_u32(g, "general.alignment", (uint32_t)alignment);
    gguf_set_val_u32(g, "llama.block_count", 1);
    gguf_set_val_u32(g, "llama.embedding_length", (uint32_t)ne0[0]);
    gguf_set_val_f32(g, "llama.rope.freq_base", 10000.0f);
    gguf_set_val_bool(g, "lk.flag", 1);
    const char *toks[3] = {"<s>", "</s>", "hello"};
    gguf_set_arr_str(g, "tokenizer.ggml.tokens", toks, 3);
    for (int i = 0; i < nt; i++) {
        struct ggml_tensor *t = ggml_new_tensor_2d(c, map_type(types[i]), ne0[i], ne1[i]);
        ggml_set_name(t, names[i]);
        memcpy(t->data, payload[i], ggml_nbytes(t));
        gguf_add_tensor(g, t);
    }
    gguf_write_to_file(g, path, 0);
    gguf_free(g);
    ggml_free(c);
    return 0;
}

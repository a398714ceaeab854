// lk_kernels.h — host-callable launchers for the sm_100a kernels.  Internal to
// the shared library; the public surface is include/llamakotlin_cuda.h.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace lk {

// Tunables for the K1 streaming GEMV (set through lk_set_option; defaults are
// the product settings chosen from measurements).
struct GemvTuning {
    int variant = 0;        // 0 = auto
    int stage_bytes = 0;    // 0 = auto: target bytes per TMA stage
    int max_stages = 0;     // 0 = auto
    int smem_budget = 0;    // 0 = auto: bytes of dynamic smem per CTA
    int pdl = 1;            // programmatic dependent launch attribute
    int ctas_per_sm = 1;
    int bitnet_copies = 0;  // BITNET_1_58: interleaved copies of the trit table in shared memory (1..16, power of two); 0 = by shape
    int act_planes = 3;     // activation fixed-point planes: 3 = 24-bit (default: as accurate as the F32 reference),
                            // 2 = 16-bit (opt-in: +8 % throughput, error 5e-5 of max instead of 2e-6)
    int row_quant = -1;     // stage rows a multiple of 4 per warp: -1 auto (dependent launches only), 0 never, 1 always
    int nibble_order = 0;   // Q4 blocks: 0 = the Kotlin accessors' order (byte j = elements 2j, 2j+1, core/GGMLTypes.kt:647-651),
                            // 1 = canonical ggml / GGUF order (byte j = elements j, j+16, ggml/src/ggml-quants.c:1524-1530)
    int fix_gap = 20;       // range fix-up: blocks holding a non-zero |x| below 2^-fix_gap of the block maximum are recomputed from the
                            // F32 activations (lk_gemv.cu, GemvSide); >= 255 switches the detection off
    bool fix_gap_set = false;   // the option was set by the caller: the opt-in 16-bit mode (act_planes = 2) runs without the fix-up
                                // otherwise (its contract is the 2^-15 absolute bound; at gap 12 most launches would take the cold path)
    int wait_mode = 1;      // 0 wait before the weight prefetch, 1 before reading x, 2 at the end (independent launch)
    void *timeline = nullptr;   // tools/gemv_timeline.cu only: per-CTA %globaltimer stamps
};

// K0: block dequantisation, bit-exact with dequantizeTensor (core/GGMLComputeOps.kt:918-964).
// nib: nibble order of Q4 blocks (GemvTuning::nibble_order)
cudaError_t launch_dequantize(int type, const void *src, int64_t n, float *dst, cudaStream_t s, int nib = 0);

// A11: quantizeTensor on the device, bit-exact with core/GGMLComputeOps.kt:1058-1087 / :1162-1203 (incl. floatToHalf's F11)
cudaError_t launch_quantize(int type, const float *x, int64_t n, void *out, cudaStream_t s, int nib = 0);
// A9: gather a strided F32 tensor into contiguous order (scatter = 0) or scatter a contiguous one through strides (1)
cudaError_t launch_strided_copy(const void *src, void *dst, const int64_t ne[4], const uint64_t nb[4], int scatter, cudaStream_t s);

// K1: decode GEMV, N == 1, K % 64 == 0.  Returns cudaErrorNotSupported when the
// shape cannot take the streaming path (caller then uses the generic kernel).
// dst_peers/npeers: optional extra destinations (peer-mapped buffers) for the
// fused all-gather; row_offset shifts the output index (sharded weights).
struct GemvPeers {
    float *ptr[8];
    int n = 0;
};
// Neighbours of the matmul in the layer graph folded into the launch; all optional.
struct GemvFusion {
    int pro_mode = 0;               // 1: the activations are RMSNorm(b, pro_w, pro_eps), computed by every CTA from b
    const float *pro_w = nullptr;
    float pro_eps = 0.f;
    float *b_out = nullptr;         // also store the activations actually used (the unfused RMSNorm's dst)
    const float *residual = nullptr;   // dst = residual + a x b  (the following ADD)
    float *dst_raw = nullptr;       // also store the bare product (the unfused matmul's dst)
    // stacked segments: up to two more weights of the same type and shape, multiplied with the same activations by
    // the same launch (Q/K/V, gate/up); no residual with nseg_extra > 0; peers / a row offset only with epi_mode 1
    // (the peers then receive epi_out: lk_compute_swiglu_sharded)
    int nseg_extra = 0;
    const void *a_extra[2] = {nullptr, nullptr};
    float *dst_extra[2] = {nullptr, nullptr};
    int epi_mode = 0;               // 1 (exactly one extra segment): SwiGLU pair, epi_out = silu(dst) * dst_extra[0]
    float *epi_act = nullptr;       //   optional: silu(dst) (the SILU node's dst)
    float *epi_out = nullptr;
};
// Cross-GPU exchange folded into a sharded launch (no separate barrier kernel, SURVEY §8e / DESIGN §6).  An exchange
// point is a 64-bit arrival counter in every rank's peer-mapped flag buffer:
//   signal: every CTA of the launch, after its peer stores and a system-scope fence, adds its weight to every rank's
//           counter (sig_peer[0..sig_n), own rank included); the weights of one launch add up to 2^20 whatever its grid;
//   wait:   before the launch reads its activations every CTA waits until the local counter has received wait_n * 2^20 of
//           its epoch (epoch = ticket / grid size, wait_ticket is a local counter); bounded, time-outs counted in *timeout.
// sig_done is unused since the per-CTA arrival replaced the last-CTA funnel (kept so that the flag-buffer layout stays).
struct GemvSync {
    int sig_n = 0;
    unsigned long long *sig_peer[8] = {};
    unsigned long long *sig_done = nullptr;
    int wait_n = 0;
    unsigned long long *wait_ctr = nullptr, *wait_ticket = nullptr;
    unsigned *timeout = nullptr;
};
cudaError_t launch_gemv(int type, const void *a, int64_t K, int64_t M, const float *b, float *dst,
                        int64_t row_offset, const GemvPeers &peers, const GemvTuning &tune, int sm_count,
                        cudaStream_t s, const GemvFusion *fusion = nullptr, const GemvSync *sync = nullptr);

// K1c: up to GEMV_CHAIN_MAX dependent GEMVs (each with its optional fusions, no peers) as ONE persistent launch with a
// grid-wide barrier between phases; the weight stream runs ahead across the hand-offs.  Same results as the launches
// issued one by one.  bar: 3 zero-initialised device words owned by the context (arrivals, generation, time-outs).
constexpr int GEMV_CHAIN_MAX = 8;
struct GemvChainLink {
    const void *a = nullptr;
    int64_t K = 0, M = 0;
    const float *b = nullptr;
    float *dst = nullptr;
    GemvFusion fu;
};
cudaError_t launch_gemv_chain(int type, int n, const GemvChainLink *links, const GemvTuning &tune, int sm_count,
                              cudaStream_t s, unsigned *bar);
// K1m (lk_gemv_mma.cu): the same chain with the plane sums on the warp-level int8 MMA (16-row strips, 16 consumer warps
// on K-slices); Q4_0, K % 256 == 0.  cudaErrorNotSupported for everything else: the caller then takes launch_gemv_chain.
// bar: the same three words as launch_gemv_chain (the 64-bit counter doubles as the launch's epoch source); mbox: zero-initialised
// device scratch of GEMV_MCHAIN_MBOX_BYTES owned by the context (activations handed from phase to phase as (value, epoch) pairs).
constexpr int GEMV_MCHAIN_KMAX = 16384;
constexpr size_t GEMV_MCHAIN_MBOX_BYTES = 2 * (size_t)GEMV_MCHAIN_KMAX * 8;
constexpr size_t CHAIN_SCRATCH_BYTES = 256 + GEMV_MCHAIN_MBOX_BYTES;      // the context's chain words, then the mailbox
cudaError_t launch_gemv_mchain(int type, int n, const GemvChainLink *links, const GemvTuning &tune, int sm_count,
                               cudaStream_t s, unsigned *bar, void *mbox);

// Generic matmul for any (K, M, N) and a.type in {F32, Q4_0, Q4_1, Q8_0}; reference layouts.
cudaError_t launch_matmul_generic(int type, const void *a, int64_t K, int64_t M, const float *b, int64_t N,
                                  float *dst, cudaStream_t s, int nib = 0);

// N4 (lk_kquants.cu): the reference's Q2_K / Q4_K / Q8_K / BITNET_1_58 formats.  Decode GEMV for N == 1 and rows of
// whole blocks, the reference's literal block addressing for every other shape; bit-exact dequantize.
cudaError_t launch_kq_matmul(int type, const void *a, int64_t K, int64_t M, const float *b, int64_t N, float *dst,
                             int sm_count, cudaStream_t s);
cudaError_t launch_kq_dequantize(int type, const void *src, int64_t n, float *dst, cudaStream_t s);
int kq_type_block_bytes(int type);    // 84 / 144 / 292 / 10, else 0
int kq_type_block_elems(int type);    // 256 / 256 / 256 / 32, else 0

// K2: tcgen05 prefill GEMM (lk_gemm_tc.cu).  workspace holds the scaled FP16 copy (hi / lo parts) of b, the column
// maxima and the column factors.
size_t gemm_tc_workspace_bytes(int64_t K, int64_t N);
// cluster: 1, 2 or 4 CTAs (consecutive row tiles) share each activation tile by TMA multicast.
// precision: 2 = A and b as FP16 hi + lo pairs (3 MMAs per step, F32-grade operands; default), 1 = A hi + lo, b single
// FP16 (2 MMAs), 0 = single FP16 operands (1 MMA).
cudaError_t launch_gemm_tc(int type, const void *a, int64_t K, int64_t M, const float *b, int64_t N, float *dst,
                           void *workspace, int sm_count, int cluster, cudaStream_t s, int nib = 0, int precision = 2);

// K3-K5 + R4
// pdl: launch with programmatic stream serialization (the kernels wait for their predecessor themselves), so the
// launch overlaps the predecessor's tail and a following GEMV can prefetch its weights under the op.
cudaError_t launch_rms_norm(const float *x, const float *w, float eps, int64_t ne0, int64_t nrows, float *dst,
                            cudaStream_t s, bool pdl = false);
cudaError_t launch_silu(const float *x, int64_t n, float *dst, cudaStream_t s, bool pdl = false);
cudaError_t launch_silu_mul(const float *g, const float *u, int64_t n, float *dst, cudaStream_t s, bool pdl = false);
cudaError_t launch_softmax(const float *x, int64_t ne0, int64_t nrows, float *dst, cudaStream_t s, bool pdl = false);
cudaError_t launch_add(const float *a, const float *b, int64_t n, float *dst, cudaStream_t s, bool pdl = false);
cudaError_t launch_mul(const float *a, const float *b, int64_t n, float *dst, cudaStream_t s, bool pdl = false);
cudaError_t launch_scale(const float *x, float factor, int64_t n, float *dst, cudaStream_t s, bool pdl = false);

// N4 attention core (lk_attention.cu).  pos: device int32 scalar (may be NULL = 0 for rope / append).
cudaError_t launch_rope(const float *x, const float *table, const int32_t *pos, int head_dim, int64_t row, int ntok,
                        int max_pos, float *dst, cudaStream_t s, bool pdl = false);
cudaError_t launch_kv_append(const float *src, float *cache, const int32_t *pos, int64_t row, int ntok, int64_t max_seq,
                             cudaStream_t s, bool pdl = false);
size_t attn_workspace_bytes(int head_dim, int nheads, int64_t max_seq);
// step != NULL: the fused decode step — q (and step->k_raw, v_raw) are this token's projections before RoPE; the kernel
// rotates q / k, appends k / v to the caches at pos (kc / vc are written) and attends, in one launch.
struct AttnStep {
    const float *k_raw, *v_raw;
    const float *table;       // [table_pos][head_dim/2][2] (cos, sin), lk_rope_table
    int table_pos;
    float *q_rope, *k_rope;   // optional: where the rotated q / k are also stored
};
cudaError_t launch_attn_decode(const float *q, const float *kc, const float *vc, const int32_t *pos, int head_dim,
                               int nheads, int64_t max_seq, float scale, float *dst, void *workspace, int sm_count,
                               cudaStream_t s, bool pdl = false, const AttnStep *step = nullptr);

// The decode attention with its scratch memory in a file-scope static device array (no allocation of any kind); see
// lk_attention.cu.  cudaErrorInvalidValue: the shape needs more than attn_static_scratch_capacity() bytes.
size_t attn_static_scratch_capacity();
size_t attn_scratch_needed(int head_dim, int nheads, int64_t max_seq, int sm_count);
cudaError_t launch_attn_decode_static(int device, const float *q, const float *kc, const float *vc, const int32_t *pos, int head_dim,
                                      int nheads, int64_t max_seq, float scale, float *dst, int sm_count, cudaStream_t s);

}  // namespace lk

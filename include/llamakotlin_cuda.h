/*
 * llamakotlin_cuda.h — C ABI of the B200 (sm_100a) implementation of
 * llama.kotlin's quantized-matmul hot path.
 *
 * This is the drop-in boundary: Kotlin/Native binds it through cinterop
 * (kotlin/llamakotlin_cuda.def), tests bind it through ctypes.  Plain C,
 * plain pointers and sizes; no CUDA, torch or C++ types in any signature.
 *
 * Every entry point names the reference interface it replaces.  Reference
 * paths are relative to
 *   /root/reference/src/nativeMain/kotlin/ai/solace/llamakotlin/
 *
 * Conventions (all from the reference):
 *   a   = weight,      a.ne  = [K, M]  (K fastest; M rows of K/32 blocks)
 *   b   = activation,  b.ne  = [N, K]  (N fastest; element (j,k) at (j+k*N)*4)
 *   dst = F32,         dst.ne= [N, M]  (element (j,i) at (j+i*N)*4)
 *   Q8_0 block 34 B: d:f16@0, qs:int8[32]@2
 *   Q4_0 block 18 B: d:f16@0, qs:u8[16]@2, element e -> (qs[e/2] >> 4*(e&1)) & 15
 *   Q4_1 block 20 B: d:f16@0, m:f16@2, qs:u8[16]@4, same nibble map
 *   (the reference's own nibble order, core/GGMLTypes.kt:647-651, NOT ggml's)
 *
 * Threading: one lk_ctx per device; calls on a ctx are serialised by the
 * caller (the reference is single-threaded).  No global mutable state.
 *
 * There is NO CPU fallback: every compute entry point returns LK_ECUDA if the
 * device path is unavailable.
 */
#ifndef LLAMAKOTLIN_CUDA_H
#define LLAMAKOTLIN_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define LK_API __attribute__((visibility("default")))
#else
#define LK_API
#endif

#define LK_ABI_VERSION 3
#define LK_MAX_DIMS 4           /* GGML_MAX_DIMS, core/GGMLTypes.kt:67 */

/* GGMLType ordinals, core/GGMLTypes.kt:145-153 */
enum {
    LK_TYPE_F32  = 0,
    LK_TYPE_F16  = 1,
    LK_TYPE_Q4_0 = 2,
    LK_TYPE_Q4_1 = 3,
    LK_TYPE_Q8_0 = 6,
    /* SURVEY §8f N4 — the reference's own K-quant formats (256-element super-blocks of 84 / 144 / 292 bytes,
     * core/GGMLTypes.kt:117-122; field meaning per its accessors :737-918 and dot products
     * core/GGMLComputeOps.kt:152-432, which is NOT ggml's) and BitNet 1.58 (32 ternary weights in 10 bytes,
     * core/GGMLTypes.kt:127, :920-1000).  lk_compute_matmul (x F32 -> F32) and lk_compute_dequantize accept them;
     * the fused / sharded / tensor-core entry points do not.  BITNET_1_58 has no GGMLType.fromValue id in the
     * reference (:145-165); 40 is this ABI's own. */
    LK_TYPE_Q2_K = 8,
    LK_TYPE_Q4_K = 10,
    LK_TYPE_Q8_K = 13,
    LK_TYPE_I32  = 17,  /* position scalars of the attention ops */
    LK_TYPE_BITNET_1_58 = 40
};

/* Status codes.  The Kotlin wrapper rethrows the same exception class the
 * reference throws at the cited line:
 *   LK_EINVAL_DIM / LK_EINVAL_TYPE / LK_EINVAL_ARG -> IllegalArgumentException
 *       (core/GGMLComputeOps.kt:1440, :1444-1446, :1449, :1463, :1517, :1530)
 *   LK_EBOUNDS   -> IndexOutOfBoundsException (core/GGMLTypes.kt:556-558)
 *   LK_ENOTIMPL  -> NotImplementedError       (core/GGMLComputeOps.kt:1563, :2548)
 *   LK_ECUDA / LK_ENCCL / LK_ENOMEM -> IllegalStateException; at backend level
 *       GGMLStatus.FAILED (core/GGMLCpuBackend.kt:168-175)                   */
typedef int32_t lk_status;
enum {
    LK_OK          = 0,
    LK_EINVAL_DIM  = 1,
    LK_EINVAL_TYPE = 2,
    LK_EBOUNDS     = 3,
    LK_ENOTIMPL    = 4,
    LK_ECUDA       = 5,
    LK_ENCCL       = 6,
    LK_EINVAL_ARG  = 7,
    LK_ENOMEM      = 8
};

typedef struct lk_ctx   lk_ctx;    /* one device + one stream */
typedef struct lk_buf   lk_buf;    /* device buffer == GGMLBackendBuffer on "CUDA" */
typedef struct lk_graph lk_graph;  /* captured op sequence (CUDA graph) */

/* Mirror of the fields of GGMLTensor (core/GGMLTypes.kt:251-267) that the hot
 * path reads, plus where its bytes live.  Exactly one of host_base / dev_buf
 * is non-NULL:
 *   host_base = address of graphAllocator.buffers[bufferId][0] while pinned
 *               (usePinned { addressOf(0) }); buffer_size = that array's size.
 *   dev_buf   = a device buffer from lk_buf_alloc (GGMLCudaBuffer).
 * data_offset = tensor.dataOffset.  nb[]: block types ignore it like the
 * reference does.  F32 tensors are addressed through nb[] like
 * GGMLTensor.getFloat / setFloat (core/GGMLTypes.kt:328-380): element
 * (i0..i3) at data_offset + sum i_d * nb[d]; all-zero nb[] means contiguous.
 * A non-contiguous F32 operand is gathered into (a result scattered from) a
 * contiguous staging copy by one extra kernel; strides must be multiples of
 * 4 bytes; not available inside a captured sequence or in the fused / sharded
 * entry points (device-resident contiguous vectors only).                     */
typedef struct lk_tensor {
    int32_t  type;
    int32_t  reserved;
    int64_t  ne[LK_MAX_DIMS];
    uint64_t nb[LK_MAX_DIMS];
    void    *host_base;
    lk_buf  *dev_buf;
    uint64_t buffer_size;
    uint64_t data_offset;
} lk_tensor;

/* ---- context ------------------------------------------------------------ */
LK_API int32_t     lk_abi_version(void);
/* bytes per block (GGMLType.byteSize, core/GGMLTypes.kt:99-133: 32-element blocks, 256-element K-quant
 * super-blocks); 4 for F32, 0 if unknown */
LK_API int32_t     lk_type_block_bytes(int32_t type);
/* calculateTensorByteSize, core/GGMLTypes.kt:1136-1190 (n/32*blockBytes for Q types) */
LK_API int64_t     lk_tensor_nbytes(const lk_tensor *t);
LK_API lk_status   lk_ctx_create(int32_t device, lk_ctx **out);
LK_API lk_status   lk_ctx_destroy(lk_ctx *ctx);
/* GGMLBackend.synchronize(), core/GGMLBackend.kt:90-157 */
LK_API lk_status   lk_sync(lk_ctx *ctx);
/* message of the last non-OK status on this ctx (or a static string if ctx==NULL) */
LK_API const char *lk_last_error(lk_ctx *ctx);
/* the ctx's cudaStream_t as an opaque pointer (interop / external timing) */
LK_API void       *lk_ctx_stream(lk_ctx *ctx);
LK_API int32_t     lk_ctx_device(lk_ctx *ctx);
LK_API int32_t     lk_ctx_sm_count(lk_ctx *ctx);
/* number of kernels this ctx has launched since creation (bench: gpu_launches) */
LK_API int64_t     lk_ctx_launch_count(lk_ctx *ctx);

/* CUDA-event timer on the ctx stream (the stream the kernels run on). */
LK_API lk_status   lk_timer_start(lk_ctx *ctx);
LK_API lk_status   lk_timer_stop(lk_ctx *ctx, float *ms_out);   /* synchronises */

/* ---- device buffers: GGMLBackendBuffer, core/GGMLBackend.kt:45-75 ------- */
LK_API lk_status   lk_buf_alloc(lk_ctx *ctx, uint64_t size, lk_buf **out);      /* allocBuffer */
LK_API lk_status   lk_buf_free(lk_buf *buf);                                    /* free */
LK_API uint64_t    lk_buf_size(const lk_buf *buf);                              /* getSize */
LK_API void       *lk_buf_device_ptr(lk_buf *buf);                              /* getBase */
LK_API lk_status   lk_buf_upload(lk_buf *buf, uint64_t offset, const void *host, uint64_t n);   /* setTensor */
LK_API lk_status   lk_buf_download(lk_buf *buf, uint64_t offset, void *host, uint64_t n);       /* getTensor */
LK_API lk_status   lk_buf_copy(lk_buf *dst, uint64_t dst_off, lk_buf *src, uint64_t src_off, uint64_t n); /* copyTensor */
/* setTensor / getTensor without the wait (core/GGMLBackend.kt:95-110 are synchronous; these are what lets the copies of
 * a step overlap its launches).  The copies run on two streams of their own beside the compute stream:
 *   upload_async   starts at once — it does NOT wait for launches already enqueued, so the caller must not overwrite
 *                  memory they still read (double-buffer, or lk_sync between steps); every lk_compute_* call enqueued
 *                  AFTER it waits for it.
 *   download_async starts when everything enqueued on the compute stream so far has finished; `host` is valid after the
 *                  next lk_sync (which waits for all three streams).
 * `host` should be pinned (lk_host_alloc), else the driver serialises the copy.  Inside lk_capture_begin / _end the
 * copies become nodes of the captured graph (the host addresses are part of it): uploads hang off the graph's root,
 * downloads off the launch that precedes them, and a replay (lk_graph_launch) finishes only when they have.            */
LK_API lk_status   lk_buf_upload_async(lk_buf *buf, uint64_t offset, const void *host, uint64_t n);
LK_API lk_status   lk_buf_download_async(lk_buf *buf, uint64_t offset, void *host, uint64_t n);
LK_API lk_status   lk_buf_clear(lk_buf *buf, uint8_t value);                    /* clear */
/* Pinned host staging memory (page-locked) for callers without their own. */
LK_API lk_status   lk_host_alloc(lk_ctx *ctx, uint64_t size, void **out);
LK_API lk_status   lk_host_free(lk_ctx *ctx, void *p);

/* ---- compute ops --------------------------------------------------------
 * Destination-based like the reference: the caller owns dst, the op fills it.
 * If every tensor is device-resident the op is stream-ordered (asynchronous;
 * lk_sync or lk_buf_download observes it).  Any host tensor is staged through
 * the ctx (H2D before, D2H after) and the call returns with dst valid, like
 * the reference (tests read dst immediately,
 * nativeTest core/GGMLComputeOpsDestinationTest.kt:256-264).
 * Validation order and error classes follow the cited reference lines.       */

/* computeMatMul(graphAllocator, context, a, b, dst) — core/GGMLComputeOps.kt:1435.
 * a.type in {F32, Q4_0, Q4_1, Q8_0} (and the K-quant / BitNet types above), b F32, dst F32.  Routing:
 *   N == 1, K % 64 == 0        -> K1 streaming decode GEMV (lk_gemv.cu)
 *   N >= 2, K % 256 == 0, M*K*N >= 2^21 (options gemm_tc_min_n = 2, gemm_tc_min_work)
 *                              -> K2 tcgen05 GEMM (lk_gemm_tc.cu); N < 32 is padded to one 32-column tile
 *   otherwise                  -> generic F32 kernel (the reference's per-element arithmetic, warp-strided sums)
 * Arithmetic contract (the reference is F32 throughout, :1516-1527; tolerance of the path: 1e-3 of max|dst|):
 *   K1: integer block dot products are exact; the activations are held as T_k = rint(x_k * 2^(22-E)) with ONE exponent E
 *       per 32-element block (E = exponent of the block's max |x|), so x_k carries an absolute error <= 2^(E-23):
 *       |dst_i - exact_i| <= 2^-23 * sum_k |w_ik| * bmax(k) + F32 summation terms (measured 1e-6 .. 4e-6 of max|dst|).
 *       A block that holds a non-zero |x| below 2^-20 of its maximum (option gemv_fix_gap) is recomputed from the F32
 *       activations (exact residuals), so elements the block scale cannot represent are not lost; non-finite
 *       activations give non-finite outputs like the reference.  The opt-in 16-bit mode (option gemv_act_planes = 2:
 *       T_k = rint(x_k * 2^(14-E)), absolute error <= 2^(E-15) per element) runs without that recomputation unless
 *       gemv_fix_gap is set as well (the gap then counts from 2^-12).
 *   K2: option gemm_tc_precision = 2 (default): d*q and b as FP16 hi + lo pairs (d*q exactly; b to 22 bits), three
 *       MMAs per K step, F32 accumulation in tensor memory: max-norm error ~4e-6; per element <= 1.5e-6 * K relative
 *       (the accumulator truncates).  1: exact weights x FP16 activations (max-norm ~2.5e-4).  0: single FP16 operands
 *       (~3e-4, up to 1.02e-3 on periodic data).  Any F32 magnitude of b is accepted (per-column power-of-two scaling,
 *       exact); weight scales that could leave the FP16 range make the affected tile take an F32 path.                */
LK_API lk_status   lk_compute_matmul(lk_ctx *ctx, const lk_tensor *a, const lk_tensor *b, const lk_tensor *dst);

/* computeMatMul with its neighbours in the layer graph folded into the launch (op order of
 * LlamaDecoderLayer.forward, model/LlamaModel.kt:253-285).  Every tensor the unfused ops would have written is
 * still written, so the graph's intermediates stay valid:
 *   norm_x   != NULL: the activations are RMSNorm(norm_x, norm_w or NULL, eps); `b` (the RMS_NORM node's dst) also
 *                     receives them.  1/rms multiply instead of the op's division: <= 2 ulp from lk_compute_rms_norm.
 *   residual != NULL: sum_dst = residual + a x b (the following ADD node's dst); dst still receives a x b.
 * N == 1, Q-type a, device-resident tensors only: anything else returns LK_ENOTIMPL and the caller issues the
 * unfused ops.                                                                                             */
typedef struct lk_matmul_fusion {
    const lk_tensor *norm_x;
    const lk_tensor *norm_w;
    float            eps;
    const lk_tensor *residual;
    const lk_tensor *sum_dst;
    int32_t          flags;      /* LK_FUSE_SKIP_B_STORE: another launch of the same RMS_NORM already stored `b` */
    /* Stacked matmuls: n_extra (0..2) more weights of a's type and shape multiplied with the same activations by the
     * same launch (Q/K/V, gate/up): dst_extra[i] = a_extra[i] x b.  Not combinable with residual.
     * With n_extra == 1 and mul_dst != NULL the SwiGLU pair of LlamaMLP.forward (model/LlamaModel.kt:159-176) is
     * finished in the epilogue: silu_dst (optional) = silu(dst), mul_dst = silu(dst) * dst_extra[0], with the
     * elementwise kernels' roundings (lk_compute_silu, lk_compute_mul).                                          */
    int32_t          n_extra;
    const lk_tensor *a_extra[2];
    const lk_tensor *dst_extra[2];
    const lk_tensor *silu_dst;
    const lk_tensor *mul_dst;
} lk_matmul_fusion;
#define LK_FUSE_SKIP_B_STORE 1
LK_API lk_status   lk_compute_matmul_fused(lk_ctx *ctx, const lk_tensor *a, const lk_tensor *b, const lk_tensor *dst,
                                           const lk_matmul_fusion *fusion);
/* A run of DEPENDENT decode matmuls — the four launches of LlamaDecoderLayer.forward (model/LlamaModel.kt:253-285:
 * [RMSNorm -> Q/K/V], [O + residual], [RMSNorm -> gate/up + SwiGLU], [down + residual]), possibly across layers — as ONE
 * persistent launch: link i is exactly the call lk_compute_matmul_fused(a, b, dst, fusion) (fusion may be NULL = plain
 * N == 1 matmul), links run in order and the weight stream runs ahead across the hand-offs.  Up to 8 links, one weight
 * type, device-resident tensors; anything else returns LK_ENOTIMPL and the caller issues the links itself.
 * Two kernels stand behind it.  K1m (lk_gemv_mma.cu; K % 256 == 0, every tensor the launch writes written by exactly one
 * link, activations either the previous link's output or from outside the launch): plane sums on the warp-level int8
 * MMA, phases handed over through a mailbox of (value, epoch) pairs instead of a barrier; results equal the links
 * issued one by one to F32 rounding (same exact integer sums per block, different order of the F32 additions; measured
 * <= 1e-6 of the largest output).  K1c (lk_gemv.cu; everything K1m declines, or everything after
 * lk_set_option(ctx, "gemv_chain_mma", 0)): DP4A plane sums, a grid-wide barrier between links, results identical to the
 * links issued one by one (a link with the RMSNorm prologue may differ in the last bit of 1/rms).
 * lk_get_option(ctx, "gemv_chain_mma_launches") counts the launches K1m took.
 * The launch is one CTA per SM and all of them must be resident at once: do not run it beside kernels of other
 * streams that hold SMs indefinitely.  The waits are bounded (the device never hangs); a time-out leaves wrong
 * results behind and is counted: lk_get_option(ctx, "gemv_chain_timeouts") synchronises, returns the count since the
 * last read (0 in a healthy run) and re-arms the counter.  lk_set_option(ctx, "gemv_chain", 0) makes this call decline. */
typedef struct lk_matmul_link {
    const lk_tensor *a, *b, *dst;
    const lk_matmul_fusion *fusion;
} lk_matmul_link;
LK_API lk_status   lk_compute_matmul_chain(lk_ctx *ctx, const lk_matmul_link *links, int32_t n);
/* dequantizeTensor(graphAllocator, tensor) — core/GGMLComputeOps.kt:918-964. Bit-exact. dst F32, same ne. */
LK_API lk_status   lk_compute_dequantize(lk_ctx *ctx, const lk_tensor *src, const lk_tensor *dst);
/* quantizeTensor(graphAllocator, tensor, targetType) — core/GGMLComputeOps.kt:1040-1433 (Q8_0 :1058-1071, Q4_0 :1072-1087,
 * Q4_1 :1162-1203), destination-based: src F32, dst Q8_0 / Q4_0 / Q4_1 with the same element count (a multiple of 32).
 * Bit-exact with the reference incl. floatToHalf's half-denormal off-by-one (core/NumericConversions.kt:83) and the
 * Kotlin quantizers' formulas (Q4_0: d = amax / 8, round-half-even), in the context's nibble_order.                */
LK_API lk_status   lk_compute_quantize(lk_ctx *ctx, const lk_tensor *src, const lk_tensor *dst);
/* computeAdd / computeMul F32 same-shape paths — core/GGMLComputeOps.kt:735-759, :811-835 */
LK_API lk_status   lk_compute_add(lk_ctx *ctx, const lk_tensor *a, const lk_tensor *b, const lk_tensor *dst);
LK_API lk_status   lk_compute_mul(lk_ctx *ctx, const lk_tensor *a, const lk_tensor *b, const lk_tensor *dst);
/* RMS_NORM node (rmsNormOp, model/GGMLIntegration.kt:181-201; src[1]=weight, opParams[0]=eps bits).
 * The reference executor leaves it a no-op (core/GGMLComputeOps.kt:2640-2645); semantics follow
 * nativeTest core/GGMLComputeOpsTest.kt:502-521 and model/LlamaModel.kt:48-63.  weight may be NULL. */
LK_API lk_status   lk_compute_rms_norm(lk_ctx *ctx, const lk_tensor *x, const lk_tensor *weight, float eps, const lk_tensor *dst);
/* SILU node (siluOp, model/GGMLIntegration.kt:206-220): x * 1/(1+exp(-x)),
 * nativeTest core/GGMLComputeOpsTest.kt:491-499 (the executor's GELU fallback is a reference defect). */
LK_API lk_status   lk_compute_silu(lk_ctx *ctx, const lk_tensor *x, const lk_tensor *dst);
/* SwiGLU: silu(gate) * up in one pass — model/LlamaModel.kt:165-175 */
LK_API lk_status   lk_compute_silu_mul(lk_ctx *ctx, const lk_tensor *gate, const lk_tensor *up, const lk_tensor *dst);
/* SOFT_MAX node over ne[0] — LlamaAttention.softmax, model/LlamaAttention.kt:242-288 */
LK_API lk_status   lk_compute_softmax(lk_ctx *ctx, const lk_tensor *x, const lk_tensor *dst);

/* ---- attention core -------------------------------------------
 * The position is an I32 tensor with one element (device-resident inside captured sequences), like ggml's `pos`
 * operand of ROPE, so that one captured layer graph serves every token.  pos == NULL means position 0.           */
/* SCALE node — LlamaAttention.scale, model/LlamaAttention.kt:166-195: value * factor */
LK_API lk_status   lk_compute_scale(lk_ctx *ctx, const lk_tensor *x, float factor, const lk_tensor *dst);
/* (cos, sin) of LlamaAttention.applyRoPE (model/LlamaAttention.kt:53-58) for positions [0, max_pos), computed on the
 * host in the reference's F64 operations and rounded to F32 as it does: out[(p * head_dim/2 + i/2) * 2 + {0, 1}].
 * Upload it once as an F32 tensor with ne = [2, head_dim/2, max_pos].                                          */
LK_API lk_status   lk_rope_table(int32_t head_dim, float rope_theta, int32_t max_pos, float *out);
/* ROPE node — applyRoPE, model/LlamaAttention.kt:29-72: x.ne[0] = numHeads * headDim, rows = consecutive positions
 * starting at pos; rotates the pairs (i, i+1) of every head: (x cos - y sin, x sin + y cos), bit-exact.          */
LK_API lk_status   lk_compute_rope(lk_ctx *ctx, const lk_tensor *x, const lk_tensor *table, const lk_tensor *pos,
                                   int32_t head_dim, const lk_tensor *dst);
/* KVCache.updateKey / updateValue copy loops, model/KVCache.kt:47-55, :88-96: rows of src (ne = [headDim, numHeads,
 * newSeqLen]) go to cache rows pos .. pos+newSeqLen-1 (cache ne = [headDim, numHeads, maxSeq], device-resident).
 * The `require(currentLength + newSeqLen <= maxSequenceLength)` of :44 is the caller's (it knows the length); the
 * kernel never writes outside the cache.                                                                         */
LK_API lk_status   lk_compute_kv_append(lk_ctx *ctx, const lk_tensor *src, const lk_tensor *cache, const lk_tensor *pos);
/* LlamaAttention.computeAttention for one query position, model/LlamaAttention.kt:104-123, per head:
 * dst[h] = softmax(scale * q[h] . K[0..pos, h]) . V[0..pos, h], softmax as :242-288; q, dst hold numHeads * headDim
 * floats, caches ne = [headDim, numHeads, maxSeq]; length = pos + 1 (the row appended for this token included).   */
LK_API lk_status   lk_compute_attention_decode(lk_ctx *ctx, const lk_tensor *q, const lk_tensor *kcache,
                                               const lk_tensor *vcache, const lk_tensor *pos, float scale,
                                               const lk_tensor *dst);
/* lk_compute_attention_decode for callers that cannot allocate (code under CUDA-graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same arguments, same kernel and same results as lk_compute_attention_decode, but the call allocates nothing — no
 * cudaMalloc, cudaFree, pinned memory or staging arena: all tensors must be device-resident and contiguous (a host
 * tensor returns LK_EINVAL_ARG), and the kernel's scratch memory (per-head tickets + split partials) is a file-scope
 * static device array of LK_ATTN_NOALLOC_SCRATCH_BYTES that every GPU which loads the library holds until the process
 * exits.  A shape that needs more (256-byte-rounded numHeads * 4 + numHeads * splits * (headDim + 4) * 4 bytes, splits
 * <= 3 * SMs / numHeads: 234 KB for 32 heads of 128) is rejected with LK_EINVAL_ARG and a message that names both
 * sizes.  All calls on one device share the array and find the previous call's contents in it (the part that must
 * start at zero is cleared in front of every launch, stream-ordered): calls that overlap on the GPU — two streams or
 * two contexts of one device — race; serialise them.                                                              */
#define LK_ATTN_NOALLOC_SCRATCH_BYTES (512 * 1024)
LK_API lk_status   lk_compute_attention_decode_noalloc(lk_ctx *ctx, const lk_tensor *q, const lk_tensor *kcache,
                                                       const lk_tensor *vcache, const lk_tensor *pos, float scale,
                                                       const lk_tensor *dst);
/* One decode step of LlamaAttention.forward's core in ONE launch, model/LlamaAttention.kt:87-99 + :104-123:
 * qRope = applyRoPE(q), kRope = applyRoPE(k), kvCache.updateKey(kRope), kvCache.updateValue(v), computeAttention.
 * q, k, v: this token's projections (numHeads * headDim floats each) before RoPE; table / pos / head_dim as
 * lk_compute_rope; caches as lk_compute_kv_append (row `pos` of both is written); q_rope_dst / k_rope_dst (optional,
 * may be NULL) also receive the rotated vectors, so a graph that names them as nodes keeps them valid.  Results are
 * bit-identical to lk_compute_rope x 2 + lk_compute_kv_append x 2 + lk_compute_attention_decode.  Device-resident
 * tensors only.                                                                                                    */
LK_API lk_status   lk_compute_attention_step(lk_ctx *ctx, const lk_tensor *q, const lk_tensor *k, const lk_tensor *v,
                                             const lk_tensor *table, const lk_tensor *pos, int32_t head_dim,
                                             const lk_tensor *kcache, const lk_tensor *vcache, float scale,
                                             const lk_tensor *q_rope_dst, const lk_tensor *k_rope_dst,
                                             const lk_tensor *dst);

/* ---- op-sequence capture: GGMLBackend.graphCompute, core/GGMLBackend.kt:145 --
 * Between begin and end, device-resident lk_compute_* calls are recorded, not
 * run; lk_graph_launch replays the whole sequence with one submission
 * (CUDA graph with programmatic dependent launch between GEMV nodes).        */
LK_API lk_status   lk_capture_begin(lk_ctx *ctx);
LK_API lk_status   lk_capture_end(lk_ctx *ctx, lk_graph **out);
LK_API lk_status   lk_graph_launch(lk_ctx *ctx, lk_graph *g);
LK_API lk_status   lk_graph_destroy(lk_graph *g);

/* ---- whole graphs: GGMLBackend.graphCompute with the fusion / chaining plan inside the library ------------------
 * GGMLComputeOps.computeGraph walks graph.nodes[0 .. nNodes) with `when (node.op)` (core/GGMLComputeOps.kt:2515-2550).
 * lk_graph_compute takes the same node list — every node with its op, its own tensor (the dst), its sources and
 * opParams — plans which neighbours fold into the decode GEMV launches (RMS_NORM -> MUL_MAT x up to 3, MUL_MAT -> ADD,
 * gate / up -> SILU -> MUL, runs of dependent launches as lk_compute_matmul_chain, ROPE ROPE CPY CPY FLASH_ATTN as
 * lk_compute_attention_step) and issues the calls on the ctx stream; a launch the library declines for its shape falls
 * back to the unfused ops in graph order.  Every folded node's dst is still written.  Call it between
 * lk_capture_begin / lk_capture_end to get a replayable lk_graph (device-resident tensors), or directly.
 * Source slots per op (the node constructors of model/GGMLIntegration.kt:118-220 and this ABI's attention ops):
 *   MUL_MAT a, b | ADD / MUL a, b | SILU / SOFT_MAX / SCALE x | RMS_NORM x, weight | ROPE x, position, table |
 *   CPY new rows, cache, position | FLASH_ATTN q, key cache, value cache, position;
 *   op_params[0]: RMS_NORM eps bits, SCALE factor bits, FLASH_ATTN scale bits, ROPE headDim.                       */
enum {
    LK_OP_NONE = 0, LK_OP_ADD = 1, LK_OP_MUL = 2, LK_OP_MUL_MAT = 3, LK_OP_RMS_NORM = 4, LK_OP_SILU = 5, LK_OP_SOFT_MAX = 6,
    LK_OP_SCALE = 7, LK_OP_ROPE = 8, LK_OP_CPY = 9, LK_OP_FLASH_ATTN = 10
};
typedef struct lk_graph_node {
    int32_t   op;
    int32_t   n_src;             /* 0 .. 4 */
    lk_tensor dst;               /* the node's own tensor */
    lk_tensor src[4];
    int32_t   op_params[4];
} lk_graph_node;
/* One launch of the plan (lk_graph_plan: the planning half alone, host logic only, no device needed).  Index fields
 * are node indices, -1 = none.  Consecutive launches with the same chain >= 0 go out as one lk_compute_matmul_chain. */
enum { LK_PLAN_NODE = 0, LK_PLAN_FUSED = 1, LK_PLAN_ATTN_STEP = 2 };
typedef struct lk_plan_launch {
    int32_t kind;
    int32_t node;                /* the node the launch is made at (LK_PLAN_ATTN_STEP: the ROPE(q) node) */
    int32_t rms, add;            /* RMS_NORM / ADD node folded into this GEMV */
    int32_t n_extra, extra[2];   /* stacked MUL_MAT nodes */
    int32_t silu, mul;           /* SwiGLU epilogue */
    int32_t store_b;             /* 1: this launch also stores the RMS_NORM's dst (the first consumer does) */
    int32_t chain;
    int32_t attn[5];             /* LK_PLAN_ATTN_STEP: ROPE(q), ROPE(k), CPY(k), CPY(v), FLASH_ATTN */
} lk_plan_launch;
/* chain_mode: -1 = where it was measured to pay (graphs with an attention step between the runs, and graphs whose decode
 * GEMVs all have K % 256 == 0: the MMA chain kernel K1m takes Q4_0 / Q4_1 and short Q8_0 rows, K1c the rest), 0 never, 1 always */
LK_API lk_status   lk_graph_plan(const lk_graph_node *nodes, int32_t n, int32_t chain_mode, lk_plan_launch *out,
                                 int32_t max_out, int32_t *n_out);
#define LK_GRAPH_NO_FUSION    1   /* node by node, like the reference's interpreter */
#define LK_GRAPH_CHAIN_ALWAYS 2
#define LK_GRAPH_CHAIN_NEVER  4
LK_API lk_status   lk_graph_compute(lk_ctx *ctx, const lk_graph_node *nodes, int32_t n, int32_t flags);
/* The same with the step's host copies scheduled inside it: every upload is issued (lk_buf_upload_async) right before
 * the first launch that reads its region, every download (lk_buf_download_async) right after the last launch that
 * writes it, so a caller whose activations arrive from and whose results go back to host memory (the reference's
 * setTensor -> graphCompute -> getTensor sequence) pays one chunk of each at the ends of the step instead of both
 * copies in full.  Split the transfers by consumer (e.g. one per layer) for the overlap to exist.                     */
enum { LK_IO_UPLOAD = 0, LK_IO_DOWNLOAD = 1 };
typedef struct lk_host_io {
    int32_t   direction;         /* LK_IO_UPLOAD | LK_IO_DOWNLOAD */
    int32_t   reserved;
    lk_buf   *buf;
    uint64_t  offset, bytes;     /* the region of buf */
    void     *host;
} lk_host_io;
LK_API lk_status   lk_graph_compute_io(lk_ctx *ctx, const lk_graph_node *nodes, int32_t n, int32_t flags,
                                       const lk_host_io *io, int32_t n_io);

/* ---- options ----------------------------------------------------------------
 * "nibble_order" (product option): how the 32 four-bit quants of a Q4_0 / Q4_1 block lie in its 16
 *   quant bytes.  0 (default) = the reference's accessors, getQ4_0NibbleWeight / getQ4_1NibbleWeight
 *   (core/GGMLTypes.kt:627-653, :706-732): byte j = elements 2j (low nibble) and 2j+1 (high).  1 = canonical ggml,
 *   i.e. what every real GGUF file holds (ggml/src/ggml-quants.c:1515-1555): byte j = elements j (low) and j+16
 *   (high).  Applies to lk_compute_matmul (all three kernels), the fused / sharded matmuls and lk_compute_dequantize
 *   of the context; Q8_0 blocks are identical in both orders.
 * everything else: tuning knobs (bench / tests only; defaults are the product settings) */
LK_API lk_status   lk_set_option(lk_ctx *ctx, const char *name, int64_t value);
LK_API int64_t     lk_get_option(lk_ctx *ctx, const char *name);

/* ---- multi-GPU: weight rows sharded over ranks, F32 output all-gathered ----
 * One process per GPU.  Rank 0 makes an id, the host side broadcasts it
 * (torch.distributed / any channel), every rank joins.  No reference
 * counterpart exists; layout rules are §8(e).                    */
#define LK_COMM_ID_BYTES 128
#define LK_IPC_HANDLE_BYTES 64
LK_API lk_status   lk_comm_unique_id(uint8_t id_out[LK_COMM_ID_BYTES]);
LK_API lk_status   lk_comm_init(lk_ctx *ctx, int32_t nranks, int32_t rank, const uint8_t id[LK_COMM_ID_BYTES]);
LK_API lk_status   lk_comm_destroy(lk_ctx *ctx);
/* in-place ncclAllGather of `count` floats per rank inside buf at off (rank r's slice at off + r*count*4) */
LK_API lk_status   lk_allgather_f32(lk_ctx *ctx, lk_buf *buf, uint64_t off, int64_t count);
/* peer-mapped buffers for the fused GEMV + all-gather (direct NVLink stores) */
LK_API lk_status   lk_buf_ipc_export(lk_buf *buf, uint8_t handle_out[LK_IPC_HANDLE_BYTES]);
LK_API lk_status   lk_buf_ipc_import(lk_ctx *ctx, const uint8_t handle[LK_IPC_HANDLE_BYTES], uint64_t size, lk_buf **out);
/* Cross-rank barrier on peer-mapped flag words (each rank's flag buffer holds 16 words, zeroed
 * before first use): publishes `epoch` to every peer after the preceding ops on the ctx stream,
 * then waits (bounded, ~2 s; word 14 of the own buffer is set on time-out) until every peer has
 * published it here.  epoch must increase by 1 per call (wrap-safe compare); epoch == 0 uses a
 * device-side counter instead, so the call can sit in a captured op sequence that is replayed.
 * The first call (which uploads the peer table) must be made outside capture. */
LK_API lk_status   lk_peer_barrier(lk_ctx *ctx, lk_buf *own_flags, lk_buf *const *peer_flags, int32_t npeers,
                            int32_t rank, uint32_t epoch);
/* All-gather as peer stores: copies `count` floats at byte offset `off` of `local` to the same offset of every
 * buffer in peer_bufs (entries equal to `local` or NULL are skipped) over NVLink, stream-ordered; follow it
 * with lk_peer_barrier.  Used where the slice comes from an elementwise op instead of a GEMV epilogue. */
LK_API lk_status   lk_peer_push_f32(lk_ctx *ctx, lk_buf *local, uint64_t off, int64_t count,
                             lk_buf *const *peer_bufs, int32_t npeers);
/* Sharded computeMatMul, N == 1: this rank owns rows [row0, row0 + a_shard.ne[1]) of the
 * M_total-row weight; writes its slice of dst_full (ne=[1, M_total]) locally and, when
 * peers != NULL, into the same offsets of every peer's buffer from the GEMV epilogue,
 * then signals flags.  npeers = nranks (entry `rank` is this rank's own dst).           */
LK_API lk_status   lk_compute_matmul_sharded(lk_ctx *ctx, const lk_tensor *a_shard, const lk_tensor *b,
                                      const lk_tensor *dst_full, int64_t row0, int64_t M_total,
                                      lk_buf *const *peer_dst, int32_t npeers);
/* Sharded SwiGLU half of LlamaMLP.forward (model/LlamaModel.kt:159-176) in ONE launch, N == 1: this rank owns rows
 * [row0, row0 + gate_shard.ne[1]) of the M_total-row gate and up weights (same type and shape).  With
 * x' = RMSNorm(x, norm_w, eps) when norm_w != NULL (x' is also stored to normed_dst when given), else x' = x:
 *   g = gate x', u = up x', silu_dst (optional) = silu(g), h = silu(g) * u
 * written to rows [row0, ...) of g_full / u_full / silu_full / h_full (ne = [1, M_total]) on this rank, and h also into
 * the same offsets of every peer's copy of h_full's buffer (peer_h, npeers as lk_compute_matmul_sharded) from the
 * GEMV epilogue over NVLink; follow with lk_peer_barrier.  Same roundings as lk_compute_matmul_fused.            */
LK_API lk_status   lk_compute_swiglu_sharded(lk_ctx *ctx, const lk_tensor *gate_shard, const lk_tensor *up_shard,
                                      const lk_tensor *x, const lk_tensor *norm_w, float eps, const lk_tensor *normed_dst,
                                      const lk_tensor *g_full, const lk_tensor *u_full, const lk_tensor *silu_full,
                                      const lk_tensor *h_full, int64_t row0, int64_t M_total,
                                      lk_buf *const *peer_h, int32_t npeers);

/* The exchange folded into the sharded launches themselves (no lk_peer_barrier launch between producer and consumer).
 * An exchange point ("slot", 0 .. LK_MAX_EXCHANGE_SLOTS-1) is a 64-bit arrival counter inside every rank's flag buffer
 * (LK_FLAG_BYTES bytes, zeroed before first use, peer-mapped like for lk_peer_barrier; the barrier's 16 words are its
 * first 64 bytes, so one buffer serves both):
 *   signal_slot >= 0: every CTA of this launch, after its peer stores and a system-scope fence, arrives on that slot of
 *                     EVERY rank's flag buffer (a weighted add; one launch adds up to 2^20 per rank whatever its grid);
 *   wait_slot   >= 0: before the launch reads its activations it waits until that slot of THIS rank's flag buffer has
 *                     received a whole launch's arrivals from every rank (counters only grow, the epoch of a launch is
 *                     counted on the device, so a captured sequence can be replayed).
 * One producer launch and one consumer launch per use of a slot on every rank, in the same order on every rank; a vector
 * that peers write may be rewritten only by a launch that comes after a wait on a LATER exchange of the same ranks
 * (then every reader is known to have finished), which the FFN pattern [SwiGLU -> slot 0 -> down -> slot 1 -> next SwiGLU]
 * satisfies.  The wait is bounded (~2 s); a time-out is counted and makes the next synchronising call return LK_ECUDA. */
#define LK_FLAG_BYTES 512
#define LK_MAX_EXCHANGE_SLOTS 8
typedef struct lk_shard_sync {
    lk_buf *const *peer_flags;   /* flag buffer of every rank, entry `rank` = this rank's own */
    int32_t nranks, rank;
    int32_t wait_slot;           /* -1: no wait */
    int32_t signal_slot;         /* -1: no signal */
} lk_shard_sync;
LK_API lk_status   lk_compute_matmul_sharded_sync(lk_ctx *ctx, const lk_tensor *a_shard, const lk_tensor *b,
                                      const lk_tensor *dst_full, int64_t row0, int64_t M_total,
                                      lk_buf *const *peer_dst, int32_t npeers, const lk_shard_sync *sync);
LK_API lk_status   lk_compute_swiglu_sharded_sync(lk_ctx *ctx, const lk_tensor *gate_shard, const lk_tensor *up_shard,
                                      const lk_tensor *x, const lk_tensor *norm_w, float eps, const lk_tensor *normed_dst,
                                      const lk_tensor *g_full, const lk_tensor *u_full, const lk_tensor *silu_full,
                                      const lk_tensor *h_full, int64_t row0, int64_t M_total,
                                      lk_buf *const *peer_h, int32_t npeers, const lk_shard_sync *sync);

#ifdef __cplusplus
}
#endif
#endif /* LLAMAKOTLIN_CUDA_H */

// lk_shim.cu — the C ABI (include/llamakotlin_cuda.h): context, device buffers,
// validation in the reference's order, host-tensor staging, op dispatch,
// op-sequence capture.  No CPU compute path exists here: if CUDA is not
// available every compute entry point fails with LK_ECUDA.
#include "lk_shim.h"

#include <string.h>

#include "lk_common.cuh"

static thread_local std::string g_noctx_err = "no error";

namespace lk {
lk_status fail(lk_ctx *ctx, lk_status st, const std::string &msg) {
    if (ctx) ctx->err = msg; else g_noctx_err = msg;
    return st;
}
lk_status cuda_fail(lk_ctx *ctx, cudaError_t e, const char *what) {
    cudaGetLastError();  // clear sticky-less error state
    return fail(ctx, e == cudaErrorMemoryAllocation ? LK_ENOMEM : LK_ECUDA,
                std::string("CUDA error: ") + cudaGetErrorString(e) + " at " + what);
}
// a non-GEMV op (elementwise kernel, copy): it starts only after everything earlier has completed (normal stream
// order, or its own griddepcontrol.wait before it releases its successor); remember what it wrote, because the next
// GEMV may start its weight prefetch while this op is still running
void chain_other(lk_ctx *ctx, const void *dst, size_t bytes) {
    ctx->chain_reads.clear();
    ctx->chain_writes.clear();
    typedef std::pair<uintptr_t, uintptr_t> Range;
    if (ctx->other_writes.size() > 64) { ctx->other_writes.clear(); ctx->other_writes.push_back(Range(0, ~(uintptr_t)0)); }
    ctx->other_writes.push_back(Range((uintptr_t)dst, (uintptr_t)dst + bytes));
}
lk_status check_async_faults(lk_ctx *ctx) {
    if (ctx->capturing) return LK_OK;
    if (ctx->chain_unchecked) {
        ctx->chain_unchecked = 0;
        unsigned w[4] = {0, 0, 0, 0};
        if (cudaMemcpy(w, ctx->chain_bar, 16, cudaMemcpyDeviceToHost) != cudaSuccess) return cuda_fail(ctx, cudaGetLastError(), "reading the chain barrier words");
        if (w[2]) {
            cudaMemset(ctx->chain_bar, 0, lk::CHAIN_SCRATCH_BYTES);   // a timed-out barrier leaves its arrival count behind: re-arm (K1m: and the mailbox epochs)
            ctx->chain_timeouts_seen += w[2];
            return fail(ctx, LK_ECUDA, "chained GEMV: the grid barrier timed out (" + std::to_string(w[2]) +
                        " CTAs gave up): not all CTAs were resident at once; the results of that launch are invalid");
        }
    }
    if (ctx->barrier_unchecked && ctx->barrier_flags) {
        ctx->barrier_unchecked = 0;
        uint32_t w = 0, w2 = 0;
        if (cudaMemcpy(&w, ctx->barrier_flags + 14, 4, cudaMemcpyDeviceToHost) != cudaSuccess) return cuda_fail(ctx, cudaGetLastError(), "reading the peer barrier words");
        if (ctx->barrier_flags_bytes >= LK_FLAG_BYTES && cudaMemcpy(&w2, ctx->barrier_flags + 64, 4, cudaMemcpyDeviceToHost) != cudaSuccess)
            return cuda_fail(ctx, cudaGetLastError(), "reading the exchange time-out word");
        if (w || w2) {
            if (w) cudaMemset(ctx->barrier_flags + 14, 0, 4);
            if (w2) cudaMemset(ctx->barrier_flags + 64, 0, 4);
            return fail(ctx, LK_ECUDA, "cross-GPU exchange: a peer did not arrive within the bounded wait; gathered outputs are invalid");
        }
    }
    return LK_OK;
}
// ---- programmatic-dependent-launch chain bookkeeping ---------------------------------------
// A GEMV launched with the PDL attribute may start while earlier GEMVs of the chain still run.
// It may skip the wait on its predecessor (GEMV_WAIT_END, the kernel then waits at its very end so
// completion order stays launch order) only if it touches nothing an in-flight launch wrote and
// writes nothing an in-flight launch read or wrote.  Any other op on the stream is a normal kernel
// or copy that serialises fully, which closes the chain.
namespace {
typedef std::pair<uintptr_t, uintptr_t> Range;
bool overlaps(const std::vector<Range> &v, Range r) {
    for (const Range &q : v)
        if (r.first < q.second && q.first < r.second) return true;
    return false;
}
void chain_close(lk_ctx *ctx) { ctx->chain_reads.clear(); ctx->chain_writes.clear(); }

}  // namespace
int gemv_wait_mode(lk_ctx *ctx, const void *a, size_t abytes, const void *b, size_t bbytes, const void *d, size_t dbytes,
                   bool pdl, const std::vector<std::pair<const void *, size_t>> *more_reads,
                   const std::vector<std::pair<const void *, size_t>> *more_writes,
                   const std::vector<std::pair<const void *, size_t>> *more_weights) {
    const Range ra((uintptr_t)a, (uintptr_t)a + abytes);
    std::vector<Range> wt;          // weights: read before the wait point (prefetch)
    wt.push_back(ra);
    if (more_weights) for (const auto &r : *more_weights) if (r.first && r.second) wt.push_back(Range((uintptr_t)r.first, (uintptr_t)r.first + r.second));
    std::vector<Range> rd, wr;      // what the launch reads after its wait point / writes
    rd.push_back(Range((uintptr_t)b, (uintptr_t)b + bbytes));
    wr.push_back(Range((uintptr_t)d, (uintptr_t)d + dbytes));
    if (more_reads) for (const auto &r : *more_reads) if (r.first && r.second) rd.push_back(Range((uintptr_t)r.first, (uintptr_t)r.first + r.second));
    if (more_writes) for (const auto &r : *more_writes) if (r.first && r.second) wr.push_back(Range((uintptr_t)r.first, (uintptr_t)r.first + r.second));
    int mode = 2;   // GEMV_WAIT_END
    if (!pdl || !ctx->overlap_mode) mode = 1;
    for (const Range &r : rd) if (overlaps(ctx->chain_writes, r)) mode = 1;
    for (const Range &r : wr) if (overlaps(ctx->chain_writes, r) || overlaps(ctx->chain_reads, r)) mode = 1;
    for (const Range &r : wt) if (overlaps(ctx->chain_writes, r)) mode = 0;
    if (!ctx->other_writes.empty()) {   // first GEMV after a normal op: it can start before that op has finished
        bool hit = false;
        for (const Range &r : wt) hit = hit || overlaps(ctx->other_writes, r);
        mode = hit ? 0 : (mode == 0 ? 0 : 1);
        ctx->other_writes.clear();
    }
    if (mode != 2) chain_close(ctx);   // everything earlier has completed once this launch passes its wait
    if (ctx->chain_reads.size() > 256) { chain_close(ctx); mode = 1; }
    for (const Range &r : wt) ctx->chain_reads.push_back(r);
    for (const Range &r : rd) ctx->chain_reads.push_back(r);
    for (const Range &r : wr) ctx->chain_writes.push_back(r);
    return mode;
}

}  // namespace lk
using lk::fail;
using lk::chain_other;

extern "C" {

int32_t lk_abi_version(void) { return LK_ABI_VERSION; }

int32_t lk_type_block_bytes(int32_t type) {
    if (type == LK_TYPE_F32 || type == LK_TYPE_I32) return 4;
    if (type == LK_TYPE_F16) return 2;
    if (lk::kq_type_block_bytes(type)) return lk::kq_type_block_bytes(type);
    return lk::block_bytes(type);
}

static bool is_block_type(int t) { return t == LK_TYPE_Q4_0 || t == LK_TYPE_Q4_1 || t == LK_TYPE_Q8_0; }
// N4: Q2_K / Q4_K / Q8_K / BITNET_1_58 (lk_kquants.cu)
static bool is_kq_type(int t) { return lk::kq_type_block_bytes(t) != 0; }

static int64_t num_elements(const lk_tensor *t) {
    int64_t n = 1;
    for (int d = 0; d < LK_MAX_DIMS; d++) n *= (t->ne[d] > 0 ? t->ne[d] : (t->ne[d] == 0 && d == 0 ? 0 : 1));
    return n;
}

int64_t lk_tensor_nbytes(const lk_tensor *t) {
    if (!t) return -1;
    const int64_t n = num_elements(t);
    if (is_block_type(t->type)) return n / LK_QK * lk::block_bytes(t->type);
    if (is_kq_type(t->type)) return n / lk::kq_type_block_elems(t->type) * lk::kq_type_block_bytes(t->type);
    if (t->type == LK_TYPE_F32 || t->type == LK_TYPE_I32) return n * 4;
    if (t->type == LK_TYPE_F16) return n * 2;
    return -1;
}

lk_status lk_ctx_create(int32_t device, lk_ctx **out) {
    if (!out) return fail(nullptr, LK_EINVAL_ARG, "lk_ctx_create: out is NULL");
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(nullptr, LK_ECUDA, "lk_ctx_create: no CUDA device (this library has no CPU fallback)");
    }
    if (device < 0 || device >= ndev) return fail(nullptr, LK_EINVAL_ARG, "lk_ctx_create: bad device index");
    lk_ctx *ctx = new lk_ctx();
    ctx->device = device;
    LK_CUDA(nullptr, cudaSetDevice(device));
    cudaDeviceProp prop;
    LK_CUDA(nullptr, cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) {
        delete ctx;
        return fail(nullptr, LK_ECUDA, "lk_ctx_create: kernels are built for sm_100a only");
    }
    ctx->sm_count = prop.multiProcessorCount;
    LK_CUDA(nullptr, cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    LK_CUDA(nullptr, cudaEventCreate(&ctx->ev0));
    LK_CUDA(nullptr, cudaEventCreate(&ctx->ev1));
    // copy streams of lk_buf_upload_async / lk_buf_download_async: made here so that a capture can record its fork event
    LK_CUDA(nullptr, cudaStreamCreateWithFlags(&ctx->copy_in, cudaStreamNonBlocking));
    LK_CUDA(nullptr, cudaStreamCreateWithFlags(&ctx->copy_out, cudaStreamNonBlocking));
    LK_CUDA(nullptr, cudaEventCreateWithFlags(&ctx->ev_in, cudaEventDisableTiming));
    LK_CUDA(nullptr, cudaEventCreateWithFlags(&ctx->ev_out, cudaEventDisableTiming));
    LK_CUDA(nullptr, cudaEventCreateWithFlags(&ctx->ev_root, cudaEventDisableTiming));
    // grid-barrier words of the chained GEMV (arrivals, generation, time-outs); allocated here because nothing may be
    // allocated-and-cleared while a launch sequence is being captured
    // ... followed, 256 bytes on, by K1m's mailbox (activations handed from phase to phase, lk_gemv_mma.cu)
    LK_CUDA(nullptr, cudaMalloc((void **)&ctx->chain_bar, lk::CHAIN_SCRATCH_BYTES));
    LK_CUDA(nullptr, cudaMemset(ctx->chain_bar, 0, lk::CHAIN_SCRATCH_BYTES));
    *out = ctx;
    return LK_OK;
}

lk_status lk_comm_destroy(lk_ctx *ctx);

lk_status lk_ctx_destroy(lk_ctx *ctx) {
    if (!ctx) return LK_OK;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    if (ctx->copy_in) { cudaStreamSynchronize(ctx->copy_in); cudaStreamDestroy(ctx->copy_in); }
    if (ctx->copy_out) { cudaStreamSynchronize(ctx->copy_out); cudaStreamDestroy(ctx->copy_out); }
    if (ctx->ev_in) cudaEventDestroy(ctx->ev_in);
    if (ctx->ev_out) cudaEventDestroy(ctx->ev_out);
    if (ctx->ev_root) cudaEventDestroy(ctx->ev_root);
    lk_comm_destroy(ctx);
    if (ctx->arena) cudaFree(ctx->arena);
    if (ctx->chain_bar) cudaFree(ctx->chain_bar);
    if (ctx->gemm_ws) cudaFree(ctx->gemm_ws);
    for (void *w : ctx->gemm_ws_retired) cudaFree(w);
    if (ctx->attn_ws) cudaFree(ctx->attn_ws);
    for (void *w : ctx->attn_ws_retired) cudaFree(w);
    cudaEventDestroy(ctx->ev0);
    cudaEventDestroy(ctx->ev1);
    cudaStreamDestroy(ctx->stream);
    delete ctx;
    return LK_OK;
}

lk_status lk_sync(lk_ctx *ctx) {
    if (!ctx) return fail(nullptr, LK_EINVAL_ARG, "lk_sync: ctx is NULL");
    LK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (ctx->copies_pending && !ctx->capturing) {     // the asynchronous copies run beside the compute stream
        if (ctx->copy_in) LK_CUDA(ctx, cudaStreamSynchronize(ctx->copy_in));
        if (ctx->copy_out) LK_CUDA(ctx, cudaStreamSynchronize(ctx->copy_out));
        ctx->copies_pending = false;
    }
    return lk::check_async_faults(ctx);
}

const char *lk_last_error(lk_ctx *ctx) { return ctx ? ctx->err.c_str() : g_noctx_err.c_str(); }
void *lk_ctx_stream(lk_ctx *ctx) { return ctx ? (void *)ctx->stream : nullptr; }
int32_t lk_ctx_device(lk_ctx *ctx) { return ctx ? ctx->device : -1; }
int32_t lk_ctx_sm_count(lk_ctx *ctx) { return ctx ? ctx->sm_count : 0; }
int64_t lk_ctx_launch_count(lk_ctx *ctx) { return ctx ? ctx->launches : 0; }

lk_status lk_timer_start(lk_ctx *ctx) {
    if (!ctx) return fail(nullptr, LK_EINVAL_ARG, "lk_timer_start: ctx is NULL");
    LK_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    return LK_OK;
}
lk_status lk_timer_stop(lk_ctx *ctx, float *ms_out) {
    if (!ctx || !ms_out) return fail(ctx, LK_EINVAL_ARG, "lk_timer_stop: NULL argument");
    LK_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    LK_CUDA(ctx, cudaEventSynchronize(ctx->ev1));
    LK_CUDA(ctx, cudaEventElapsedTime(ms_out, ctx->ev0, ctx->ev1));
    if (cudaStreamQuery(ctx->stream) == cudaSuccess) return lk::check_async_faults(ctx);
    return LK_OK;
}

// ---- device buffers ---------------------------------------------------------
lk_status lk_buf_alloc(lk_ctx *ctx, uint64_t size, lk_buf **out) {
    if (!ctx || !out) return fail(ctx, LK_EINVAL_ARG, "lk_buf_alloc: NULL argument");
    *out = nullptr;
    LK_CUDA(ctx, cudaSetDevice(ctx->device));
    void *p = nullptr;
    LK_CUDA(ctx, cudaMalloc(&p, size ? size : 16));
    lk_buf *b = new lk_buf();
    b->ctx = ctx; b->ptr = p; b->size = size;
    *out = b;
    return LK_OK;
}

lk_status lk_buf_free(lk_buf *buf) {
    if (!buf) return LK_OK;
    lk_ctx *ctx = buf->ctx;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    if (buf->imported) cudaIpcCloseMemHandle(buf->ptr); else cudaFree(buf->ptr);
    delete buf;
    return LK_OK;
}

uint64_t lk_buf_size(const lk_buf *buf) { return buf ? buf->size : 0; }
void *lk_buf_device_ptr(lk_buf *buf) { return buf ? buf->ptr : nullptr; }

static lk_status buf_range(lk_buf *buf, uint64_t off, uint64_t n, const char *what) {
    if (!buf) return fail(nullptr, LK_EINVAL_ARG, std::string(what) + ": buf is NULL");
    if (off > buf->size || n > buf->size - off)
        return fail(buf->ctx, LK_EBOUNDS, std::string(what) + ": range outside the buffer");
    return LK_OK;
}

lk_status lk_buf_upload(lk_buf *buf, uint64_t off, const void *host, uint64_t n) {
    lk_status st = buf_range(buf, off, n, "lk_buf_upload");
    if (st != LK_OK) return st;
    if (n == 0) return LK_OK;
    if (!host) return fail(buf->ctx, LK_EINVAL_ARG, "lk_buf_upload: host is NULL");
    lk_ctx *ctx = buf->ctx;
    LK_CUDA(ctx, cudaMemcpyAsync((uint8_t *)buf->ptr + off, host, n, cudaMemcpyHostToDevice, ctx->stream));
    chain_other(ctx, (uint8_t *)buf->ptr + off, n);
    LK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // host memory may be unpinned right after return
    return LK_OK;
}

lk_status lk_buf_download(lk_buf *buf, uint64_t off, void *host, uint64_t n) {
    lk_status st = buf_range(buf, off, n, "lk_buf_download");
    if (st != LK_OK) return st;
    if (n == 0) return LK_OK;
    if (!host) return fail(buf->ctx, LK_EINVAL_ARG, "lk_buf_download: host is NULL");
    lk_ctx *ctx = buf->ctx;
    LK_CUDA(ctx, cudaMemcpyAsync(host, (uint8_t *)buf->ptr + off, n, cudaMemcpyDeviceToHost, ctx->stream));
    chain_other(ctx, nullptr, 0);
    LK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return lk::check_async_faults(ctx);
}

// ---- asynchronous host copies (pipelined host I/O of a step, include/llamakotlin_cuda.h) ----
lk_status lk_buf_upload_async(lk_buf *buf, uint64_t off, const void *host, uint64_t n) {
    lk_status st = buf_range(buf, off, n, "lk_buf_upload_async");
    if (st != LK_OK) return st;
    if (n == 0) return LK_OK;
    if (!host) return fail(buf->ctx, LK_EINVAL_ARG, "lk_buf_upload_async: host is NULL");
    lk_ctx *ctx = buf->ctx;
    if (ctx->capturing && !ctx->cap_in) {     // fork from the root of the capture: the copy depends on nothing recorded so far
        LK_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_in, ctx->ev_root, 0));
        ctx->cap_in = true;
    }
    LK_CUDA(ctx, cudaMemcpyAsync((uint8_t *)buf->ptr + off, host, n, cudaMemcpyHostToDevice, ctx->copy_in));
    LK_CUDA(ctx, cudaEventRecord(ctx->ev_in, ctx->copy_in));
    LK_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_in, 0));   // a full dependency of whatever is enqueued next
    ctx->copies_pending = true;
    return LK_OK;
}

lk_status lk_buf_download_async(lk_buf *buf, uint64_t off, void *host, uint64_t n) {
    lk_status st = buf_range(buf, off, n, "lk_buf_download_async");
    if (st != LK_OK) return st;
    if (n == 0) return LK_OK;
    if (!host) return fail(buf->ctx, LK_EINVAL_ARG, "lk_buf_download_async: host is NULL");
    lk_ctx *ctx = buf->ctx;
    LK_CUDA(ctx, cudaEventRecord(ctx->ev_out, ctx->stream));          // everything enqueued on the compute stream so far
    LK_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_out, ctx->ev_out, 0));
    if (ctx->capturing) ctx->cap_out = true;
    LK_CUDA(ctx, cudaMemcpyAsync(host, (uint8_t *)buf->ptr + off, n, cudaMemcpyDeviceToHost, ctx->copy_out));
    ctx->copies_pending = true;
    return LK_OK;
}

lk_status lk_buf_copy(lk_buf *dst, uint64_t doff, lk_buf *src, uint64_t soff, uint64_t n) {
    lk_status st = buf_range(dst, doff, n, "lk_buf_copy(dst)");
    if (st != LK_OK) return st;
    st = buf_range(src, soff, n, "lk_buf_copy(src)");
    if (st != LK_OK) return st;
    lk_ctx *ctx = dst->ctx;
    if (n) LK_CUDA(ctx, cudaMemcpyAsync((uint8_t *)dst->ptr + doff, (uint8_t *)src->ptr + soff, n, cudaMemcpyDeviceToDevice, ctx->stream));
    chain_other(ctx, (uint8_t *)dst->ptr + doff, n);
    return LK_OK;
}

lk_status lk_buf_clear(lk_buf *buf, uint8_t value) {
    if (!buf) return fail(nullptr, LK_EINVAL_ARG, "lk_buf_clear: buf is NULL");
    lk_ctx *ctx = buf->ctx;
    if (buf->size) LK_CUDA(ctx, cudaMemsetAsync(buf->ptr, value, buf->size, ctx->stream));
    chain_other(ctx, buf->ptr, buf->size);
    return LK_OK;
}

lk_status lk_host_alloc(lk_ctx *ctx, uint64_t size, void **out) {
    if (!ctx || !out) return fail(ctx, LK_EINVAL_ARG, "lk_host_alloc: NULL argument");
    LK_CUDA(ctx, cudaSetDevice(ctx->device));
    LK_CUDA(ctx, cudaHostAlloc(out, size ? size : 16, cudaHostAllocDefault));
    return LK_OK;
}
lk_status lk_host_free(lk_ctx *ctx, void *p) {
    if (p) LK_CUDA(ctx, cudaFreeHost(p));
    return LK_OK;
}

}  // extern "C"

// ---- tensor resolution: host tensors are staged through a device arena -------
namespace {

struct Staged {
    void *dev = nullptr;       // device address the kernel uses (contiguous)
    void *host = nullptr;      // non-NULL: copy back to here after the op (outputs only)
    int64_t bytes = 0;         // bytes copied back to `host`
    // strided F32 output (A9): the kernel wrote `dev` in contiguous order; scatter it through the strides into
    // scatter_base (the tensor's first byte on the device: its own buffer, or the staged copy of its host byte range)
    void *scatter_base = nullptr;
    int64_t ne[4] = {1, 1, 1, 1};
    uint64_t nb[4] = {0, 0, 0, 0};
};

struct OpScope {
    lk_ctx *ctx;
    bool any_host = false;
    std::vector<Staged> outs;
    explicit OpScope(lk_ctx *c) : ctx(c) { c->arena_used = 0; }
};

lk_status arena_reserve(lk_ctx *ctx, size_t need) {
    if (need <= ctx->arena_size) return LK_OK;
    LK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (ctx->arena) cudaFree(ctx->arena);
    ctx->arena = nullptr; ctx->arena_size = 0;
    size_t sz = need + need / 4 + (1 << 20);
    LK_CUDA(ctx, cudaMalloc((void **)&ctx->arena, sz));
    ctx->arena_size = sz;
    return LK_OK;
}

// K2 workspace (FP16 activation tiles).  It may have to grow inside a capture: cudaMalloc is legal there only
// in relaxed capture mode, and the old block stays alive because an earlier captured graph may name it.
lk_status gemm_ws_reserve(lk_ctx *ctx, size_t need) {
    if (need <= ctx->gemm_ws_size) return LK_OK;
    cudaStreamCaptureMode mode = cudaStreamCaptureModeRelaxed;
    LK_CUDA(ctx, cudaThreadExchangeStreamCaptureMode(&mode));
    void *w = nullptr;
    const size_t sz = need + need / 4 + 4096;      // cudaMalloc returns >= 256-byte alignment; the tiles need 1024
    cudaError_t e = cudaMalloc(&w, sz);
    cudaThreadExchangeStreamCaptureMode(&mode);
    if (e != cudaSuccess) return lk::cuda_fail(ctx, e, "workspace allocation");
    if (ctx->gemm_ws) ctx->gemm_ws_retired.push_back(ctx->gemm_ws);
    ctx->gemm_ws = w;
    ctx->gemm_ws_size = sz;
    return LK_OK;
}

// decode-attention workspace; grows like gemm_ws.  The tickets must be zero before first use: the memset is ordered on
// the ctx stream (inside a capture it becomes a node of the graph, harmless on replay: tickets are zero between launches)
lk_status attn_ws_reserve(lk_ctx *ctx, size_t need) {
    if (need <= ctx->attn_ws_size) return LK_OK;
    cudaStreamCaptureMode mode = cudaStreamCaptureModeRelaxed;
    LK_CUDA(ctx, cudaThreadExchangeStreamCaptureMode(&mode));
    void *w = nullptr;
    const size_t sz = need + need / 4 + 4096;
    cudaError_t e = cudaMalloc(&w, sz);
    cudaThreadExchangeStreamCaptureMode(&mode);
    if (e != cudaSuccess) return lk::cuda_fail(ctx, e, "attention workspace allocation");
    if (ctx->attn_ws) ctx->attn_ws_retired.push_back(ctx->attn_ws);
    ctx->attn_ws = w;
    ctx->attn_ws_size = sz;
    LK_CUDA(ctx, cudaMemsetAsync(w, 0, sz, ctx->stream));
    return LK_OK;
}

bool contiguous_f32(const lk_tensor *t) {
    bool allzero = true;
    for (int d = 0; d < LK_MAX_DIMS; d++) allzero = allzero && t->nb[d] == 0;
    if (allzero) return true;
    if (t->nb[0] != 4) return false;
    uint64_t expect = 4;
    for (int d = 1; d < LK_MAX_DIMS; d++) {
        expect *= (uint64_t)(t->ne[d - 1] > 0 ? t->ne[d - 1] : 1);
        if ((t->ne[d] > 1) && t->nb[d] != expect) return false;
    }
    return true;
}
bool strided_f32(const lk_tensor *t) { return t && (t->type == LK_TYPE_F32 || t->type == LK_TYPE_I32) && !contiguous_f32(t); }
// bytes from the tensor's first byte to one past its last element when addressed through nb[] (getElementByteOffset,
// core/GGMLTypes.kt:328-357); strides must be multiples of 4 (aligned F32 accesses on the device)
int64_t strided_extent(const lk_tensor *t) {
    int64_t ext = 4;
    for (int d = 0; d < LK_MAX_DIMS; d++) {
        const int64_t n = t->ne[d] > 0 ? t->ne[d] : 1;
        if (t->nb[d] % 4) return -1;
        ext += (n - 1) * (int64_t)t->nb[d];
    }
    return ext;
}

// total staging bytes the host tensors of one op need (each rounded to 256); a strided F32 tensor also needs its
// contiguous copy (and, when host resident, the copy of the byte range it spans)
size_t host_need(std::initializer_list<const lk_tensor *> ts) {
    size_t need = 0;
    for (const lk_tensor *t : ts) {
        if (!t) continue;
        if (strided_f32(t)) {
            need += ((size_t)lk_tensor_nbytes(t) + 255) & ~(size_t)255;
            if (!t->dev_buf) { const int64_t e = strided_extent(t); if (e > 0) need += ((size_t)e + 255) & ~(size_t)255; }
        } else if (!t->dev_buf) {
            need += ((size_t)lk_tensor_nbytes(t) + 255) & ~(size_t)255;
        }
    }
    return need;
}

lk_status resolve(OpScope &sc, const lk_tensor *t, bool is_output, void **dev_out, const char *name) {
    lk_ctx *ctx = sc.ctx;
    const int64_t nbytes = lk_tensor_nbytes(t);
    if (nbytes < 0) return fail(ctx, LK_ENOTIMPL, std::string(name) + ": unsupported tensor type");
    if ((t->dev_buf != nullptr) == (t->host_base != nullptr))
        return fail(ctx, LK_EINVAL_ARG, std::string(name) + ": exactly one of host_base / dev_buf must be set");
    const bool strided = strided_f32(t);
    const int64_t span = strided ? strided_extent(t) : nbytes;      // bytes of the buffer the tensor touches
    if (strided && span < 0) return fail(ctx, LK_EINVAL_ARG, std::string(name) + ": F32 strides must be multiples of 4 bytes");
    const uint64_t bsize = t->dev_buf ? t->dev_buf->size : t->buffer_size;
    if (t->data_offset > bsize || (uint64_t)span > bsize - t->data_offset)      // IndexOutOfBounds semantics of the accessors (core/GGMLTypes.kt:556-558)
        return fail(ctx, LK_EBOUNDS, std::string(name) + (t->dev_buf ? ": tensor bytes exceed the device buffer" : ": tensor bytes exceed the host buffer"));
    uint8_t *base;             // the tensor's first byte on the device
    if (t->dev_buf) {
        if (t->data_offset % (strided ? 4 : 16))  // GGMLDynTensorAllocator alignment, core/GGMLAlloc.kt:122,149
            return fail(ctx, LK_EINVAL_ARG, std::string(name) + ": dataOffset must be 16-byte aligned on device");
        base = (uint8_t *)t->dev_buf->ptr + t->data_offset;
        if (!strided) { *dev_out = base; return LK_OK; }
        if (ctx->capturing) return fail(ctx, LK_EINVAL_ARG, std::string(name) + ": strided tensors cannot be used inside a captured sequence");
    } else {
        if (ctx->capturing)
            return fail(ctx, LK_EINVAL_ARG, std::string(name) + ": host tensors cannot be used inside a captured sequence");
        sc.any_host = true;
        base = ctx->arena + ctx->arena_used;
        ctx->arena_used += ((size_t)span + 255) & ~(size_t)255;
        uint8_t *h = (uint8_t *)t->host_base + t->data_offset;
        // inputs are uploaded; so is the byte range of a strided output (the bytes between its elements must survive)
        if ((!is_output || strided) && span) {
            LK_CUDA(ctx, cudaMemcpyAsync(base, h, (size_t)span, cudaMemcpyHostToDevice, ctx->stream));
            chain_other(ctx, base, (size_t)span);
        }
        if (is_output) {
            Staged s; s.dev = base; s.host = h; s.bytes = span;
            if (!strided) { sc.outs.push_back(s); *dev_out = base; return LK_OK; }
        } else if (!strided) {
            *dev_out = base;
            return LK_OK;
        }
    }
    // strided F32 (A9): a contiguous copy in the arena is what the kernel sees
    uint8_t *c = ctx->arena + ctx->arena_used;
    ctx->arena_used += ((size_t)nbytes + 255) & ~(size_t)255;
    if (is_output) {
        Staged s; s.dev = c; s.scatter_base = base;
        for (int d = 0; d < LK_MAX_DIMS; d++) { s.ne[d] = t->ne[d] > 0 ? t->ne[d] : 1; s.nb[d] = t->nb[d]; }
        if (!t->dev_buf) { s.host = (uint8_t *)t->host_base + t->data_offset; s.bytes = span; }
        sc.outs.push_back(s);
    } else {
        int64_t ne[4];
        for (int d = 0; d < LK_MAX_DIMS; d++) ne[d] = t->ne[d] > 0 ? t->ne[d] : 1;
        const cudaError_t e = lk::launch_strided_copy(base, c, ne, t->nb, 0, ctx->stream);
        if (e != cudaSuccess) return lk::cuda_fail(ctx, e, "strided gather");
        chain_other(ctx, c, (size_t)nbytes);
    }
    *dev_out = c;
    return LK_OK;
}

// scatter strided outputs, copy staged outputs back and make dst valid on return (reference semantics)
lk_status finish(OpScope &sc) {
    lk_ctx *ctx = sc.ctx;
    for (const Staged &s : sc.outs) {
        if (s.scatter_base) {
            const cudaError_t e = lk::launch_strided_copy(s.dev, s.scatter_base, s.ne, s.nb, 1, ctx->stream);
            if (e != cudaSuccess) return lk::cuda_fail(ctx, e, "strided scatter");
            chain_other(ctx, nullptr, 0);
            if (s.host && s.bytes) LK_CUDA(ctx, cudaMemcpyAsync(s.host, s.scatter_base, (size_t)s.bytes, cudaMemcpyDeviceToHost, ctx->stream));
        } else if (s.bytes) {
            LK_CUDA(ctx, cudaMemcpyAsync(s.host, s.dev, (size_t)s.bytes, cudaMemcpyDeviceToHost, ctx->stream));
        }
    }
    if (sc.any_host) {
        LK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        chain_other(ctx, nullptr, 0);
        return lk::check_async_faults(ctx);
    }
    return LK_OK;
}

void count_launch(lk_ctx *ctx, int n = 1) {
    if (ctx->capturing) ctx->capture_launches += n; else ctx->launches += n;
}

// elementwise ops join the programmatic-dependent-launch chain under the same policy as the GEMV (gemv_pdl option:
// 1 = inside captured sequences, 2 = always); host-staged operands keep plain stream order
bool op_pdl(const lk_ctx *ctx, const OpScope &sc) {
    return !sc.any_host && ((ctx->pdl_mode == 2) || (ctx->pdl_mode == 1 && ctx->capturing));
}

bool same_shape(const lk_tensor *a, const lk_tensor *b) {
    for (int d = 0; d < LK_MAX_DIMS; d++) {
        const int64_t x = a->ne[d] > 0 ? a->ne[d] : (d == 0 ? 0 : 1), y = b->ne[d] > 0 ? b->ne[d] : (d == 0 ? 0 : 1);
        if (x != y) return false;
    }
    return true;
}

}  // namespace

extern "C" {

lk_status lk_compute_matmul(lk_ctx *ctx, const lk_tensor *a, const lk_tensor *b, const lk_tensor *dst) {
    if (!ctx || !a || !b || !dst) return fail(ctx, LK_EINVAL_ARG, "lk_compute_matmul: NULL argument");
    // validation in the order of core/GGMLComputeOps.kt:1436-1446
    const int64_t M = a->ne[1], K = a->ne[0], N = b->ne[0], Kb = b->ne[1];
    if (K != Kb) return fail(ctx, LK_EINVAL_DIM, "Dim mismatch K: a.ne[0] != b.ne[1]");
    if (dst->ne[0] != N || dst->ne[1] != M)
        return fail(ctx, LK_EINVAL_DIM, "Result tensor dimensions must match expected output size [N, M]");
    if ((is_block_type(a->type) || is_kq_type(a->type)) && b->type == LK_TYPE_F32) {
        if (dst->type != LK_TYPE_F32) return fail(ctx, LK_EINVAL_TYPE, "Result tensor type must be F32 for Q x F32 matmul");
    } else {
        // general fallback, :1530-1565
        if (dst->type != a->type) return fail(ctx, LK_EINVAL_TYPE, "Result tensor type must match first input type for general matmul");
        if (a->type != LK_TYPE_F32 || b->type != LK_TYPE_F32)
            return fail(ctx, LK_ENOTIMPL, "computeMatMul on this device path supports {F32,Q4_0,Q4_1,Q8_0} x F32 only");
    }
    if (M < 0 || K < 0 || N < 0) return fail(ctx, LK_EINVAL_DIM, "negative dimension");
    if (is_block_type(a->type) && ((M * K) % LK_QK) != 0)
        return fail(ctx, LK_EINVAL_DIM, "quantized tensor element count must be a multiple of 32");
    if (is_kq_type(a->type) && ((M * K) % lk::kq_type_block_elems(a->type)) != 0)   // getNumBlocks, core/GGMLTypes.kt:507-535
        return fail(ctx, LK_EINVAL_DIM, "quantized tensor element count must be a multiple of the block size");
    LK_CUDA(ctx, cudaSetDevice(ctx->device));

    OpScope sc(ctx);
    lk_status st = arena_reserve(ctx, host_need({a, b, dst}));
    if (st != LK_OK) return st;
    void *da, *db, *dd;
    if ((st = resolve(sc, a, false, &da, "a")) != LK_OK) return st;
    if ((st = resolve(sc, b, false, &db, "b")) != LK_OK) return st;
    if ((st = resolve(sc, dst, true, &dd, "dst")) != LK_OK) return st;
    if (M == 0 || N == 0) return finish(sc);

    if (is_kq_type(a->type)) {     // core/GGMLComputeOps.kt:1481-1513 (+ computeDotProductBitNet158F32, :635-660)
        if (N == 1 && K > 0 && (a->type == LK_TYPE_BITNET_1_58 ? K % 64 == 0 : K % 256 == 0)) {
            // whole super-blocks per row: the streaming DP4A GEMV (K1) takes the three K-quant formats and BitNet 1.58 as well
            lk::GemvTuning tune = ctx->tune;
            tune.pdl = (ctx->pdl_mode == 2) || (ctx->pdl_mode == 1 && ctx->capturing);
            lk::GemvPeers none;
            tune.wait_mode = lk::gemv_wait_mode(ctx, da, (size_t)lk_tensor_nbytes(a), db, (size_t)K * 4, dd, (size_t)M * 4, tune.pdl != 0);
            if (sc.any_host) tune.wait_mode = 0;
            const cudaError_t eg = lk::launch_gemv(a->type, da, K, M, (const float *)db, (float *)dd, 0, none, tune, ctx->sm_count, ctx->stream);
            if (eg == cudaSuccess) { count_launch(ctx); return finish(sc); }
            if (eg != cudaErrorNotSupported) return lk::cuda_fail(ctx, eg, "K-quant GEMV launch");
            cudaGetLastError();
        }
        chain_other(ctx, dd, (size_t)M * N * 4);
        const cudaError_t ek = lk::launch_kq_matmul(a->type, da, K, M, (const float *)db, N, (float *)dd, ctx->sm_count, ctx->stream);
        if (ek != cudaSuccess) return lk::cuda_fail(ctx, ek, "K-quant matmul launch");
        count_launch(ctx);
        return finish(sc);
    }
    cudaError_t e = cudaErrorNotSupported;
    if (N == 1 && is_block_type(a->type)) {
        lk::GemvTuning tune = ctx->tune;
        tune.pdl = (ctx->pdl_mode == 2) || (ctx->pdl_mode == 1 && ctx->capturing);
        lk::GemvPeers none;
        // a weight that a preceding normal op wrote is covered by that op's full serialisation only if this
        // launch waits before its prefetch; the first GEMV after a non-GEMV op therefore never uses mode END
        tune.wait_mode = lk::gemv_wait_mode(ctx, da, (size_t)lk_tensor_nbytes(a), db, (size_t)K * 4, dd, (size_t)M * 4, tune.pdl != 0);
        if (sc.any_host) tune.wait_mode = 0;
        e = lk::launch_gemv(a->type, da, K, M, (const float *)db, (float *)dd, 0, none, tune, ctx->sm_count, ctx->stream);
    }
    int launched = 1;
    // Batched case (N >= 2): the tcgen05 GEMM whenever the packed rows are TMA-addressable (K % 256 == 0: 16-byte row pitch)
    // and the product is big enough to pay for its three launches; small N is padded to one 32-column tile, so the
    // weights are still streamed and dequantised exactly once.  Everything else: the generic F32 kernel.
    if (N > 1 && is_block_type(a->type) && ctx->gemm_tc_min_n > 0 && N >= ctx->gemm_tc_min_n && K > 0 && K % 256 == 0 &&
        (double)M * (double)K * (double)N >= (double)ctx->gemm_tc_min_work) {
        if ((st = gemm_ws_reserve(ctx, lk::gemm_tc_workspace_bytes(K, N))) != LK_OK) return st;
        chain_other(ctx, dd, (size_t)M * N * 4);
        void *ws = (void *)(((uintptr_t)ctx->gemm_ws + 1023) & ~(uintptr_t)1023);
        e = lk::launch_gemm_tc(a->type, da, K, M, (const float *)db, N, (float *)dd, ws, ctx->sm_count, ctx->gemm_tc_cluster, ctx->stream, ctx->tune.nibble_order, ctx->gemm_tc_precision);
        if (e == cudaSuccess) { launched = 3; ctx->gemm_tc_launches++; }   // colmax_kernel + pack_b_kernel + gemm_tc_kernel
    }
    if (e == cudaErrorNotSupported) {
        cudaGetLastError();
        chain_other(ctx, dd, (size_t)M * N * 4);
        e = lk::launch_matmul_generic(a->type, da, K, M, (const float *)db, N, (float *)dd, ctx->stream, ctx->tune.nibble_order);
    }
    if (e != cudaSuccess) return lk::cuda_fail(ctx, e, "matmul launch");
    count_launch(ctx, launched);
    return finish(sc);
}

// One fused GEMV, validated and resolved to device pointers: what lk_compute_matmul_fused launches on its own and
// lk_compute_matmul_chain launches as one phase of a chain.
namespace {
struct FusedLaunch {
    const void *a = nullptr;
    int64_t K = 0, M = 0;
    size_t abytes = 0;
    const float *x = nullptr;
    float *out = nullptr;
    int type = 0;
    lk::GemvFusion fu;
    std::vector<std::pair<const void *, size_t>> rd, wr, wt;   // further reads / writes / weights (bookkeeping)
};
}  // namespace

static lk_status fused_prepare(lk_ctx *ctx, const lk_tensor *a, const lk_tensor *b, const lk_tensor *dst,
                               const lk_matmul_fusion *f, FusedLaunch *L) {
    if (!ctx || !a || !b || !dst || !f) return fail(ctx, LK_EINVAL_ARG, "lk_compute_matmul_fused: NULL argument");
    const int64_t M = a->ne[1], K = a->ne[0], N = b->ne[0];
    if (K != b->ne[1]) return fail(ctx, LK_EINVAL_DIM, "Dim mismatch K: a.ne[0] != b.ne[1]");
    if (dst->ne[0] != N || dst->ne[1] != M)
        return fail(ctx, LK_EINVAL_DIM, "Result tensor dimensions must match expected output size [N, M]");
    if (!is_block_type(a->type) || b->type != LK_TYPE_F32 || dst->type != LK_TYPE_F32)
        return fail(ctx, LK_ENOTIMPL, "fused matmul: Q4_0/Q4_1/Q8_0 x F32 -> F32 only");
    if (N != 1 || M <= 0 || K <= 0) return fail(ctx, LK_ENOTIMPL, "fused matmul: N == 1 only");
    if (f->n_extra < 0 || f->n_extra > 2) return fail(ctx, LK_EINVAL_ARG, "fused matmul: n_extra must be 0..2");
    if (f->n_extra > 0 && f->residual) return fail(ctx, LK_EINVAL_ARG, "fused matmul: stacked matmuls take no residual");
    if (f->mul_dst && f->n_extra != 1) return fail(ctx, LK_EINVAL_ARG, "fused matmul: the SwiGLU epilogue needs exactly one extra weight");
    if (f->silu_dst && !f->mul_dst) return fail(ctx, LK_EINVAL_ARG, "fused matmul: silu_dst goes with mul_dst");
    const lk_tensor *all[] = {a, b, dst, f->norm_x, f->norm_w, f->residual, f->sum_dst, f->n_extra > 0 ? f->a_extra[0] : nullptr,
                              f->n_extra > 1 ? f->a_extra[1] : nullptr, f->n_extra > 0 ? f->dst_extra[0] : nullptr,
                              f->n_extra > 1 ? f->dst_extra[1] : nullptr, f->silu_dst, f->mul_dst};
    for (int i = 0; i < f->n_extra; i++) {
        if (!f->a_extra[i] || !f->dst_extra[i]) return fail(ctx, LK_EINVAL_ARG, "fused matmul: NULL extra tensor");
        if (f->a_extra[i]->type != a->type || f->a_extra[i]->ne[0] != K || f->a_extra[i]->ne[1] != M)
            return fail(ctx, LK_EINVAL_DIM, "fused matmul: extra weights must have a's type and shape");
        if (num_elements(f->dst_extra[i]) != M) return fail(ctx, LK_EINVAL_DIM, "fused matmul: extra dst must hold M elements");
    }
    if (f->mul_dst && num_elements(f->mul_dst) != M) return fail(ctx, LK_EINVAL_DIM, "fused matmul: mul_dst must hold M elements");
    if (f->silu_dst && num_elements(f->silu_dst) != M) return fail(ctx, LK_EINVAL_DIM, "fused matmul: silu_dst must hold M elements");
    for (const lk_tensor *t : all) {
        if (!t) continue;
        if (!t->dev_buf) return fail(ctx, LK_ENOTIMPL, "fused matmul: device-resident tensors only");
        const int64_t nb = lk_tensor_nbytes(t);
        if (nb < 0 || t->data_offset > t->dev_buf->size || (uint64_t)nb > t->dev_buf->size - t->data_offset)
            return fail(ctx, LK_EBOUNDS, "fused matmul: tensor bytes exceed the device buffer");
        if (t != a && t != f->a_extra[0] && t != f->a_extra[1] && t->type != LK_TYPE_F32)
            return fail(ctx, LK_EINVAL_TYPE, "fused matmul: F32 vectors expected");
    }
    if ((f->residual != nullptr) != (f->sum_dst != nullptr)) return fail(ctx, LK_EINVAL_ARG, "fused matmul: residual and sum_dst go together");
    if (f->norm_x && num_elements(f->norm_x) != K) return fail(ctx, LK_EINVAL_DIM, "fused matmul: norm_x must hold K elements");
    if (f->norm_w && num_elements(f->norm_w) != K) return fail(ctx, LK_EINVAL_DIM, "fused matmul: norm_w must hold K elements");
    if (f->residual && (num_elements(f->residual) != M || num_elements(f->sum_dst) != M))
        return fail(ctx, LK_EINVAL_DIM, "fused matmul: residual and sum_dst must hold M elements");
    auto ptr = [](const lk_tensor *t) { return t ? (void *)((uint8_t *)t->dev_buf->ptr + t->data_offset) : nullptr; };
    lk::GemvFusion &fu = L->fu;
    auto &rd = L->rd; auto &wr = L->wr; auto &wt = L->wt;
    const float *x = (const float *)ptr(b);
    if (f->norm_x) {
        fu.pro_mode = 1; fu.pro_w = (const float *)ptr(f->norm_w); fu.pro_eps = f->eps;
        fu.b_out = (f->flags & LK_FUSE_SKIP_B_STORE) ? nullptr : (float *)ptr(b);
        x = (const float *)ptr(f->norm_x);
    }
    float *out = (float *)ptr(dst);
    if (f->residual) { fu.residual = (const float *)ptr(f->residual); fu.dst_raw = out; out = (float *)ptr(f->sum_dst); }
    rd.push_back({fu.pro_w, (size_t)K * 4});
    rd.push_back({fu.residual, (size_t)M * 4});
    wr.push_back({fu.b_out, (size_t)K * 4});
    wr.push_back({fu.dst_raw, (size_t)M * 4});
    fu.nseg_extra = f->n_extra;
    for (int i = 0; i < f->n_extra; i++) {
        fu.a_extra[i] = ptr(f->a_extra[i]);
        fu.dst_extra[i] = (float *)ptr(f->dst_extra[i]);
        wt.push_back({fu.a_extra[i], (size_t)lk_tensor_nbytes(f->a_extra[i])});
        wr.push_back({fu.dst_extra[i], (size_t)M * 4});
    }
    if (f->mul_dst) {
        fu.epi_mode = 1; fu.epi_out = (float *)ptr(f->mul_dst); fu.epi_act = (float *)ptr(f->silu_dst);
        wr.push_back({fu.epi_out, (size_t)M * 4});
        wr.push_back({fu.epi_act, (size_t)M * 4});
    }
    L->a = ptr(a); L->abytes = (size_t)lk_tensor_nbytes(a); L->K = K; L->M = M; L->x = x; L->out = out; L->type = a->type;
    return LK_OK;
}

lk_status lk_compute_matmul_fused(lk_ctx *ctx, const lk_tensor *a, const lk_tensor *b, const lk_tensor *dst,
                                  const lk_matmul_fusion *f) {
    FusedLaunch L;
    const lk_status st = fused_prepare(ctx, a, b, dst, f, &L);
    if (st != LK_OK) return st;
    LK_CUDA(ctx, cudaSetDevice(ctx->device));
    lk::GemvTuning tune = ctx->tune;
    tune.pdl = (ctx->pdl_mode == 2) || (ctx->pdl_mode == 1 && ctx->capturing);
    tune.wait_mode = lk::gemv_wait_mode(ctx, L.a, L.abytes, L.x, (size_t)L.K * 4, L.out, (size_t)L.M * 4, tune.pdl != 0, &L.rd, &L.wr, &L.wt);
    lk::GemvPeers none;
    cudaError_t e = lk::launch_gemv(L.type, L.a, L.K, L.M, L.x, L.out, 0, none, tune, ctx->sm_count, ctx->stream, &L.fu);
    if (e == cudaErrorNotSupported) {
        cudaGetLastError();
        chain_other(ctx, nullptr, 0);      // nothing was launched; the unfused ops that follow do their own bookkeeping
        return fail(ctx, LK_ENOTIMPL, "fused matmul: shape not supported by the streaming GEMV");
    }
    if (e != cudaSuccess) return lk::cuda_fail(ctx, e, "fused matmul launch");
    count_launch(ctx);
    return LK_OK;
}

// A run of dependent fused GEMVs as ONE persistent launch (K1c, lk_gemv.cu): same results as lk_compute_matmul_fused
// called once per link, in order.
lk_status lk_compute_matmul_chain(lk_ctx *ctx, const lk_matmul_link *links, int32_t n) {
    if (!ctx || !links) return fail(ctx, LK_EINVAL_ARG, "lk_compute_matmul_chain: NULL argument");
    if (n < 1) return fail(ctx, LK_EINVAL_ARG, "lk_compute_matmul_chain: n must be >= 1");
    if (n > lk::GEMV_CHAIN_MAX) return fail(ctx, LK_ENOTIMPL, "matmul chain: more links than one launch takes");
    if (!ctx->chain_mode) return fail(ctx, LK_ENOTIMPL, "matmul chain: switched off (gemv_chain = 0)");
    static const lk_matmul_fusion no_fusion = {};
    std::vector<FusedLaunch> L((size_t)n);
    for (int i = 0; i < n; i++) {
        const lk_status st = fused_prepare(ctx, links[i].a, links[i].b, links[i].dst, links[i].fusion ? links[i].fusion : &no_fusion, &L[(size_t)i]);
        if (st != LK_OK) return st;
        if (L[(size_t)i].type != L[0].type) return fail(ctx, LK_ENOTIMPL, "matmul chain: one weight type per chain");
    }
    LK_CUDA(ctx, cudaSetDevice(ctx->device));
    lk::GemvTuning tune = ctx->tune;
    tune.pdl = (ctx->pdl_mode == 2) || (ctx->pdl_mode == 1 && ctx->capturing);
    // The consumers wait for everything earlier before they read phase 0's activations; the producer prefetches weights
    // before that wait unless an op that may still be running wrote one of them.
    std::vector<lk::GemvChainLink> links_d((size_t)n);
    bool weight_hazard = false;
    typedef std::pair<uintptr_t, uintptr_t> Range;
    auto hit = [&](const void *q, size_t bytes) {
        const Range r((uintptr_t)q, (uintptr_t)q + bytes);
        for (const auto *v : {&ctx->chain_writes, &ctx->other_writes})
            for (const Range &w : *v)
                if (r.first < w.second && w.first < r.second) return true;
        return false;
    };
    for (int i = 0; i < n; i++) {
        const FusedLaunch &F = L[(size_t)i];
        lk::GemvChainLink &d = links_d[(size_t)i];
        d.a = F.a; d.K = F.K; d.M = F.M; d.b = F.x; d.dst = F.out; d.fu = F.fu;
        weight_hazard = weight_hazard || hit(F.a, F.abytes);
        for (const auto &w : F.wt) if (w.first) weight_hazard = weight_hazard || hit(w.first, w.second);
        // a weight that an earlier link of this chain writes cannot be prefetched at all: not one launch
        for (int j = 0; j < i; j++) {
            std::vector<std::pair<const void *, size_t>> wts = F.wt, outs = L[(size_t)j].wr;
            wts.push_back({F.a, F.abytes});
            outs.push_back({L[(size_t)j].out, (size_t)L[(size_t)j].M * 4});
            for (const auto &w : wts)
                for (const auto &o : outs)
                    if (w.first && o.first && (uintptr_t)w.first < (uintptr_t)o.first + o.second && (uintptr_t)o.first < (uintptr_t)w.first + w.second)
                        return fail(ctx, LK_ENOTIMPL, "matmul chain: a link's weight is written by an earlier link");
        }
    }
    tune.wait_mode = weight_hazard ? 0 : 1;
    // K1m first (plane sums on the int8 MMA, hand-offs by mailbox: lk_gemv_mma.cu); what it declines — other types, K % 256 != 0,
    // phases that share an output tensor — goes to K1c
    cudaError_t e = cudaErrorNotSupported;
    if (ctx->chain_mma) {
        lk::GemvTuning tm = tune;
        tm.max_stages = 0;
        e = lk::launch_gemv_mchain(L[0].type, n, links_d.data(), tm, ctx->sm_count, ctx->stream, ctx->chain_bar, (uint8_t *)ctx->chain_bar + 256);
        if (e == cudaErrorNotSupported) cudaGetLastError(); else if (e == cudaSuccess) ctx->chain_mma_launches++;
    }
    if (e == cudaErrorNotSupported) e = lk::launch_gemv_chain(L[0].type, n, links_d.data(), tune, ctx->sm_count, ctx->stream, ctx->chain_bar);
    if (e == cudaErrorNotSupported) {
        cudaGetLastError();
        return fail(ctx, LK_ENOTIMPL, "matmul chain: shapes not supported as one launch");
    }
    if (e != cudaSuccess) return lk::cuda_fail(ctx, e, "matmul chain launch");
    // to whatever follows, the chain is one op that waits for everything before it and completes as a whole
    ctx->other_writes.clear();
    for (int i = 0; i < n; i++) {
        const FusedLaunch &F = L[(size_t)i];
        chain_other(ctx, F.out, (size_t)F.M * 4);
        for (const auto &w : F.wr) if (w.first) chain_other(ctx, w.first, w.second);
    }
    ctx->chain_launches++;
    ctx->chain_unchecked++;
    count_launch(ctx);
    return LK_OK;
}

lk_status lk_compute_dequantize(lk_ctx *ctx, const lk_tensor *src, const lk_tensor *dst) {
    if (!ctx || !src || !dst) return fail(ctx, LK_EINVAL_ARG, "lk_compute_dequantize: NULL argument");
    if (!is_block_type(src->type) && !is_kq_type(src->type))
        return fail(ctx, LK_ENOTIMPL, "dequantize: source must be Q4_0, Q4_1, Q8_0, Q2_K, Q4_K, Q8_K or BITNET_1_58");
    if (dst->type != LK_TYPE_F32) return fail(ctx, LK_EINVAL_TYPE, "dequantize: destination must be F32");
    const int64_t n = num_elements(src);
    if (n != num_elements(dst)) return fail(ctx, LK_EINVAL_DIM, "dequantize: element counts differ");
    if (n % (is_kq_type(src->type) ? lk::kq_type_block_elems(src->type) : LK_QK))
        return fail(ctx, LK_EINVAL_DIM, "dequantize: element count must be a multiple of the block size");
    LK_CUDA(ctx, cudaSetDevice(ctx->device));
    OpScope sc(ctx);
    lk_status st = arena_reserve(ctx, host_need({src, dst}));
    if (st != LK_OK) return st;
    void *ds, *dd;
    if ((st = resolve(sc, src, false, &ds, "src")) != LK_OK) return st;
    if ((st = resolve(sc, dst, true, &dd, "dst")) != LK_OK) return st;
    chain_other(ctx, dd, (size_t)n * 4);
    cudaError_t e = is_kq_type(src->type) ? lk::launch_kq_dequantize(src->type, ds, n, (float *)dd, ctx->stream)
                                          : lk::launch_dequantize(src->type, ds, n, (float *)dd, ctx->stream, ctx->tune.nibble_order);
    if (e != cudaSuccess) return lk::cuda_fail(ctx, e, "dequantize launch");
    count_launch(ctx);
    return finish(sc);
}

// quantizeTensor(graphAllocator, tensor, targetType) — core/GGMLComputeOps.kt:1040-1433, destination-based: src F32, dst
// Q8_0 / Q4_0 / Q4_1 with the same element count (a multiple of 32, require(...) at :1059 / :1073 / :1163).  Bit-exact.
lk_status lk_compute_quantize(lk_ctx *ctx, const lk_tensor *src, const lk_tensor *dst) {
    if (!ctx || !src || !dst) return fail(ctx, LK_EINVAL_ARG, "lk_compute_quantize: NULL argument");
    if (src->type != LK_TYPE_F32) return fail(ctx, LK_EINVAL_TYPE, "quantize: source must be F32");
    if (!is_block_type(dst->type)) return fail(ctx, LK_ENOTIMPL, "quantize: destination must be Q8_0, Q4_0 or Q4_1 on this device path");
    const int64_t n = num_elements(src);
    if (n != num_elements(dst)) return fail(ctx, LK_EINVAL_DIM, "quantize: element counts differ");
    if (n % LK_QK) return fail(ctx, LK_EINVAL_DIM, "quantize: element count must be a multiple of 32");
    LK_CUDA(ctx, cudaSetDevice(ctx->device));
    OpScope sc(ctx);
    lk_status st = arena_reserve(ctx, host_need({src, dst}));
    if (st != LK_OK) return st;
    void *ds, *dd;
    if ((st = resolve(sc, src, false, &ds, "src")) != LK_OK) return st;
    if ((st = resolve(sc, dst, true, &dd, "dst")) != LK_OK) return st;
    chain_other(ctx, dd, (size_t)lk_tensor_nbytes(dst));
    const cudaError_t e = lk::launch_quantize(dst->type, (const float *)ds, n, dd, ctx->stream, ctx->tune.nibble_order);
    if (e != cudaSuccess) return lk::cuda_fail(ctx, e, "quantize launch");
    count_launch(ctx);
    return finish(sc);
}

typedef cudaError_t (*binary_fn)(const float *, const float *, int64_t, float *, cudaStream_t, bool);

static lk_status binary_op(lk_ctx *ctx, const lk_tensor *a, const lk_tensor *b, const lk_tensor *dst, binary_fn fn,
                           const char *name) {
    if (!ctx || !a || !b || !dst) return fail(ctx, LK_EINVAL_ARG, std::string(name) + ": NULL argument");
    // core/GGMLComputeOps.kt:742-750: dims must match (no broadcast), dst dims and type must match
    if (!same_shape(a, b)) return fail(ctx, LK_EINVAL_DIM, std::string(name) + ": input dimensions differ");
    if (!same_shape(a, dst)) return fail(ctx, LK_EINVAL_DIM, std::string(name) + ": result dimensions must match input dimensions");
    if (dst->type != a->type || b->type != a->type) return fail(ctx, LK_EINVAL_TYPE, std::string(name) + ": tensor types must match");
    if (a->type != LK_TYPE_F32) return fail(ctx, LK_ENOTIMPL, std::string(name) + ": F32 only on this device path");
    LK_CUDA(ctx, cudaSetDevice(ctx->device));
    OpScope sc(ctx);
    lk_status st = arena_reserve(ctx, host_need({a, b, dst}));
    if (st != LK_OK) return st;
    void *da, *db, *dd;
    if ((st = resolve(sc, a, false, &da, "a")) != LK_OK) return st;
    if ((st = resolve(sc, b, false, &db, "b")) != LK_OK) return st;
    if ((st = resolve(sc, dst, true, &dd, "dst")) != LK_OK) return st;
    chain_other(ctx, dd, (size_t)num_elements(a) * 4);
    cudaError_t e = fn((const float *)da, (const float *)db, num_elements(a), (float *)dd, ctx->stream, op_pdl(ctx, sc));
    if (e != cudaSuccess) return lk::cuda_fail(ctx, e, name);
    count_launch(ctx);
    return finish(sc);
}

lk_status lk_compute_add(lk_ctx *ctx, const lk_tensor *a, const lk_tensor *b, const lk_tensor *dst) {
    return binary_op(ctx, a, b, dst, lk::launch_add, "computeAdd");
}
lk_status lk_compute_mul(lk_ctx *ctx, const lk_tensor *a, const lk_tensor *b, const lk_tensor *dst) {
    return binary_op(ctx, a, b, dst, lk::launch_mul, "computeMul");
}
lk_status lk_compute_silu_mul(lk_ctx *ctx, const lk_tensor *g, const lk_tensor *u, const lk_tensor *dst) {
    return binary_op(ctx, g, u, dst, lk::launch_silu_mul, "siluMul");
}

}  // extern "C"

extern "C" {

static lk_status unary_check(lk_ctx *ctx, const lk_tensor *x, const lk_tensor *dst, const char *name) {
    if (!ctx || !x || !dst) return fail(ctx, LK_EINVAL_ARG, std::string(name) + ": NULL argument");
    if (!same_shape(x, dst)) return fail(ctx, LK_EINVAL_DIM, std::string(name) + ": result dimensions must match input dimensions");
    if (dst->type != x->type) return fail(ctx, LK_EINVAL_TYPE, std::string(name) + ": result type must match input type");
    if (x->type != LK_TYPE_F32) return fail(ctx, LK_ENOTIMPL, std::string(name) + ": F32 only on this device path");
    return LK_OK;
}

lk_status lk_compute_silu(lk_ctx *ctx, const lk_tensor *x, const lk_tensor *dst) {
    lk_status st = unary_check(ctx, x, dst, "silu");
    if (st != LK_OK) return st;
    LK_CUDA(ctx, cudaSetDevice(ctx->device));
    OpScope sc(ctx);
    if ((st = arena_reserve(ctx, host_need({x, dst}))) != LK_OK) return st;
    void *dx, *dd;
    if ((st = resolve(sc, x, false, &dx, "x")) != LK_OK) return st;
    if ((st = resolve(sc, dst, true, &dd, "dst")) != LK_OK) return st;
    chain_other(ctx, dd, (size_t)num_elements(x) * 4);
    cudaError_t e = lk::launch_silu((const float *)dx, num_elements(x), (float *)dd, ctx->stream, op_pdl(ctx, sc));
    if (e != cudaSuccess) return lk::cuda_fail(ctx, e, "silu launch");
    count_launch(ctx);
    return finish(sc);
}

lk_status lk_compute_rms_norm(lk_ctx *ctx, const lk_tensor *x, const lk_tensor *w, float eps, const lk_tensor *dst) {
    lk_status st = unary_check(ctx, x, dst, "rmsNorm");
    if (st != LK_OK) return st;
    const int64_t ne0 = x->ne[0];
    if (w) {
        if (w->type != LK_TYPE_F32) return fail(ctx, LK_EINVAL_TYPE, "rmsNorm: weight must be F32");
        if (w->ne[0] != ne0 || num_elements(w) != ne0) return fail(ctx, LK_EINVAL_DIM, "rmsNorm: weight must have ne = [ne0]");
    }
    LK_CUDA(ctx, cudaSetDevice(ctx->device));
    OpScope sc(ctx);
    if ((st = arena_reserve(ctx, host_need({x, w, dst}))) != LK_OK) return st;
    void *dx, *dw = nullptr, *dd;
    if ((st = resolve(sc, x, false, &dx, "x")) != LK_OK) return st;
    if (w && (st = resolve(sc, w, false, &dw, "weight")) != LK_OK) return st;
    if ((st = resolve(sc, dst, true, &dd, "dst")) != LK_OK) return st;
    const int64_t n = num_elements(x);
    chain_other(ctx, dd, (size_t)n * 4);
    cudaError_t e = lk::launch_rms_norm((const float *)dx, (const float *)dw, eps, ne0, ne0 ? n / ne0 : 0, (float *)dd, ctx->stream, op_pdl(ctx, sc));
    if (e != cudaSuccess) return lk::cuda_fail(ctx, e, "rmsNorm launch");
    count_launch(ctx);
    return finish(sc);
}

lk_status lk_compute_softmax(lk_ctx *ctx, const lk_tensor *x, const lk_tensor *dst) {
    lk_status st = unary_check(ctx, x, dst, "softmax");
    if (st != LK_OK) return st;
    LK_CUDA(ctx, cudaSetDevice(ctx->device));
    OpScope sc(ctx);
    if ((st = arena_reserve(ctx, host_need({x, dst}))) != LK_OK) return st;
    void *dx, *dd;
    if ((st = resolve(sc, x, false, &dx, "x")) != LK_OK) return st;
    if ((st = resolve(sc, dst, true, &dd, "dst")) != LK_OK) return st;
    const int64_t ne0 = x->ne[0], n = num_elements(x);
    chain_other(ctx, dd, (size_t)n * 4);
    cudaError_t e = lk::launch_softmax((const float *)dx, ne0, ne0 ? n / ne0 : 0, (float *)dd, ctx->stream, op_pdl(ctx, sc));
    if (e != cudaSuccess) return lk::cuda_fail(ctx, e, "softmax launch");
    count_launch(ctx);
    return finish(sc);
}

// ---- N4: attention core (lk_attention.cu) -------------------------------------------------------------------
lk_status lk_compute_scale(lk_ctx *ctx, const lk_tensor *x, float factor, const lk_tensor *dst) {
    lk_status st = unary_check(ctx, x, dst, "scale");
    if (st != LK_OK) return st;
    LK_CUDA(ctx, cudaSetDevice(ctx->device));
    OpScope sc(ctx);
    if ((st = arena_reserve(ctx, host_need({x, dst}))) != LK_OK) return st;
    void *dx, *dd;
    if ((st = resolve(sc, x, false, &dx, "x")) != LK_OK) return st;
    if ((st = resolve(sc, dst, true, &dd, "dst")) != LK_OK) return st;
    chain_other(ctx, dd, (size_t)num_elements(x) * 4);
    cudaError_t e = lk::launch_scale((const float *)dx, factor, num_elements(x), (float *)dd, ctx->stream, op_pdl(ctx, sc));
    if (e != cudaSuccess) return lk::cuda_fail(ctx, e, "scale launch");
    count_launch(ctx);
    return finish(sc);
}

lk_status lk_rope_table(int32_t head_dim, float rope_theta, int32_t max_pos, float *out) {
    if (!out || head_dim <= 0 || (head_dim & 1) || max_pos <= 0) return LK_EINVAL_ARG;
    // applyRoPE, model/LlamaAttention.kt:53-58, in the same F64 operations
    for (int p = 0; p < max_pos; p++)
        for (int i = 0; i < head_dim; i += 2) {
            const double exponent = ((double)i / 2.0) / ((double)head_dim / 2.0);
            const double inv_freq = 1.0 / pow((double)rope_theta, exponent);
            const double angle = (double)p * inv_freq;
            float *o = out + ((size_t)p * (head_dim / 2) + i / 2) * 2;
            o[0] = (float)cos(angle);
            o[1] = (float)sin(angle);
        }
    return LK_OK;
}

static lk_status pos_check(lk_ctx *ctx, const lk_tensor *pos, const char *name) {
    if (pos && (pos->type != LK_TYPE_I32 || num_elements(pos) < 1))
        return fail(ctx, LK_EINVAL_TYPE, std::string(name) + ": position must be an I32 tensor with one element");
    return LK_OK;
}

lk_status lk_compute_rope(lk_ctx *ctx, const lk_tensor *x, const lk_tensor *table, const lk_tensor *pos, int32_t head_dim,
                          const lk_tensor *dst) {
    lk_status st = unary_check(ctx, x, dst, "rope");
    if (st != LK_OK) return st;
    if (!table) return fail(ctx, LK_EINVAL_ARG, "rope: table is NULL");
    if ((st = pos_check(ctx, pos, "rope")) != LK_OK) return st;
    const int64_t row = x->ne[0];
    if (head_dim <= 0 || (head_dim & 1) || row % head_dim) return fail(ctx, LK_EINVAL_DIM, "rope: ne[0] must be a multiple of an even head_dim");
    if (table->type != LK_TYPE_F32 || table->ne[0] != 2 || table->ne[1] != head_dim / 2 || table->ne[2] < 1)
        return fail(ctx, LK_EINVAL_DIM, "rope: table must be F32 with ne = [2, head_dim/2, max_pos]");
    const int64_t ntok = row ? num_elements(x) / row : 0;
    if (ntok > table->ne[2]) return fail(ctx, LK_EINVAL_DIM, "rope: more tokens than table positions");
    LK_CUDA(ctx, cudaSetDevice(ctx->device));
    OpScope sc(ctx);
    if ((st = arena_reserve(ctx, host_need({x, table, pos, dst}))) != LK_OK) return st;
    void *dx, *dt, *dp = nullptr, *dd;
    if ((st = resolve(sc, x, false, &dx, "x")) != LK_OK) return st;
    if ((st = resolve(sc, table, false, &dt, "table")) != LK_OK) return st;
    if (pos && (st = resolve(sc, pos, false, &dp, "pos")) != LK_OK) return st;
    if ((st = resolve(sc, dst, true, &dd, "dst")) != LK_OK) return st;
    chain_other(ctx, dd, (size_t)num_elements(x) * 4);
    cudaError_t e = lk::launch_rope((const float *)dx, (const float *)dt, (const int32_t *)dp, head_dim, row, (int)ntok,
                                    (int)table->ne[2], (float *)dd, ctx->stream, op_pdl(ctx, sc));
    if (e == cudaErrorNotSupported) return fail(ctx, LK_ENOTIMPL, "rope: unsupported shape / alignment");
    if (e != cudaSuccess) return lk::cuda_fail(ctx, e, "rope launch");
    count_launch(ctx);
    return finish(sc);
}

lk_status lk_compute_kv_append(lk_ctx *ctx, const lk_tensor *src, const lk_tensor *cache, const lk_tensor *pos) {
    if (!ctx || !src || !cache) return fail(ctx, LK_EINVAL_ARG, "kvAppend: NULL argument");
    if (src->type != LK_TYPE_F32 || cache->type != LK_TYPE_F32) return fail(ctx, LK_EINVAL_TYPE, "kvAppend: F32 only");
    if (!cache->dev_buf) return fail(ctx, LK_EINVAL_ARG, "kvAppend: the cache must be device-resident");
    lk_status st = pos_check(ctx, pos, "kvAppend");
    if (st != LK_OK) return st;
    // KVCache: cache ne = [headDim, numHeads, maxSeq, 1], new rows ne = [headDim, numHeads, newSeqLen, 1] (model/KVCache.kt:19-34, :43)
    if (src->ne[0] != cache->ne[0] || src->ne[1] != cache->ne[1]) return fail(ctx, LK_EINVAL_DIM, "kvAppend: [headDim, numHeads] differ");
    const int64_t row = cache->ne[0] * cache->ne[1], ntok = src->ne[2] > 0 ? src->ne[2] : 1, max_seq = cache->ne[2] > 0 ? cache->ne[2] : 1;
    if (ntok > max_seq) return fail(ctx, LK_EINVAL_DIM, "Sequence length would exceed maximum");    // model/KVCache.kt:44
    LK_CUDA(ctx, cudaSetDevice(ctx->device));
    OpScope sc(ctx);
    if ((st = arena_reserve(ctx, host_need({src, pos}))) != LK_OK) return st;
    void *ds, *dc, *dp = nullptr;
    if ((st = resolve(sc, src, false, &ds, "src")) != LK_OK) return st;
    if ((st = resolve(sc, cache, false, &dc, "cache")) != LK_OK) return st;
    if (pos && (st = resolve(sc, pos, false, &dp, "pos")) != LK_OK) return st;
    chain_other(ctx, dc, (size_t)lk_tensor_nbytes(cache));
    cudaError_t e = lk::launch_kv_append((const float *)ds, (float *)dc, (const int32_t *)dp, row, (int)ntok, max_seq, ctx->stream, op_pdl(ctx, sc));
    if (e == cudaErrorNotSupported) return fail(ctx, LK_ENOTIMPL, "kvAppend: unsupported shape / alignment");
    if (e != cudaSuccess) return lk::cuda_fail(ctx, e, "kvAppend launch");
    count_launch(ctx);
    return finish(sc);
}

lk_status lk_compute_attention_decode(lk_ctx *ctx, const lk_tensor *q, const lk_tensor *kcache, const lk_tensor *vcache,
                                      const lk_tensor *pos, float scale, const lk_tensor *dst) {
    if (!ctx || !q || !kcache || !vcache || !pos || !dst) return fail(ctx, LK_EINVAL_ARG, "attention: NULL argument");
    if (q->type != LK_TYPE_F32 || kcache->type != LK_TYPE_F32 || vcache->type != LK_TYPE_F32 || dst->type != LK_TYPE_F32)
        return fail(ctx, LK_EINVAL_TYPE, "attention: F32 only");
    lk_status st = pos_check(ctx, pos, "attention");
    if (st != LK_OK) return st;
    if (!kcache->dev_buf || !vcache->dev_buf) return fail(ctx, LK_EINVAL_ARG, "attention: the caches must be device-resident");
    const int64_t D = kcache->ne[0], H = kcache->ne[1], max_seq = kcache->ne[2] > 0 ? kcache->ne[2] : 1;
    if (!same_shape(kcache, vcache)) return fail(ctx, LK_EINVAL_DIM, "attention: key and value caches differ in shape");
    if (num_elements(q) != D * H || !same_shape(q, dst)) return fail(ctx, LK_EINVAL_DIM, "attention: q / dst must hold headDim * numHeads elements");
    LK_CUDA(ctx, cudaSetDevice(ctx->device));
    OpScope sc(ctx);
    if ((st = arena_reserve(ctx, host_need({q, pos, dst}))) != LK_OK) return st;
    if ((st = attn_ws_reserve(ctx, lk::attn_workspace_bytes((int)D, (int)H, max_seq))) != LK_OK) return st;
    void *dq, *dk, *dv, *dp, *dd;
    if ((st = resolve(sc, q, false, &dq, "q")) != LK_OK) return st;
    if ((st = resolve(sc, kcache, false, &dk, "kcache")) != LK_OK) return st;
    if ((st = resolve(sc, vcache, false, &dv, "vcache")) != LK_OK) return st;
    if ((st = resolve(sc, pos, false, &dp, "pos")) != LK_OK) return st;
    if ((st = resolve(sc, dst, true, &dd, "dst")) != LK_OK) return st;
    chain_other(ctx, dd, (size_t)D * H * 4);
    cudaError_t e = lk::launch_attn_decode((const float *)dq, (const float *)dk, (const float *)dv, (const int32_t *)dp, (int)D, (int)H,
                                           max_seq, scale, (float *)dd, ctx->attn_ws, ctx->sm_count, ctx->stream, op_pdl(ctx, sc));
    if (e == cudaErrorNotSupported) return fail(ctx, LK_ENOTIMPL, "attention: headDim must be 64 or 128, tensors 16-byte aligned");
    if (e != cudaSuccess) return lk::cuda_fail(ctx, e, "attention launch");
    count_launch(ctx);
    return finish(sc);
}

// lk_compute_attention_decode without any allocation (no counterpart in the reference):
// device-resident contiguous tensors only — a host tensor would need the staging arena — and the scratch
// memory is the static device array of lk_attention.cu.  Calls no cudaMalloc / cudaFree and nothing that allocates device
// or pinned memory; the only host-side growth possible is the small std::vector of the launch bookkeeping.
lk_status lk_compute_attention_decode_noalloc(lk_ctx *ctx, const lk_tensor *q, const lk_tensor *kcache, const lk_tensor *vcache,
                                              const lk_tensor *pos, float scale, const lk_tensor *dst) {
    if (!ctx || !q || !kcache || !vcache || !pos || !dst) return fail(ctx, LK_EINVAL_ARG, "attention (noalloc): NULL argument");
    if (q->type != LK_TYPE_F32 || kcache->type != LK_TYPE_F32 || vcache->type != LK_TYPE_F32 || dst->type != LK_TYPE_F32)
        return fail(ctx, LK_EINVAL_TYPE, "attention (noalloc): F32 only");
    lk_status st = pos_check(ctx, pos, "attention (noalloc)");
    if (st != LK_OK) return st;
    const int64_t D = kcache->ne[0], H = kcache->ne[1], max_seq = kcache->ne[2] > 0 ? kcache->ne[2] : 1;
    if (!same_shape(kcache, vcache)) return fail(ctx, LK_EINVAL_DIM, "attention (noalloc): key and value caches differ in shape");
    if (num_elements(q) != D * H || !same_shape(q, dst)) return fail(ctx, LK_EINVAL_DIM, "attention (noalloc): q / dst must hold headDim * numHeads elements");
    const lk_tensor *all[] = {q, kcache, vcache, pos, dst};
    void *ptr[5];
    for (int i = 0; i < 5; i++) {
        const lk_tensor *t = all[i];
        if (!t->dev_buf || t->host_base)
            return fail(ctx, LK_EINVAL_ARG, "attention (noalloc): device-resident tensors only (this entry point allocates nothing, so it cannot stage host tensors)");
        if (!contiguous_f32(t)) return fail(ctx, LK_EINVAL_ARG, "attention (noalloc): contiguous tensors only");
        const int64_t nb = lk_tensor_nbytes(t);
        if (nb < 0 || t->data_offset > t->dev_buf->size || (uint64_t)nb > t->dev_buf->size - t->data_offset)
            return fail(ctx, LK_EBOUNDS, "attention (noalloc): tensor bytes exceed the device buffer");
        if (t->data_offset % 16) return fail(ctx, LK_EINVAL_ARG, "attention (noalloc): dataOffset must be 16-byte aligned on device");
        ptr[i] = (uint8_t *)t->dev_buf->ptr + t->data_offset;
    }
    if (D != 64 && D != 128) return fail(ctx, LK_ENOTIMPL, "attention (noalloc): headDim must be 64 or 128");
    if (H <= 0 || H > (1 << 20) || max_seq > (1 << 24)) return fail(ctx, LK_EINVAL_DIM, "attention (noalloc): bad numHeads / maxSeq");
    const size_t need = lk::attn_scratch_needed((int)D, (int)H, max_seq, ctx->sm_count), cap = lk::attn_static_scratch_capacity();
    if (need > cap)
        return fail(ctx, LK_EINVAL_ARG, "attention (noalloc): this shape needs " + std::to_string(need) + " bytes of scratch memory, the static array holds " +
                    std::to_string(cap) + " (LK_ATTN_NOALLOC_SCRATCH_BYTES); use lk_compute_attention_decode");
    LK_CUDA(ctx, cudaSetDevice(ctx->device));
    chain_other(ctx, ptr[4], (size_t)D * H * 4);
    const cudaError_t e = lk::launch_attn_decode_static(ctx->device, (const float *)ptr[0], (const float *)ptr[1], (const float *)ptr[2],
                                                        (const int32_t *)ptr[3], (int)D, (int)H, max_seq, scale, (float *)ptr[4], ctx->sm_count, ctx->stream);
    if (e == cudaErrorNotSupported) return fail(ctx, LK_ENOTIMPL, "attention (noalloc): unsupported shape / alignment");
    if (e != cudaSuccess) return lk::cuda_fail(ctx, e, "attention (noalloc) launch");
    count_launch(ctx);
    return LK_OK;
}

lk_status lk_compute_attention_step(lk_ctx *ctx, const lk_tensor *q, const lk_tensor *k, const lk_tensor *v,
                                    const lk_tensor *table, const lk_tensor *pos, int32_t head_dim, const lk_tensor *kcache,
                                    const lk_tensor *vcache, float scale, const lk_tensor *q_rope_dst,
                                    const lk_tensor *k_rope_dst, const lk_tensor *dst) {
    if (!ctx || !q || !k || !v || !table || !pos || !kcache || !vcache || !dst) return fail(ctx, LK_EINVAL_ARG, "attentionStep: NULL argument");
    const lk_tensor *all[] = {q, k, v, table, kcache, vcache, q_rope_dst, k_rope_dst, dst};
    for (const lk_tensor *t : all) {
        if (!t) continue;
        if (t->type != LK_TYPE_F32) return fail(ctx, LK_EINVAL_TYPE, "attentionStep: F32 only");
        if (!t->dev_buf) return fail(ctx, LK_ENOTIMPL, "attentionStep: device-resident tensors only");
    }
    lk_status st = pos_check(ctx, pos, "attentionStep");
    if (st != LK_OK) return st;
    if (!pos->dev_buf) return fail(ctx, LK_ENOTIMPL, "attentionStep: device-resident tensors only");
    const int64_t D = kcache->ne[0], H = kcache->ne[1], max_seq = kcache->ne[2] > 0 ? kcache->ne[2] : 1;
    if (!same_shape(kcache, vcache)) return fail(ctx, LK_EINVAL_DIM, "attentionStep: key and value caches differ in shape");
    if (head_dim != D) return fail(ctx, LK_EINVAL_DIM, "attentionStep: head_dim must equal the caches' ne[0]");
    for (const lk_tensor *t : {q, k, v, q_rope_dst, k_rope_dst, dst})
        if (t && num_elements(t) != D * H) return fail(ctx, LK_EINVAL_DIM, "attentionStep: vectors must hold headDim * numHeads elements");
    if (table->ne[0] != 2 || table->ne[1] != head_dim / 2 || table->ne[2] < 1)
        return fail(ctx, LK_EINVAL_DIM, "attentionStep: table must be F32 with ne = [2, head_dim/2, max_pos]");
    LK_CUDA(ctx, cudaSetDevice(ctx->device));
    OpScope sc(ctx);
    if ((st = attn_ws_reserve(ctx, lk::attn_workspace_bytes((int)D, (int)H, max_seq))) != LK_OK) return st;
    void *dq, *dk, *dv, *dt, *dp, *dkc, *dvc, *dqr = nullptr, *dkr = nullptr, *dd;
    if ((st = resolve(sc, q, false, &dq, "q")) != LK_OK) return st;
    if ((st = resolve(sc, k, false, &dk, "k")) != LK_OK) return st;
    if ((st = resolve(sc, v, false, &dv, "v")) != LK_OK) return st;
    if ((st = resolve(sc, table, false, &dt, "table")) != LK_OK) return st;
    if ((st = resolve(sc, pos, false, &dp, "pos")) != LK_OK) return st;
    if ((st = resolve(sc, kcache, false, &dkc, "kcache")) != LK_OK) return st;
    if ((st = resolve(sc, vcache, false, &dvc, "vcache")) != LK_OK) return st;
    if (q_rope_dst && (st = resolve(sc, q_rope_dst, true, &dqr, "q_rope_dst")) != LK_OK) return st;
    if (k_rope_dst && (st = resolve(sc, k_rope_dst, true, &dkr, "k_rope_dst")) != LK_OK) return st;
    if ((st = resolve(sc, dst, true, &dd, "dst")) != LK_OK) return st;
    // a fully serialising op for the GEMV chain bookkeeping: it writes dst, both rotated vectors and a row of each cache
    chain_other(ctx, dd, (size_t)D * H * 4);
    chain_other(ctx, dkc, (size_t)lk_tensor_nbytes(kcache));
    chain_other(ctx, dvc, (size_t)lk_tensor_nbytes(vcache));
    if (dqr) chain_other(ctx, dqr, (size_t)D * H * 4);
    if (dkr) chain_other(ctx, dkr, (size_t)D * H * 4);
    lk::AttnStep step;
    step.k_raw = (const float *)dk; step.v_raw = (const float *)dv; step.table = (const float *)dt;
    step.table_pos = (int)table->ne[2]; step.q_rope = (float *)dqr; step.k_rope = (float *)dkr;
    cudaError_t e = lk::launch_attn_decode((const float *)dq, (const float *)dkc, (const float *)dvc, (const int32_t *)dp, (int)D, (int)H,
                                           max_seq, scale, (float *)dd, ctx->attn_ws, ctx->sm_count, ctx->stream, op_pdl(ctx, sc), &step);
    if (e == cudaErrorNotSupported) return fail(ctx, LK_ENOTIMPL, "attentionStep: headDim must be 64 or 128, tensors 16-byte aligned");
    if (e != cudaSuccess) return lk::cuda_fail(ctx, e, "attentionStep launch");
    count_launch(ctx);
    return finish(sc);
}

// ---- op-sequence capture -----------------------------------------------------
lk_status lk_capture_begin(lk_ctx *ctx) {
    if (!ctx) return fail(nullptr, LK_EINVAL_ARG, "lk_capture_begin: ctx is NULL");
    if (ctx->capturing) return fail(ctx, LK_EINVAL_ARG, "lk_capture_begin: already capturing");
    LK_CUDA(ctx, cudaSetDevice(ctx->device));
    LK_CUDA(ctx, cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeThreadLocal));
    ctx->capturing = true;
    ctx->cap_in = ctx->cap_out = false;
    LK_CUDA(ctx, cudaEventRecord(ctx->ev_root, ctx->stream));   // where a copy stream forks from
    ctx->capture_launches = 0;
    ctx->capture_chain0 = ctx->chain_unchecked;
    ctx->capture_barrier0 = ctx->barrier_unchecked;
    chain_other(ctx, nullptr, 0);
    return LK_OK;
}

lk_status lk_capture_end(lk_ctx *ctx, lk_graph **out) {
    if (!ctx || !out) return fail(ctx, LK_EINVAL_ARG, "lk_capture_end: NULL argument");
    if (!ctx->capturing) return fail(ctx, LK_EINVAL_ARG, "lk_capture_end: not capturing");
    ctx->capturing = false;
    chain_other(ctx, nullptr, 0);
    cudaGraph_t g = nullptr;
    // copy streams that joined the capture come back to the compute stream before it ends
    if (ctx->cap_in) { cudaEventRecord(ctx->ev_in, ctx->copy_in); cudaStreamWaitEvent(ctx->stream, ctx->ev_in, 0); }
    if (ctx->cap_out) { cudaEventRecord(ctx->ev_out, ctx->copy_out); cudaStreamWaitEvent(ctx->stream, ctx->ev_out, 0); }
    ctx->cap_in = ctx->cap_out = false;
    LK_CUDA(ctx, cudaStreamEndCapture(ctx->stream, &g));
    cudaGraphExec_t ex = nullptr;
    cudaError_t e = cudaGraphInstantiate(&ex, g, 0);
    if (e != cudaSuccess) { cudaGraphDestroy(g); return lk::cuda_fail(ctx, e, "cudaGraphInstantiate"); }
    lk_graph *gr = new lk_graph();
    gr->ctx = ctx; gr->graph = g; gr->exec = ex; gr->launches = ctx->capture_launches;
    gr->chain_launches = ctx->chain_unchecked - ctx->capture_chain0;       // what a replay of this graph contains
    gr->barrier_launches = ctx->barrier_unchecked - ctx->capture_barrier0;
    ctx->chain_unchecked = ctx->capture_chain0;                             // capturing launched nothing
    ctx->barrier_unchecked = ctx->capture_barrier0;
    *out = gr;
    return LK_OK;
}

lk_status lk_graph_launch(lk_ctx *ctx, lk_graph *g) {
    if (!ctx || !g) return fail(ctx, LK_EINVAL_ARG, "lk_graph_launch: NULL argument");
    LK_CUDA(ctx, cudaGraphLaunch(g->exec, ctx->stream));
    ctx->chain_unchecked += g->chain_launches;
    ctx->barrier_unchecked += g->barrier_launches;
    chain_other(ctx, nullptr, 0);
    ctx->launches += g->launches;
    return LK_OK;
}

lk_status lk_graph_destroy(lk_graph *g) {
    if (!g) return LK_OK;
    cudaStreamSynchronize(g->ctx->stream);
    cudaGraphExecDestroy(g->exec);
    cudaGraphDestroy(g->graph);
    delete g;
    return LK_OK;
}

// ---- options -------------------------------------------------------------------
lk_status lk_set_option(lk_ctx *ctx, const char *name, int64_t v) {
    if (!ctx || !name) return fail(ctx, LK_EINVAL_ARG, "lk_set_option: NULL argument");
    const std::string n(name);
    if (n == "gemv_pdl") ctx->pdl_mode = (int)v;
    else if (n == "gemv_warps") ctx->tune.variant = (int)v;
    else if (n == "gemv_stage_bytes") ctx->tune.stage_bytes = (int)v;
    else if (n == "gemv_row_quant") ctx->tune.row_quant = (int)v;
    else if (n == "gemv_max_stages") ctx->tune.max_stages = (int)v;
    else if (n == "gemv_smem_budget") ctx->tune.smem_budget = (int)v;
    else if (n == "gemv_ctas_per_sm") ctx->tune.ctas_per_sm = (int)v;
    else if (n == "gemv_overlap") ctx->overlap_mode = (int)v;
    else if (n == "gemv_chain") ctx->chain_mode = v ? 1 : 0;
    else if (n == "gemv_chain_mma") ctx->chain_mma = v ? 1 : 0;
    else if (n == "gemv_act_planes") ctx->tune.act_planes = (v == 2) ? 2 : 3;
    else if (n == "gemv_bitnet_copies") ctx->tune.bitnet_copies = (v == 1 || v == 2 || v == 4 || v == 8 || v == 16) ? (int)v : 0;
    else if (n == "nibble_order") ctx->tune.nibble_order = v ? 1 : 0;
    else if (n == "gemv_fix_gap") { ctx->tune.fix_gap = v < 2 ? 2 : v > 255 ? 255 : (int)v; ctx->tune.fix_gap_set = true; }
    else if (n == "gemm_tc_min_n") ctx->gemm_tc_min_n = (int)v;
    else if (n == "gemm_tc_min_work") ctx->gemm_tc_min_work = v < 0 ? 0 : v;
    else if (n == "gemm_tc_cluster") ctx->gemm_tc_cluster = (int)v;
    else if (n == "gemm_tc_precision") ctx->gemm_tc_precision = v < 0 ? 0 : v > 2 ? 2 : (int)v;
    else return fail(ctx, LK_EINVAL_ARG, "lk_set_option: unknown option " + n);
    return LK_OK;
}

int64_t lk_get_option(lk_ctx *ctx, const char *name) {
    if (!ctx || !name) return -1;
    const std::string n(name);
    if (n == "gemv_pdl") return ctx->pdl_mode;
    if (n == "gemv_warps") return ctx->tune.variant;
    if (n == "gemv_stage_bytes") return ctx->tune.stage_bytes;
    if (n == "gemv_row_quant") return ctx->tune.row_quant;
    if (n == "gemv_max_stages") return ctx->tune.max_stages;
    if (n == "gemv_smem_budget") return ctx->tune.smem_budget;
    if (n == "gemv_ctas_per_sm") return ctx->tune.ctas_per_sm;
    if (n == "gemv_overlap") return ctx->overlap_mode;
    if (n == "gemv_chain") return ctx->chain_mode;
    if (n == "gemv_chain_mma") return ctx->chain_mma;
    if (n == "gemv_chain_launches") return ctx->chain_launches;
    if (n == "gemv_chain_mma_launches") return ctx->chain_mma_launches;
    if (n == "gemv_chain_timeouts") {   // grid-barrier time-outs since the context was created (0 in a healthy run); reading synchronises
        unsigned w[4] = {0, 0, 0, 0};
        if (cudaStreamSynchronize(ctx->stream) != cudaSuccess || cudaMemcpy(w, ctx->chain_bar, 16, cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
        if (w[2]) cudaMemset(ctx->chain_bar, 0, lk::CHAIN_SCRATCH_BYTES);   // a timed-out barrier leaves its arrival count behind
        ctx->chain_unchecked = 0;
        const int64_t seen = ctx->chain_timeouts_seen + (int64_t)w[2];   // incl. those lk_sync already reported
        ctx->chain_timeouts_seen = 0;
        return seen;
    }
    if (n == "gemv_act_planes") return ctx->tune.act_planes;
    if (n == "gemv_bitnet_copies") return ctx->tune.bitnet_copies;
    if (n == "nibble_order") return ctx->tune.nibble_order;
    if (n == "gemv_fix_gap") return ctx->tune.fix_gap;
    if (n == "gemm_tc_min_n") return ctx->gemm_tc_min_n;
    if (n == "gemm_tc_min_work") return ctx->gemm_tc_min_work;
    if (n == "gemm_tc_cluster") return ctx->gemm_tc_cluster;
    if (n == "gemm_tc_launches") return ctx->gemm_tc_launches;
    if (n == "gemm_tc_precision") return ctx->gemm_tc_precision;
    return -1;
}

}  // extern "C"

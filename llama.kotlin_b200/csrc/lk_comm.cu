// lk_comm.cu — multi-GPU: weight rows sharded over ranks (one process per GPU),
// F32 outputs gathered either by ncclAllGather (baseline) or by direct NVLink
// peer stores from the GEMV epilogue followed by a flag barrier (fused path).
// NCCL is bound at run time (dlopen) so the library loads without it; inside a
// torch process the already-loaded libnccl.so.2 is the one that resolves.
#include <dlfcn.h>
#include <string.h>

#include "lk_common.cuh"
#include "lk_shim.h"

namespace {
typedef struct { char internal[128]; } nccl_uid;
typedef void *nccl_comm;
typedef int (*fn_get_uid)(nccl_uid *);
typedef int (*fn_init_rank)(nccl_comm *, int, nccl_uid, int);
typedef int (*fn_allgather)(const void *, void *, size_t, int, nccl_comm, cudaStream_t);
typedef int (*fn_destroy)(nccl_comm);
typedef const char *(*fn_errstr)(int);
struct Nccl {
    void *h = nullptr;
    fn_get_uid get_uid = nullptr; fn_init_rank init_rank = nullptr; fn_allgather allgather = nullptr;
    fn_destroy destroy = nullptr; fn_errstr errstr = nullptr;
    bool tried = false;
} g_nccl;

bool nccl_load(std::string &why) {
    if (g_nccl.h) return true;
    if (g_nccl.tried) { why = "libnccl not found"; return false; }
    g_nccl.tried = true;
    const char *names[] = {"libnccl.so.2", "libnccl.so", nullptr};
    for (int i = 0; names[i] && !g_nccl.h; i++) g_nccl.h = dlopen(names[i], RTLD_NOW | RTLD_GLOBAL);
    if (!g_nccl.h) { why = std::string("dlopen(libnccl.so.2) failed: ") + dlerror(); return false; }
    g_nccl.get_uid = (fn_get_uid)dlsym(g_nccl.h, "ncclGetUniqueId");
    g_nccl.init_rank = (fn_init_rank)dlsym(g_nccl.h, "ncclCommInitRank");
    g_nccl.allgather = (fn_allgather)dlsym(g_nccl.h, "ncclAllGather");
    g_nccl.destroy = (fn_destroy)dlsym(g_nccl.h, "ncclCommDestroy");
    g_nccl.errstr = (fn_errstr)dlsym(g_nccl.h, "ncclGetErrorString");
    if (!g_nccl.get_uid || !g_nccl.init_rank || !g_nccl.allgather || !g_nccl.destroy) {
        why = "libnccl is missing required symbols"; g_nccl.h = nullptr; return false;
    }
    return true;
}
lk_status nccl_fail(lk_ctx *ctx, int rc, const char *what) {
    return lk::fail(ctx, LK_ENCCL, std::string("NCCL error in ") + what + ": " + (g_nccl.errstr ? g_nccl.errstr(rc) : "?"));
}

// One warp: publish "my slice is written" to every peer, then wait until every
// peer has published the same epoch locally.  Runs after the GEMV on the same
// stream, so the GEMV's peer stores are complete; the system-scope fence orders
// them before the flag stores.  Each rank owns its GPU, so the spin is safe; it
// is bounded (~2 s) so a dead peer cannot hang the device: on time-out word 14
// of the flag buffer is set and the kernel returns.
// epoch == 0: the epoch is a device-side counter (word 15), so the same captured
// launch can be replayed; every rank must then issue the same number of barriers.
__global__ void peer_barrier_kernel(uint32_t *own, uint32_t *const *peers, int n, int rank, uint32_t epoch) {
    const int q = threadIdx.x;
    if (epoch == 0) {
        uint32_t e = 0;
        if (q == 0) { e = own[15] + 1; own[15] = e; }
        epoch = __shfl_sync(0xffffffffu, e, 0);
    }
    __threadfence_system();
    if (q < n) {
        volatile uint32_t *f = (q == rank ? own : peers[q]) + rank;
        *f = epoch;
    }
    __threadfence_system();
    if (q < n) {
        volatile uint32_t *f = own + q;
        const long long t0 = clock64();
        while ((int32_t)(*f - epoch) < 0) {
            if (clock64() - t0 > 4000000000ll) { ((volatile uint32_t *)own)[14] = 1; break; }
        }
    }
    __threadfence_system();
}
}  // namespace

extern "C" {

lk_status lk_comm_unique_id(uint8_t id_out[LK_COMM_ID_BYTES]) {
    std::string why;
    if (!nccl_load(why)) return lk::fail(nullptr, LK_ENCCL, why);
    nccl_uid id;
    int rc = g_nccl.get_uid(&id);
    if (rc) return nccl_fail(nullptr, rc, "ncclGetUniqueId");
    memcpy(id_out, id.internal, LK_COMM_ID_BYTES);
    return LK_OK;
}

lk_status lk_comm_init(lk_ctx *ctx, int32_t nranks, int32_t rank, const uint8_t id[LK_COMM_ID_BYTES]) {
    if (!ctx || !id || nranks < 1 || rank < 0 || rank >= nranks) return lk::fail(ctx, LK_EINVAL_ARG, "lk_comm_init: bad argument");
    std::string why;
    if (!nccl_load(why)) return lk::fail(ctx, LK_ENCCL, why);
    LK_CUDA(ctx, cudaSetDevice(ctx->device));
    nccl_uid uid;
    memcpy(uid.internal, id, LK_COMM_ID_BYTES);
    nccl_comm c = nullptr;
    int rc = g_nccl.init_rank(&c, nranks, uid, rank);
    if (rc) return nccl_fail(ctx, rc, "ncclCommInitRank");
    ctx->comm = c; ctx->nranks = nranks; ctx->rank = rank;
    return LK_OK;
}

lk_status lk_comm_destroy(lk_ctx *ctx) {
    if (ctx && ctx->comm && g_nccl.destroy) { g_nccl.destroy((nccl_comm)ctx->comm); ctx->comm = nullptr; }
    if (ctx && ctx->peer_flags) { cudaFree(ctx->peer_flags); ctx->peer_flags = nullptr; }
    return LK_OK;
}

lk_status lk_allgather_f32(lk_ctx *ctx, lk_buf *buf, uint64_t off, int64_t count) {
    if (!ctx || !buf) return lk::fail(ctx, LK_EINVAL_ARG, "lk_allgather_f32: NULL argument");
    if (!ctx->comm) return lk::fail(ctx, LK_ENCCL, "lk_allgather_f32: lk_comm_init has not been called");
    const uint64_t total = (uint64_t)count * 4 * ctx->nranks;
    if (off > buf->size || total > buf->size - off) return lk::fail(ctx, LK_EBOUNDS, "lk_allgather_f32: range outside the buffer");
    float *base = (float *)((uint8_t *)buf->ptr + off);
    lk::chain_other(ctx, base, (size_t)total);
    int rc = g_nccl.allgather(base + (int64_t)ctx->rank * count, base, (size_t)count, 7 /* ncclFloat */, (nccl_comm)ctx->comm, ctx->stream);
    if (rc) return nccl_fail(ctx, rc, "ncclAllGather");
    return LK_OK;
}

lk_status lk_buf_ipc_export(lk_buf *buf, uint8_t handle_out[LK_IPC_HANDLE_BYTES]) {
    if (!buf || !handle_out) return lk::fail(nullptr, LK_EINVAL_ARG, "lk_buf_ipc_export: NULL argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == LK_IPC_HANDLE_BYTES, "IPC handle size");
    cudaIpcMemHandle_t h;
    LK_CUDA(buf->ctx, cudaIpcGetMemHandle(&h, buf->ptr));
    memcpy(handle_out, &h, sizeof(h));
    return LK_OK;
}

lk_status lk_buf_ipc_import(lk_ctx *ctx, const uint8_t handle[LK_IPC_HANDLE_BYTES], uint64_t size, lk_buf **out) {
    if (!ctx || !handle || !out) return lk::fail(ctx, LK_EINVAL_ARG, "lk_buf_ipc_import: NULL argument");
    LK_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    void *p = nullptr;
    LK_CUDA(ctx, cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    lk_buf *b = new lk_buf();
    b->ctx = ctx; b->ptr = p; b->size = size; b->imported = true;
    *out = b;
    return LK_OK;
}

lk_status lk_peer_barrier(lk_ctx *ctx, lk_buf *own_flags, lk_buf *const *peer_flags, int32_t n, int32_t rank, uint32_t epoch) {
    if (!ctx || !own_flags || !peer_flags || n < 1 || n > 8 || rank < 0 || rank >= n)
        return lk::fail(ctx, LK_EINVAL_ARG, "lk_peer_barrier: bad argument");
    if (own_flags->size < 64) return lk::fail(ctx, LK_EBOUNDS, "lk_peer_barrier: flag buffer must hold 16 words");
    uint32_t *ptrs[8] = {};
    for (int q = 0; q < n; q++) ptrs[q] = (uint32_t *)(q == rank ? own_flags->ptr : (peer_flags[q] ? peer_flags[q]->ptr : nullptr));
    for (int q = 0; q < n; q++) if (!ptrs[q]) return lk::fail(ctx, LK_EINVAL_ARG, "lk_peer_barrier: missing peer flag buffer");
    if (!ctx->peer_flags || memcmp(ptrs, ctx->peer_flags_host, sizeof(ptrs)) != 0) {
        // the pointer table is uploaded once (synchronously, never inside a captured sequence)
        if (ctx->capturing) return lk::fail(ctx, LK_EINVAL_ARG, "lk_peer_barrier: call once outside capture first (sets up the peer table)");
        if (!ctx->peer_flags) LK_CUDA(ctx, cudaMalloc((void **)&ctx->peer_flags, sizeof(ptrs)));
        LK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        LK_CUDA(ctx, cudaMemcpy(ctx->peer_flags, ptrs, sizeof(ptrs), cudaMemcpyHostToDevice));
        memcpy(ctx->peer_flags_host, ptrs, sizeof(ptrs));
    }
    lk::chain_other(ctx, own_flags->ptr, own_flags->size);
    ctx->barrier_flags = (uint32_t *)own_flags->ptr;
    ctx->barrier_flags_bytes = own_flags->size;
    ctx->barrier_unchecked++;
    peer_barrier_kernel<<<1, 32, 0, ctx->stream>>>((uint32_t *)own_flags->ptr, (uint32_t *const *)ctx->peer_flags, n, rank, epoch);
    LK_CUDA(ctx, cudaGetLastError());
    if (ctx->capturing) ctx->capture_launches++; else ctx->launches++;
    return LK_OK;
}


// ---- exchange points inside the flag buffers (lk_shard_sync, include/llamakotlin_cuda.h) ------------------------------
// bytes 0..63: the 16 words of lk_peer_barrier; 64 + 8 s: arrival counter of slot s (the only words peers write);
// 128 + 8 s: local CTA counter of the signalling launch; 192 + 8 s: local ticket counter of the waiting launch; 256: time-outs
static lk_status build_sync(lk_ctx *ctx, const lk_shard_sync *sy, lk::GemvSync *out, const char *who) {
    if (!sy) return LK_OK;
    if (!sy->peer_flags || sy->nranks < 1 || sy->nranks > 8 || sy->rank < 0 || sy->rank >= sy->nranks)
        return lk::fail(ctx, LK_EINVAL_ARG, std::string(who) + ": bad lk_shard_sync");
    if (sy->wait_slot >= LK_MAX_EXCHANGE_SLOTS || sy->signal_slot >= LK_MAX_EXCHANGE_SLOTS)
        return lk::fail(ctx, LK_EINVAL_ARG, std::string(who) + ": exchange slot out of range");
    for (int q = 0; q < sy->nranks; q++)
        if (!sy->peer_flags[q] || sy->peer_flags[q]->size < LK_FLAG_BYTES)
            return lk::fail(ctx, LK_EBOUNDS, std::string(who) + ": flag buffers must hold LK_FLAG_BYTES");
    uint8_t *own = (uint8_t *)sy->peer_flags[sy->rank]->ptr;
    out->timeout = (unsigned *)(own + 256);
    if (sy->signal_slot >= 0) {
        out->sig_n = sy->nranks;
        for (int q = 0; q < sy->nranks; q++) out->sig_peer[q] = (unsigned long long *)((uint8_t *)sy->peer_flags[q]->ptr + 64 + 8 * sy->signal_slot);
        out->sig_done = (unsigned long long *)(own + 128 + 8 * sy->signal_slot);
    }
    if (sy->wait_slot >= 0) {
        out->wait_n = sy->nranks;
        out->wait_ctr = (unsigned long long *)(own + 64 + 8 * sy->wait_slot);
        out->wait_ticket = (unsigned long long *)(own + 192 + 8 * sy->wait_slot);
    }
    ctx->barrier_flags = (uint32_t *)own;
    ctx->barrier_flags_bytes = sy->peer_flags[sy->rank]->size;
    ctx->barrier_unchecked++;
    return LK_OK;
}

struct PushPeers { float *ptr[8]; int n; };
__global__ void __launch_bounds__(256) peer_push_kernel(const float *__restrict__ src, int64_t n, PushPeers peers) {
    lk::pdl_wait();                // the slice may be the previous kernel's output
    lk::pdl_launch_dependents();
    const int64_t nthr = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += nthr) {
        const float v = src[i];
        for (int q = 0; q < peers.n; q++) peers.ptr[q][i] = v;
    }
}

lk_status lk_peer_push_f32(lk_ctx *ctx, lk_buf *local, uint64_t off, int64_t count, lk_buf *const *peer_bufs, int32_t npeers) {
    if (!ctx || !local || count < 0 || npeers < 0 || npeers > 8) return lk::fail(ctx, LK_EINVAL_ARG, "lk_peer_push_f32: bad argument");
    if ((off & 3) || off > local->size || (uint64_t)count * 4 > local->size - off) return lk::fail(ctx, LK_EBOUNDS, "lk_peer_push_f32: range outside the buffer");
    PushPeers pp = {};
    for (int q = 0; q < npeers; q++) {
        if (!peer_bufs || !peer_bufs[q] || peer_bufs[q] == local) continue;
        if (off + (uint64_t)count * 4 > peer_bufs[q]->size) return lk::fail(ctx, LK_EBOUNDS, "lk_peer_push_f32: peer buffer too small");
        pp.ptr[pp.n++] = (float *)((uint8_t *)peer_bufs[q]->ptr + off);
    }
    if (count == 0 || pp.n == 0) return LK_OK;
    LK_CUDA(ctx, cudaSetDevice(ctx->device));
    const float *src = (const float *)((uint8_t *)local->ptr + off);
    lk::chain_other(ctx, nullptr, 0);      // writes land in peer memory only
    const bool pdl = (ctx->pdl_mode == 2) || (ctx->pdl_mode == 1 && ctx->capturing);
    int grid = (int)((count + 255) / 256);
    if (grid > ctx->sm_count) grid = ctx->sm_count;
    cudaError_t e = lk::launch_kernel(peer_push_kernel, dim3(grid), dim3(256), 0, ctx->stream, pdl, src, count, pp);
    if (e != cudaSuccess) return lk::cuda_fail(ctx, e, "peer push launch");
    if (ctx->capturing) ctx->capture_launches++; else ctx->launches++;
    return LK_OK;
}

lk_status lk_compute_matmul_sharded(lk_ctx *ctx, const lk_tensor *a, const lk_tensor *b, const lk_tensor *dst,
                                    int64_t row0, int64_t M_total, lk_buf *const *peer_dst, int32_t npeers) {
    return lk_compute_matmul_sharded_sync(ctx, a, b, dst, row0, M_total, peer_dst, npeers, nullptr);
}

lk_status lk_compute_matmul_sharded_sync(lk_ctx *ctx, const lk_tensor *a, const lk_tensor *b, const lk_tensor *dst,
                                         int64_t row0, int64_t M_total, lk_buf *const *peer_dst, int32_t npeers,
                                         const lk_shard_sync *sync) {
    if (!ctx || !a || !b || !dst) return lk::fail(ctx, LK_EINVAL_ARG, "lk_compute_matmul_sharded: NULL argument");
    const int64_t Ms = a->ne[1], K = a->ne[0];
    if (K != b->ne[1]) return lk::fail(ctx, LK_EINVAL_DIM, "Dim mismatch K: a.ne[0] != b.ne[1]");
    if (b->ne[0] != 1 || dst->ne[0] != 1) return lk::fail(ctx, LK_EINVAL_DIM, "sharded matmul supports N == 1");
    if (dst->ne[1] != M_total || row0 < 0 || row0 + Ms > M_total) return lk::fail(ctx, LK_EINVAL_DIM, "shard rows outside dst");
    if (lk::block_bytes(a->type) == 0 || b->type != LK_TYPE_F32 || dst->type != LK_TYPE_F32)
        return lk::fail(ctx, LK_EINVAL_TYPE, "sharded matmul: Q4_0/Q4_1/Q8_0 x F32 -> F32");
    if (!a->dev_buf || !b->dev_buf || !dst->dev_buf) return lk::fail(ctx, LK_EINVAL_ARG, "sharded matmul: tensors must be device resident");
    if (npeers < 0 || npeers > 8) return lk::fail(ctx, LK_EINVAL_ARG, "sharded matmul: at most 8 peers");
    const int64_t abytes = lk_tensor_nbytes(a);
    if (a->data_offset + (uint64_t)abytes > a->dev_buf->size || b->data_offset + (uint64_t)K * 4 > b->dev_buf->size ||
        dst->data_offset + (uint64_t)M_total * 4 > dst->dev_buf->size)
        return lk::fail(ctx, LK_EBOUNDS, "sharded matmul: tensor bytes exceed the device buffer");
    lk::GemvPeers peers;
    for (int q = 0; q < npeers; q++) {
        if (!peer_dst || !peer_dst[q] || peer_dst[q] == dst->dev_buf) continue;
        if (dst->data_offset + (uint64_t)M_total * 4 > peer_dst[q]->size) return lk::fail(ctx, LK_EBOUNDS, "peer dst too small");
        peers.ptr[peers.n++] = (float *)((uint8_t *)peer_dst[q]->ptr + dst->data_offset);
    }
    lk::GemvTuning tune = ctx->tune;
    tune.pdl = (ctx->pdl_mode == 2) || (ctx->pdl_mode == 1 && ctx->capturing);
    float *d = (float *)((uint8_t *)dst->dev_buf->ptr + dst->data_offset);
    // this launch writes rows [row0, row0 + Ms) of dst here and at the same offsets on every peer
    tune.wait_mode = lk::gemv_wait_mode(ctx, (uint8_t *)a->dev_buf->ptr + a->data_offset, (size_t)abytes,
                                        (uint8_t *)b->dev_buf->ptr + b->data_offset, (size_t)K * 4, d + row0, (size_t)Ms * 4, tune.pdl != 0);
    lk::GemvSync gs;
    const lk_status sst = build_sync(ctx, sync, &gs, "sharded matmul");
    if (sst != LK_OK) return sst;
    if (gs.wait_n && tune.wait_mode == 2) tune.wait_mode = 1;      // a launch that waits for its peers is not independent
    cudaError_t e = lk::launch_gemv(a->type, (uint8_t *)a->dev_buf->ptr + a->data_offset, K, Ms,
                                    (const float *)((uint8_t *)b->dev_buf->ptr + b->data_offset), d, row0, peers, tune,
                                    ctx->sm_count, ctx->stream, nullptr, sync ? &gs : nullptr);
    if (e == cudaErrorNotSupported) return lk::fail(ctx, LK_ENOTIMPL, "sharded matmul: shape not supported by the streaming GEMV");
    if (e != cudaSuccess) return lk::cuda_fail(ctx, e, "sharded gemv launch");
    if (ctx->capturing) ctx->capture_launches++; else ctx->launches++;
    return LK_OK;
}

lk_status lk_compute_swiglu_sharded(lk_ctx *ctx, const lk_tensor *gate, const lk_tensor *up, const lk_tensor *x,
                                    const lk_tensor *norm_w, float eps, const lk_tensor *normed_dst, const lk_tensor *g_full,
                                    const lk_tensor *u_full, const lk_tensor *silu_full, const lk_tensor *h_full, int64_t row0,
                                    int64_t M_total, lk_buf *const *peer_h, int32_t npeers) {
    return lk_compute_swiglu_sharded_sync(ctx, gate, up, x, norm_w, eps, normed_dst, g_full, u_full, silu_full, h_full, row0, M_total,
                                          peer_h, npeers, nullptr);
}

lk_status lk_compute_swiglu_sharded_sync(lk_ctx *ctx, const lk_tensor *gate, const lk_tensor *up, const lk_tensor *x,
                                         const lk_tensor *norm_w, float eps, const lk_tensor *normed_dst, const lk_tensor *g_full,
                                         const lk_tensor *u_full, const lk_tensor *silu_full, const lk_tensor *h_full, int64_t row0,
                                         int64_t M_total, lk_buf *const *peer_h, int32_t npeers, const lk_shard_sync *sync) {
    if (!ctx || !gate || !up || !x || !g_full || !u_full || !h_full) return lk::fail(ctx, LK_EINVAL_ARG, "lk_compute_swiglu_sharded: NULL argument");
    const int64_t Ms = gate->ne[1], K = gate->ne[0];
    if (up->type != gate->type || up->ne[0] != K || up->ne[1] != Ms) return lk::fail(ctx, LK_EINVAL_DIM, "sharded SwiGLU: gate and up shards must have one type and shape");
    if (lk::block_bytes(gate->type) == 0) return lk::fail(ctx, LK_EINVAL_TYPE, "sharded SwiGLU: Q4_0/Q4_1/Q8_0 weights");
    if (row0 < 0 || row0 + Ms > M_total) return lk::fail(ctx, LK_EINVAL_DIM, "shard rows outside dst");
    if (npeers < 0 || npeers > 8) return lk::fail(ctx, LK_EINVAL_ARG, "sharded SwiGLU: at most 8 peers");
    auto count = [](const lk_tensor *t) { int64_t n = 1; for (int d = 0; d < LK_MAX_DIMS; d++) n *= t->ne[d] > 0 ? t->ne[d] : 1; return n; };
    struct Chk { const lk_tensor *t; int64_t n; };
    const Chk chk[] = {{x, K}, {norm_w, K}, {normed_dst, K}, {g_full, M_total}, {u_full, M_total}, {silu_full, M_total}, {h_full, M_total}};
    for (const Chk &c : chk) {
        if (!c.t) continue;
        if (c.t->type != LK_TYPE_F32) return lk::fail(ctx, LK_EINVAL_TYPE, "sharded SwiGLU: F32 vectors expected");
        if (!c.t->dev_buf) return lk::fail(ctx, LK_EINVAL_ARG, "sharded SwiGLU: tensors must be device resident");
        if (count(c.t) != c.n) return lk::fail(ctx, LK_EINVAL_DIM, "sharded SwiGLU: vector length mismatch");
        if (c.t->data_offset + (uint64_t)c.n * 4 > c.t->dev_buf->size) return lk::fail(ctx, LK_EBOUNDS, "sharded SwiGLU: tensor bytes exceed the device buffer");
    }
    if (normed_dst && !norm_w) return lk::fail(ctx, LK_EINVAL_ARG, "sharded SwiGLU: normed_dst goes with norm_w");
    for (const lk_tensor *t : {gate, up}) {
        if (!t->dev_buf) return lk::fail(ctx, LK_EINVAL_ARG, "sharded SwiGLU: tensors must be device resident");
        if (t->data_offset + (uint64_t)lk_tensor_nbytes(t) > t->dev_buf->size) return lk::fail(ctx, LK_EBOUNDS, "sharded SwiGLU: weight bytes exceed the device buffer");
    }
    auto ptr = [](const lk_tensor *t) { return t ? (uint8_t *)t->dev_buf->ptr + t->data_offset : nullptr; };
    lk::GemvPeers peers;
    for (int q = 0; q < npeers; q++) {
        if (!peer_h || !peer_h[q] || peer_h[q] == h_full->dev_buf) continue;
        if (h_full->data_offset + (uint64_t)M_total * 4 > peer_h[q]->size) return lk::fail(ctx, LK_EBOUNDS, "peer dst too small");
        peers.ptr[peers.n++] = (float *)((uint8_t *)peer_h[q]->ptr + h_full->data_offset);
    }
    lk::GemvFusion fu;
    if (norm_w) { fu.pro_mode = 1; fu.pro_w = (const float *)ptr(norm_w); fu.pro_eps = eps; fu.b_out = (float *)ptr(normed_dst); }
    fu.nseg_extra = 1;
    fu.a_extra[0] = ptr(up);
    fu.dst_extra[0] = (float *)ptr(u_full);
    fu.epi_mode = 1;
    fu.epi_act = (float *)ptr(silu_full);
    fu.epi_out = (float *)ptr(h_full);
    lk::GemvTuning tune = ctx->tune;
    tune.pdl = (ctx->pdl_mode == 2) || (ctx->pdl_mode == 1 && ctx->capturing);
    const size_t abytes = (size_t)lk_tensor_nbytes(gate);
    std::vector<std::pair<const void *, size_t>> rd, wr, wt;
    if (norm_w) rd.push_back({ptr(norm_w), (size_t)K * 4});
    if (normed_dst) wr.push_back({ptr(normed_dst), (size_t)K * 4});
    wr.push_back({ptr(u_full) + row0 * 4, (size_t)Ms * 4});
    wr.push_back({ptr(h_full) + row0 * 4, (size_t)Ms * 4});
    if (silu_full) wr.push_back({ptr(silu_full) + row0 * 4, (size_t)Ms * 4});
    wt.push_back({ptr(up), abytes});
    tune.wait_mode = lk::gemv_wait_mode(ctx, ptr(gate), abytes, ptr(x), (size_t)K * 4, ptr(g_full) + row0 * 4, (size_t)Ms * 4,
                                        tune.pdl != 0, &rd, &wr, &wt);
    // the pair keeps both segments' partial sums of a CTA's rows in shared memory: a shard with many rows per CTA
    // (G = 2 of the 70B FFN: 97 rows x 8192) is spread over 2 or 4 CTAs per SM instead
    lk::GemvSync gs;
    const lk_status sst = build_sync(ctx, sync, &gs, "sharded SwiGLU");
    if (sst != LK_OK) return sst;
    if (gs.wait_n && tune.wait_mode == 2) tune.wait_mode = 1;      // a launch that waits for its peers is not independent
    cudaError_t e = cudaErrorNotSupported;
    for (int cps = tune.ctas_per_sm > 0 ? tune.ctas_per_sm : 1; cps <= 4 && e == cudaErrorNotSupported; cps *= 2) {
        tune.ctas_per_sm = cps;
        e = lk::launch_gemv(gate->type, ptr(gate), K, Ms, (const float *)ptr(x), (float *)ptr(g_full), row0, peers, tune,
                            ctx->sm_count, ctx->stream, &fu, sync ? &gs : nullptr);
        if (e == cudaErrorNotSupported) cudaGetLastError();
    }
    if (e == cudaErrorNotSupported) return lk::fail(ctx, LK_ENOTIMPL, "sharded SwiGLU: shape not supported by the streaming GEMV");
    if (e != cudaSuccess) return lk::cuda_fail(ctx, e, "sharded SwiGLU launch");
    if (ctx->capturing) ctx->capture_launches++; else ctx->launches++;
    return LK_OK;
}

}  // extern "C"

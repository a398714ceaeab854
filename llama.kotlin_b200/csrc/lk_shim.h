// lk_shim.h — internal structs behind the opaque handles of llamakotlin_cuda.h
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <utility>
#include <vector>

#include "../../include/llamakotlin_cuda.h"
#include "lk_kernels.h"

struct lk_ctx {
    int device = 0;
    int sm_count = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    // lk_buf_upload_async / lk_buf_download_async: one stream per direction beside the compute stream (created on first
    // use), one event per direction (recorded and waited on back to back), the fork event of a capture
    cudaStream_t copy_in = nullptr, copy_out = nullptr;
    cudaEvent_t ev_in = nullptr, ev_out = nullptr, ev_root = nullptr;
    bool cap_in = false, cap_out = false;      // the copy streams have joined the capture in progress
    bool copies_pending = false;               // asynchronous copies were issued since the last lk_sync
    std::string err;
    int64_t launches = 0;
    lk::GemvTuning tune;
    int pdl_mode = 1;  // 0 never, 1 inside captured sequences, 2 always
    bool capturing = false;
    int64_t capture_launches = 0;
    // Byte ranges touched by the GEMV launches that may still be in flight when the next one starts
    // (programmatic dependent launch chain); see gemv_wait_mode() in lk_shim.cu.
    std::vector<std::pair<uintptr_t, uintptr_t>> chain_reads, chain_writes;
    std::vector<std::pair<uintptr_t, uintptr_t>> other_writes;   // outputs of non-GEMV ops since the last GEMV
    uint32_t *chain_bar = nullptr;   // K1c grid barrier: arrivals, generation, time-outs
    int chain_mode = 1;              // 0: lk_compute_matmul_chain declines (the caller issues the links one by one)
    int64_t chain_launches = 0;
    int chain_mma = 1;               // 1: chains go to K1m (lk_gemv_mma.cu) where it takes them, else to K1c
    int64_t chain_mma_launches = 0;
    int64_t chain_unchecked = 0;     // chained launches since the time-out word was last read (lk::check_async_faults)
    uint32_t *barrier_flags = nullptr;   // this rank's flag words of lk_peer_barrier (word 14 = time-out), NULL before the first barrier
    uint64_t barrier_flags_bytes = 0;
    int64_t barrier_unchecked = 0;
    int64_t capture_chain0 = 0, capture_barrier0 = 0;   // the two counters at lk_capture_begin
    int64_t chain_timeouts_seen = 0;     // time-outs already reported through a status, not yet read through the option
    int overlap_mode = 1;  // 0: every GEMV waits for its predecessor before reading x; 1: independent GEMVs overlap
    // device staging arena for host-resident tensors (bump allocated per op)
    uint8_t *arena = nullptr;
    size_t arena_size = 0, arena_used = 0;
    // K2 workspace (FP16 copy of the activations)
    void *gemm_ws = nullptr;
    size_t gemm_ws_size = 0;
    std::vector<void *> gemm_ws_retired;   // smaller workspaces that captured graphs may still name; freed with the ctx
    // decode-attention workspace (split partials + per-head tickets); same retire rule as gemm_ws
    void *attn_ws = nullptr;
    size_t attn_ws_size = 0;
    std::vector<void *> attn_ws_retired;
    int gemm_tc_min_n = 2;                 // N >= this takes the tcgen05 GEMM (0 = never)
    int64_t gemm_tc_min_work = 1 << 21;    // ... if M * K * N is at least this many multiply-adds (three launches must pay)
    int gemm_tc_precision = 2;             // FP16 parts per operand of the tcgen05 GEMM: 2 = A hi+lo x b hi+lo (F32-grade, default), 1 = A hi+lo x b, 0 = single
    int gemm_tc_cluster = 1;               // CTAs per cluster sharing an activation tile by multicast (1, 2, 4); 1 measured fastest
    int64_t gemm_tc_launches = 0;          // tcgen05 GEMM launches so far (tests assert the path taken)
    // multi-GPU
    void *comm = nullptr;  // ncclComm_t
    int nranks = 1, rank = 0;
    uint32_t *peer_flags = nullptr;          // device copy of the peer flag pointer table
    void *peer_flags_host[8] = {};           // what that table holds
};

struct lk_buf {
    lk_ctx *ctx = nullptr;
    void *ptr = nullptr;
    uint64_t size = 0;
    bool imported = false;  // opened from an IPC handle (peer memory)
};

struct lk_graph {
    lk_ctx *ctx = nullptr;
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    int64_t launches = 0;  // kernels per replay
    int64_t chain_launches = 0, barrier_launches = 0;   // chained GEMVs / peer barriers per replay (their time-out words are checked at the next sync)
};

namespace lk {
lk_status fail(lk_ctx *ctx, lk_status st, const std::string &msg);
lk_status cuda_fail(lk_ctx *ctx, cudaError_t e, const char *what);
void chain_other(lk_ctx *ctx, const void *dst, size_t bytes);
// After a synchronisation point: did a device-side bounded spin give up since the last check (grid barrier of the
// chained GEMV, cross-GPU flag barrier)?  Such a launch leaves wrong results behind, so every entry point that
// synchronises (lk_sync, lk_buf_download, lk_timer_stop, host-staged ops) returns LK_ECUDA once and re-arms the barrier.
lk_status check_async_faults(lk_ctx *ctx);
// where a PDL-launched GEMV must wait (GEMV_WAIT_*), and chain bookkeeping for it
int gemv_wait_mode(lk_ctx *ctx, const void *a, size_t abytes, const void *b, size_t bbytes, const void *d, size_t dbytes,
                   bool pdl, const std::vector<std::pair<const void *, size_t>> *more_reads = nullptr,
                   const std::vector<std::pair<const void *, size_t>> *more_writes = nullptr,
                   const std::vector<std::pair<const void *, size_t>> *more_weights = nullptr);   // a non-GEMV op wrote [dst, dst+bytes)
}  // namespace lk

#define LK_CUDA(ctx, call)                                             \
    do {                                                               \
        cudaError_t e__ = (call);                                      \
        if (e__ != cudaSuccess) return lk::cuda_fail(ctx, e__, #call); \
    } while (0)

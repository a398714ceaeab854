// lk_attention.cu — the attention core around the decode matmuls:
//   ROPE        LlamaAttention.applyRoPE            model/LlamaAttention.kt:29-72
//   KV append   KVCache.updateKey / updateValue     model/KVCache.kt:41-112   (cache ne = [headDim, numHeads, maxSeq])
//   SCALE       LlamaAttention.scale                model/LlamaAttention.kt:166-195
//   attention   LlamaAttention.computeAttention     model/LlamaAttention.kt:77-124: per head  softmax(q.K^T * 1/sqrt(d)) . V
//               with the softmax of :242-288 (max, exp(x - max), sum, divide)
//
// All of it is HBM-bound byte moving: per decoded token the attention reads 2 * len * hidden * 4 bytes of F32 cache
// once; RoPE / append touch a few KB.  The position lives in DEVICE memory (an int32 scalar, as ggml's `pos` tensor
// does) so that a captured CUDA graph of the layer can be replayed for every new token without re-capture.
//
// Decode attention (one query token), flash-decoding shape: grid = (heads, splits); a CTA owns `chunk` consecutive
// cache positions of one head, 8 lanes x 16 floats per K row (coalesced 512-byte rows, 3 shuffles per 4 rows), scores
// of the chunk in shared memory, local max / exp / sum, then thread d accumulates column d of V (coalesced rows
// again).  Partial (max, sum, acc[d]) go to a workspace; the last CTA of a head to arrive (atomic ticket) merges them
// with the usual exp(m_i - M) rescaling, so there is no second launch and no atomics on data (run-to-run bit
// identical for a given split).  CTAs whose chunk lies past the current length exit at once.
#include "lk_common.cuh"
#include "lk_kernels.h"

namespace lk {

// ---- ROPE: pairs (i, i+1) inside each head, angle = (pos + s) * theta^-(i / headDim) ---------------------------
// table[(p * (D/2) + i/2) * 2 + {0, 1}] = (float)cos(angle), (float)sin(angle), computed by the host in F64 exactly as
// applyRoPE does (lk_rope_table): the device does no transcendental and no FP64 work (vestigial on B200), and the
// result is bit-exact with the reference's arithmetic: x*cos - y*sin, x*sin + y*cos, each operation rounded once.
__global__ void __launch_bounds__(256) rope_kernel(const float *__restrict__ x, const float2 *__restrict__ table,
                                                   const int32_t *__restrict__ pos, int head_dim, int row, int ntok,
                                                   int max_pos, float *__restrict__ dst) {
    pdl_wait();
    pdl_launch_dependents();
    const int half = head_dim >> 1;
    const int npairs = (row >> 1) * ntok;
    const int p0 = pos ? pos[0] : 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < npairs; i += gridDim.x * blockDim.x) {
        const int s = i / (row >> 1), pr = i - s * (row >> 1);
        int p = p0 + s;
        p = p < 0 ? 0 : (p >= max_pos ? max_pos - 1 : p);     // host checks the range (KVCache / lk_compute_rope)
        const float2 cs = table[(size_t)p * half + (pr % half)];
        const float2 v = ((const float2 *)x)[i];
        float2 r;
        r.x = __fsub_rn(__fmul_rn(v.x, cs.x), __fmul_rn(v.y, cs.y));
        r.y = __fadd_rn(__fmul_rn(v.x, cs.y), __fmul_rn(v.y, cs.x));
        ((float2 *)dst)[i] = r;
    }
}

cudaError_t launch_rope(const float *x, const float *table, const int32_t *pos, int head_dim, int64_t row, int ntok,
                        int max_pos, float *dst, cudaStream_t s, bool pdl) {
    if (row == 0 || ntok == 0) return cudaSuccess;
    if (head_dim <= 0 || (head_dim & 1) || row % head_dim || (((uintptr_t)x | (uintptr_t)dst | (uintptr_t)table) & 7)) return cudaErrorNotSupported;
    const int64_t npairs = row / 2 * ntok;
    int grid = (int)((npairs + 255) / 256);
    if (grid > 148 * 8) grid = 148 * 8;
    return launch_kernel(rope_kernel, dim3(grid), dim3(256), 0, s, pdl, x, (const float2 *)table, pos, head_dim, (int)row, ntok, max_pos, dst);
}

// ---- KV append: cache[(pos + s) * row + e] = src[s * row + e]  (KVCache.updateKey :47-55, updateValue :88-96) ----
__global__ void __launch_bounds__(256) kv_append_kernel(const float *__restrict__ src, float *__restrict__ cache,
                                                        const int32_t *__restrict__ pos, int row, int ntok, int max_seq) {
    pdl_wait();
    pdl_launch_dependents();
    const int p0 = pos ? pos[0] : 0;
    const int n4 = (row >> 2) * ntok;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += gridDim.x * blockDim.x) {
        const int s = i / (row >> 2), e = i - s * (row >> 2);
        const int p = p0 + s;
        if (p < 0 || p >= max_seq) continue;                  // never write outside the cache
        ((float4 *)cache)[(size_t)p * (row >> 2) + e] = ((const float4 *)src)[i];
    }
}

cudaError_t launch_kv_append(const float *src, float *cache, const int32_t *pos, int64_t row, int ntok, int64_t max_seq,
                             cudaStream_t s, bool pdl) {
    if (row == 0 || ntok == 0) return cudaSuccess;
    if ((row & 3) || (((uintptr_t)src | (uintptr_t)cache) & 15)) return cudaErrorNotSupported;
    const int64_t n4 = row / 4 * ntok;
    int grid = (int)((n4 + 255) / 256);
    if (grid > 148 * 8) grid = 148 * 8;
    return launch_kernel(kv_append_kernel, dim3(grid), dim3(256), 0, s, pdl, src, cache, pos, (int)row, ntok, (int)max_seq);
}

// ---- decode attention -------------------------------------------------------------------------------------------
constexpr int ATT_THREADS = 128;
constexpr int ATT_MAX_SPLITS = 256;

struct AttnParams {
    const float *q;        // [D, H] (RoPE applied)
    const float *kc, *vc;  // [D, H, max_seq]
    const int32_t *pos;    // device scalar: index of the newest cache row; length = pos + 1
    float *dst;            // [D, H]
    float *ws;             // per (head, split): m, l, pad[2], acc[D]
    unsigned *tickets;     // per head, zero between launches
    int H, max_seq, chunk, nsplit;
    float scale;
    // fused decode step (lk_compute_attention_step): q_raw != NULL => q / k / v are this token's projections BEFORE RoPE;
    // the kernel rotates q and k itself (rope_kernel's arithmetic), appends k / v to the caches at `pos` and, where
    // asked, also stores the rotated vectors, i.e. it does the work of ROPE, ROPE, CPY, CPY, FLASH_ATTN in one launch
    const float *q_raw, *k_raw, *v_raw;
    const float2 *table;   // [table_pos][D/2] (cos, sin)
    float *q_rope, *k_rope;   // optional stores of the rotated q / k (the graph's intermediates)
    float *kc_w, *vc_w;    // the caches again, writable
    int table_pos;
};

__device__ __forceinline__ float2 rope_pair(float2 v, float2 cs) {   // applyRoPE, model/LlamaAttention.kt:60-66
    float2 r;
    r.x = __fsub_rn(__fmul_rn(v.x, cs.x), __fmul_rn(v.y, cs.y));
    r.y = __fadd_rn(__fmul_rn(v.x, cs.y), __fmul_rn(v.y, cs.x));
    return r;
}

template <int D>
__global__ void __launch_bounds__(ATT_THREADS) attn_decode_kernel(const AttnParams p) {
    constexpr int NV = D / 32;                 // float4 per lane per K row (8 lanes per row)
    extern __shared__ float sc[];              // [chunk] scores -> probabilities
    __shared__ float red[ATT_THREADS / 32];
    __shared__ int last_s;
    __shared__ __align__(16) float qs[D];      // fused step: this head's rotated q
    pdl_wait();
    pdl_launch_dependents();
    const int h = blockIdx.x, j = blockIdx.y;
    const int pos_raw = p.pos[0];
    int len = pos_raw + 1;
    len = len < 1 ? 1 : (len > p.max_seq ? p.max_seq : len);
    const int nact = (len + p.chunk - 1) / p.chunk;
    if (j >= nact) return;
    const int s0 = j * p.chunk, n = min(p.chunk, len - s0);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const size_t rowf = (size_t)p.H * D;       // floats between consecutive positions
    const bool fused = p.q_raw != nullptr;
    if (fused) {
        constexpr int HALF = D / 2;            // D <= ATT_THREADS: threads [0, HALF) rotate q, [HALF, D) rotate k
        const int pt = pos_raw < 0 ? 0 : (pos_raw >= p.table_pos ? p.table_pos - 1 : pos_raw);   // as rope_kernel clamps
        const bool owner = j == nact - 1 && pos_raw >= 0 && pos_raw < p.max_seq;   // my chunk holds this token's row
        if (tid < HALF) {
            const float2 r = rope_pair(((const float2 *)(p.q_raw + (size_t)h * D))[tid], p.table[(size_t)pt * HALF + tid]);
            ((float2 *)qs)[tid] = r;
            if (j == 0 && p.q_rope) ((float2 *)(p.q_rope + (size_t)h * D))[tid] = r;
        } else if (tid < D && owner) {
            const int e = tid - HALF;
            const float2 r = rope_pair(((const float2 *)(p.k_raw + (size_t)h * D))[e], p.table[(size_t)pt * HALF + e]);
            ((float2 *)(p.kc_w + (size_t)pos_raw * rowf + (size_t)h * D))[e] = r;     // KVCache.updateKey :47-55
            if (p.k_rope) ((float2 *)(p.k_rope + (size_t)h * D))[e] = r;
        }
        if (owner && tid < D / 4)                                                      // KVCache.updateValue :88-96
            ((float4 *)(p.vc_w + (size_t)pos_raw * rowf + (size_t)h * D))[tid] = ((const float4 *)(p.v_raw + (size_t)h * D))[tid];
        __syncthreads();                       // qs, and this CTA's own cache-row stores, are visible to all its threads
    }

    // pass 1: scores.  lane = (row-in-quad r, sub-lane g); row float4 index g + 8*i
    {
        const int g = lane & 7, r = lane >> 3;
        float4 qv[NV];
#pragma unroll
        for (int i = 0; i < NV; i++) qv[i] = fused ? ((const float4 *)qs)[g + 8 * i] : ((const float4 *)(p.q + (size_t)h * D))[g + 8 * i];
        const float *kbase = p.kc + (size_t)h * D;
#pragma unroll 2
        for (int tb = warp * 4; tb < n; tb += (ATT_THREADS / 32) * 4) {          // warp-uniform trip count (shuffles inside)
            const int t = tb + r;
            const bool on = t < n;
            const float4 *krow = (const float4 *)(kbase + (size_t)(s0 + (on ? t : 0)) * rowf);
            float acc = 0.f;
#pragma unroll
            for (int i = 0; i < NV; i++) {
                const float4 kv = __ldcs(krow + g + 8 * i);      // streamed once
                acc = fmaf(qv[i].x, kv.x, acc); acc = fmaf(qv[i].y, kv.y, acc);
                acc = fmaf(qv[i].z, kv.z, acc); acc = fmaf(qv[i].w, kv.w, acc);
            }
            acc += __shfl_xor_sync(0xffffffffu, acc, 1);
            acc += __shfl_xor_sync(0xffffffffu, acc, 2);
            acc += __shfl_xor_sync(0xffffffffu, acc, 4);
            if (on && g == 0) sc[t] = __fmul_rn(acc, p.scale);
        }
    }
    __syncthreads();
    // local max, exp, sum (LlamaAttention.softmax :262-276 on this chunk)
    float m = -INFINITY;
    for (int t = tid; t < n; t += ATT_THREADS) m = fmaxf(m, sc[t]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (lane == 0) red[warp] = m;
    __syncthreads();
    m = fmaxf(fmaxf(red[0], red[1]), fmaxf(red[2], red[3]));
    __syncthreads();
    float l = 0.f;
    for (int t = tid; t < n; t += ATT_THREADS) {
        const float e = expf(__fsub_rn(sc[t], m));
        sc[t] = e;
        l += e;
    }
    l = warp_sum(l);
    if (lane == 0) red[warp] = l;
    __syncthreads();
    l = (red[0] + red[1]) + (red[2] + red[3]);

    // pass 2: thread d accumulates column d of V over the chunk (rows are contiguous: coalesced)
    float acc = 0.f;
    if (tid < D) {
        const float *vcol = p.vc + (size_t)s0 * rowf + (size_t)h * D + tid;
        int t = 0;
        for (; t + 16 <= n; t += 16) {          // 16 rows in flight per thread: a 64-row chunk is 4 DRAM round trips, not 8
            float v[16];
#pragma unroll
            for (int u = 0; u < 16; u++) v[u] = __ldcs(vcol + (size_t)(t + u) * rowf);
#pragma unroll
            for (int u = 0; u < 16; u++) acc = fmaf(sc[t + u], v[u], acc);
        }
        for (; t + 4 <= n; t += 4) {
            float v[4];
#pragma unroll
            for (int u = 0; u < 4; u++) v[u] = __ldcs(vcol + (size_t)(t + u) * rowf);
#pragma unroll
            for (int u = 0; u < 4; u++) acc = fmaf(sc[t + u], v[u], acc);
        }
        for (; t < n; t++) acc = fmaf(sc[t], __ldcs(vcol + (size_t)t * rowf), acc);
    }
    if (nact == 1) {                            // short context: this CTA has the whole row
        if (tid < D) p.dst[(size_t)h * D + tid] = __fdiv_rn(acc, l);
        return;
    }
    float *w = p.ws + ((size_t)h * p.nsplit + j) * (D + 4);
    if (tid == 0) { w[0] = m; w[1] = l; }
    if (tid < D) w[4 + tid] = acc;
    __threadfence();
    __syncthreads();
    if (tid == 0) last_s = (atomicAdd(&p.tickets[h], 1u) == (unsigned)(nact - 1));
    __syncthreads();
    if (!last_s) return;
    __threadfence();
    // merge the partials of this head: M = max m_i, out = sum acc_i e^(m_i - M) / sum l_i e^(m_i - M)
    const float *wh = p.ws + (size_t)h * p.nsplit * (D + 4);
    float M = -INFINITY;
    for (int i = 0; i < nact; i++) M = fmaxf(M, __ldcg(wh + (size_t)i * (D + 4)));
    float num = 0.f, den = 0.f;
    if (tid < D) {
        for (int i = 0; i < nact; i++) {
            const float *wi = wh + (size_t)i * (D + 4);
            const float f = expf(__fsub_rn(__ldcg(wi), M));
            den = fmaf(__ldcg(wi + 1), f, den);
            num = fmaf(__ldcg(wi + 4 + tid), f, num);
        }
        p.dst[(size_t)h * D + tid] = __fdiv_rn(num, den);
    }
    if (tid == 0) p.tickets[h] = 0;             // ready for the next launch / graph replay
}

// chunk / split plan for a cache of max_seq positions (fixed at launch so that the grid of a captured graph is valid for
// every length up to max_seq)
static void attn_plan(int64_t max_seq, int *chunk, int *nsplit) {
    int c = 64;
    while ((max_seq + c - 1) / c > ATT_MAX_SPLITS) c *= 2;
    *chunk = c;
    *nsplit = (int)((max_seq + c - 1) / c);
}

size_t attn_workspace_bytes(int head_dim, int nheads, int64_t max_seq) {
    int chunk, nsplit;
    attn_plan(max_seq, &chunk, &nsplit);
    return (((size_t)nheads * 4 + 255) & ~(size_t)255) + (size_t)nheads * nsplit * (head_dim + 4) * 4;
}

cudaError_t launch_attn_decode(const float *q, const float *kc, const float *vc, const int32_t *pos, int head_dim,
                               int nheads, int64_t max_seq, float scale, float *dst, void *workspace, cudaStream_t s,
                               bool pdl, const AttnStep *step) {
    if (nheads <= 0 || max_seq <= 0 || max_seq > (1 << 24)) return cudaErrorNotSupported;
    if (head_dim != 64 && head_dim != 128) return cudaErrorNotSupported;
    if ((((uintptr_t)q | (uintptr_t)kc | (uintptr_t)vc | (uintptr_t)dst) & 15) || ((uintptr_t)workspace & 255)) return cudaErrorNotSupported;
    AttnParams p = {};
    if (step) {
        if ((((uintptr_t)step->k_raw | (uintptr_t)step->v_raw | (uintptr_t)step->q_rope | (uintptr_t)step->k_rope) & 15) ||
            ((uintptr_t)step->table & 7) || step->table_pos <= 0)
            return cudaErrorNotSupported;
        p.q_raw = q; p.k_raw = step->k_raw; p.v_raw = step->v_raw; p.table = (const float2 *)step->table;
        p.q_rope = step->q_rope; p.k_rope = step->k_rope; p.kc_w = (float *)kc; p.vc_w = (float *)vc; p.table_pos = step->table_pos;
    }
    p.q = q; p.kc = kc; p.vc = vc; p.pos = pos; p.dst = dst;
    p.H = nheads; p.max_seq = (int)max_seq; p.scale = scale;
    attn_plan(max_seq, &p.chunk, &p.nsplit);
    p.tickets = (unsigned *)workspace;                              // [H], zeroed when the workspace is created
    p.ws = (float *)((uint8_t *)workspace + (((size_t)nheads * 4 + 255) & ~(size_t)255));
    const dim3 grid((unsigned)nheads, (unsigned)p.nsplit);
    const size_t smem = (size_t)p.chunk * 4;
    if (head_dim == 128) return launch_kernel(attn_decode_kernel<128>, grid, dim3(ATT_THREADS), smem, s, pdl, p);
    return launch_kernel(attn_decode_kernel<64>, grid, dim3(ATT_THREADS), smem, s, pdl, p);
}

}  // namespace lk

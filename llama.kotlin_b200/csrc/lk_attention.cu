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
// cache positions of one head and walks them in 64-position tiles staged in shared memory by TMA (see
// attn_decode_kernel).  Partial (max, sum, acc[d]) per tile go to a workspace; the last CTA of a head to arrive (atomic
// ticket) merges them with the usual exp(m_i - M) rescaling, so there is no second launch and no atomics on data
// (run-to-run bit identical).  CTAs whose chunk lies past the current length exit at once.  The fused step
// (lk_compute_attention_step) also rotates q / k and appends this token's k / v rows in the same launch.
#include <cuda.h>   // CUtensorMap types only; the encoder is fetched through cudaGetDriverEntryPoint (no libcuda link)

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
constexpr int ATT_TILE = 32;            // cache positions per TMA box (two boxes of K and of V are staged per CTA)

struct AttnParams {
    const float *q;        // [D, H] (RoPE applied)
    const float *kc, *vc;  // [D, H, max_seq]
    const int32_t *pos;    // device scalar: index of the newest cache row; length = pos + 1
    float *dst;            // [D, H]
    float *ws;             // per (head, CTA): m, l, pad[2], acc[D]
    unsigned *tickets;     // per head, zero between launches
    int H, max_seq;
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

// One tile = ATT_TILE consecutive cache positions of one head.  The caches are 2-D tensors {H * D, max_seq} to the TMA
// unit: ONE cp.async.bulk.tensor.2d (UTMALDG) per tile brings the {D, ATT_TILE} box of K, another the box of V, into
// shared memory (dense [ATT_TILE][D] rows; positions past max_seq are zero-filled).  Two stages: a CTA has its next
// tile in flight while it computes on the current one (64 KB in flight per CTA from its first instructions).  (128
// separate 512-byte 1-D bulk copies per 64-row tile were measured first: 30 % slower than plain loads at 8 k positions
// — the copy engine's per-request cost; single-buffered boxes were still 17 % slower there.)
// Tiles are dealt round-robin at run time: with `ntiles` active tiles (from the device-side length) the first
// nact = min(grid.y, ntiles) CTAs of a head take tiles j, j + nact, ...: a short context gets one CTA per tile, a long
// one a few tiles per CTA with the loads pipelined.  Scores: a warp per K row (lane = float4 of the row: conflict-free),
// shuffle reduction; softmax of the tile by warp 0; V accumulation: thread d = column d (conflict-free).  A CTA folds its
// tiles into one running (M, L, acc[D]) (online softmax) and writes one partial; the last CTA of a head to arrive
// merges the nact partials.
__device__ __forceinline__ void tma_load_box(void *smem_dst, const CUtensorMap *tm, int c0, int c1, uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
        ::"r"(smem_u32(smem_dst)), "l"(tm), "r"(c0), "r"(c1), "r"(smem_u32(bar))
        : "memory");
}

template <int D>
__global__ void __launch_bounds__(ATT_THREADS) attn_decode_kernel(const AttnParams p, const __grid_constant__ CUtensorMap tm_k,
                                                                  const __grid_constant__ CUtensorMap tm_v) {
    constexpr int LPR = D / 4;                 // lanes per K row (one float4 each): 32 or 16
    constexpr int RPS = 32 / LPR;              // rows per warp step
    constexpr int STAGE = 2 * ATT_TILE * D;    // floats per stage: K box, then V box
    extern __shared__ __align__(128) float smem[];   // [2 stages][K box | V box]
    __shared__ float sc[ATT_TILE];             // scores -> probabilities of the current tile
    __shared__ float red[2];
    __shared__ int last_s;
    __shared__ __align__(16) float qs[D];      // this head's (rotated) q
    __shared__ __align__(8) uint64_t bar[2];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
        fence_mbar_init();
    }
    pdl_wait();
    pdl_launch_dependents();
    const int h = blockIdx.x, j = blockIdx.y;
    const int pos_raw = p.pos[0];
    int len = pos_raw + 1;
    len = len < 1 ? 1 : (len > p.max_seq ? p.max_seq : len);
    const int ntiles = (len + ATT_TILE - 1) / ATT_TILE;      // active tiles of this head
    const int nact = min((int)gridDim.y, ntiles);            // active CTAs of this head
    if (j >= nact) return;
    const int cnt = (ntiles - j + nact - 1) / nact;          // my tiles: j, j + nact, ...
    const size_t rowf = (size_t)p.H * D;       // floats between consecutive positions
    const bool fused = p.q_raw != nullptr;
    // the row this launch appends (fused step): written by the CTA whose tile holds it, into shared memory as well
    const int own_pos = (fused && pos_raw >= 0 && pos_raw < p.max_seq) ? pos_raw : -1;
    constexpr uint32_t TILE_BYTES = (uint32_t)STAGE * 4u;

    __syncthreads();                           // barriers initialised
    if (tid == 0) {
#pragma unroll
        for (int i = 0; i < 2; i++)
            if (i < cnt) {
                mbar_arrive_expect_tx(&bar[i], TILE_BYTES);
                tma_load_box(smem + i * STAGE, &tm_k, h * D, (j + i * nact) * ATT_TILE, &bar[i]);
                tma_load_box(smem + i * STAGE + ATT_TILE * D, &tm_v, h * D, (j + i * nact) * ATT_TILE, &bar[i]);
            }
    }

    // q of this head -> shared memory (rotated here in the fused step: rope_kernel's arithmetic)
    constexpr int HALF = D / 2;
    const int pt = pos_raw < 0 ? 0 : (pos_raw >= p.table_pos ? p.table_pos - 1 : pos_raw);   // as rope_kernel clamps
    if (fused) {
        if (tid < HALF) {
            const float2 r = rope_pair(((const float2 *)(p.q_raw + (size_t)h * D))[tid], p.table[(size_t)pt * HALF + tid]);
            ((float2 *)qs)[tid] = r;
            if (j == 0 && p.q_rope) ((float2 *)(p.q_rope + (size_t)h * D))[tid] = r;
        }
    } else if (tid < D / 4) {
        ((float4 *)qs)[tid] = ((const float4 *)(p.q + (size_t)h * D))[tid];
    }
    __syncthreads();
    const float4 qv = ((const float4 *)qs)[lane % LPR];

    float Mr = -INFINITY, Lr = 0.f, accr = 0.f;               // running max, sum, column d of the output
    for (int i = 0; i < cnt; i++) {
        const int tl = j + i * nact, st = i & 1;
        float *Ks = smem + st * STAGE, *Vs = Ks + ATT_TILE * D;
        const int s0 = tl * ATT_TILE, n = min(ATT_TILE, len - s0);
        const int own = (own_pos >= s0 && own_pos < s0 + n) ? own_pos - s0 : -1;
        float2 kr = make_float2(0.f, 0.f);
        float4 vv = make_float4(0.f, 0.f, 0.f, 0.f);
        if (own >= 0) {                        // KVCache.updateKey :47-55 / updateValue :88-96 for this head
            if (tid >= HALF && tid < D) {
                const int e = tid - HALF;
                kr = rope_pair(((const float2 *)(p.k_raw + (size_t)h * D))[e], p.table[(size_t)pt * HALF + e]);
                ((float2 *)(p.kc_w + (size_t)own_pos * rowf + (size_t)h * D))[e] = kr;
                if (p.k_rope) ((float2 *)(p.k_rope + (size_t)h * D))[e] = kr;
            } else if (tid < D / 4) {
                vv = ((const float4 *)(p.v_raw + (size_t)h * D))[tid];
                ((float4 *)(p.vc_w + (size_t)own_pos * rowf + (size_t)h * D))[tid] = vv;
            }
        }
        mbar_wait(&bar[st], (uint32_t)((i >> 1) & 1));
        if (own >= 0) {                        // the box held the row's old bytes: replace them (after the copy landed)
            if (tid >= HALF && tid < D) ((float2 *)(Ks + own * D))[tid - HALF] = kr;
            else if (tid < D / 4) ((float4 *)(Vs + own * D))[tid] = vv;
            __syncthreads();
        }
        // pass 1: scores, a warp per row (RPS rows per step for D = 64)
        {
            constexpr int ROWS_PER_WARP = ATT_TILE / (ATT_THREADS / 32);
#pragma unroll
            for (int s2 = 0; s2 < ROWS_PER_WARP / RPS; s2++) {
                const int t = warp * ROWS_PER_WARP + s2 * RPS + lane / LPR;
                const float4 kv = ((const float4 *)(Ks + t * D))[lane % LPR];
                float acc = fmaf(qv.x, kv.x, fmaf(qv.y, kv.y, fmaf(qv.z, kv.z, __fmul_rn(qv.w, kv.w))));
#pragma unroll
                for (int o = LPR / 2; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
                if (lane % LPR == 0 && t < n) sc[t] = __fmul_rn(acc, p.scale);
            }
        }
        __syncthreads();
        // max, exp, sum of the tile (LlamaAttention.softmax :262-276), by warp 0 (ATT_TILE == 32: one score per lane)
        if (warp == 0) {
            const float a0 = lane < n ? sc[lane] : -INFINITY;
            float m = a0;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
            const float e0 = lane < n ? expf(__fsub_rn(a0, m)) : 0.f;
            sc[lane] = e0;
            const float l = warp_sum(e0);
            if (lane == 0) { red[0] = m; red[1] = l; }
        }
        __syncthreads();
        const float m = red[0], l = red[1];
        // pass 2: thread d accumulates column d of V (rows past n carry probability 0)
        float acc = 0.f;
        if (tid < D) {
            const float *vcol = Vs + tid;
            if (n == ATT_TILE) {
#pragma unroll
                for (int t = 0; t < ATT_TILE; t++) acc = fmaf(sc[t], vcol[t * D], acc);
            } else {
                for (int t = 0; t < n; t++) acc = fmaf(sc[t], vcol[t * D], acc);
            }
        }
        // fold the tile into the running result (online softmax)
        const float Mn = fmaxf(Mr, m);
        const float fo = expf(__fsub_rn(Mr, Mn)), fn = expf(__fsub_rn(m, Mn));     // exp(-inf) = 0 for the first tile
        accr = fmaf(accr, fo, __fmul_rn(acc, fn));
        Lr = fmaf(Lr, fo, __fmul_rn(l, fn));
        Mr = Mn;
        __syncthreads();                       // everyone is done with this stage, sc and red
        if (tid == 0 && i + 2 < cnt) {
            mbar_arrive_expect_tx(&bar[st], TILE_BYTES);
            tma_load_box(Ks, &tm_k, h * D, (tl + 2 * nact) * ATT_TILE, &bar[st]);
            tma_load_box(Vs, &tm_v, h * D, (tl + 2 * nact) * ATT_TILE, &bar[st]);
        }
    }
    if (nact == 1) {                           // short context: this CTA has the whole row
        if (tid < D) p.dst[(size_t)h * D + tid] = __fdiv_rn(accr, Lr);
        return;
    }
    float *w = p.ws + ((size_t)h * gridDim.y + j) * (D + 4);
    if (tid == 0) { w[0] = Mr; w[1] = Lr; }
    if (tid < D) w[4 + tid] = accr;
    __threadfence();
    __syncthreads();
    if (tid == 0) last_s = (atomicAdd(&p.tickets[h], 1u) == (unsigned)(nact - 1));
    __syncthreads();
    if (!last_s) return;
    __threadfence();
    // merge the partials of this head: M = max m_i, out = sum acc_i e^(m_i - M) / sum l_i e^(m_i - M)
    const float *wh = p.ws + (size_t)h * gridDim.y * (D + 4);
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

// the cache as a 2-D F32 tensor {H * D, max_seq} (row pitch H * D * 4 bytes); one box = {D, ATT_TILE}
typedef CUresult (*AttnEncodeFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                 const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static cudaError_t make_cache_map(CUtensorMap *tm, const float *cache, int head_dim, int nheads, int64_t max_seq) {
    static AttnEncodeFn enc = [] {
        void *ptr = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
            ptr = nullptr;
        return (AttnEncodeFn)ptr;
    }();
    if (!enc) return cudaErrorNotSupported;
    const cuuint64_t gdim[2] = {(cuuint64_t)head_dim * nheads, (cuuint64_t)max_seq};
    const cuuint64_t gstride[1] = {(cuuint64_t)head_dim * nheads * 4};
    const cuuint32_t box[2] = {(cuuint32_t)head_dim, (cuuint32_t)ATT_TILE};
    const cuuint32_t estride[2] = {1, 1};
    const CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float *>(cache), gdim, gstride, box, estride,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? cudaSuccess : cudaErrorInvalidValue;
}

// CTAs per head (grid.y), fixed at launch so that a captured graph is valid for every length up to max_seq: one per
// tile, capped at ONE wave of resident CTAs (3 per SM at 64 KB of shared memory each) over all heads: a second, partly
// filled wave doubled the time at 512 positions (512 CTAs on 444 slots), and longer contexts want several pipelined
// tiles per CTA anyway
static int attn_splits(int nheads, int64_t max_seq, int sm_count) {
    const int64_t ntiles = (max_seq + ATT_TILE - 1) / ATT_TILE;
    int64_t cap = ((int64_t)sm_count * 3) / nheads;
    if (cap < 1) cap = 1;
    if (cap > ATT_MAX_SPLITS) cap = ATT_MAX_SPLITS;
    return (int)(ntiles < cap ? ntiles : cap);
}

static size_t attn_smem_bytes(int head_dim) { return (size_t)2 * 2 * ATT_TILE * head_dim * 4; }

size_t attn_workspace_bytes(int head_dim, int nheads, int64_t max_seq) {
    (void)max_seq;
    return (((size_t)nheads * 4 + 255) & ~(size_t)255) + (size_t)nheads * ATT_MAX_SPLITS * (head_dim + 4) * 4;
}

cudaError_t launch_attn_decode(const float *q, const float *kc, const float *vc, const int32_t *pos, int head_dim,
                               int nheads, int64_t max_seq, float scale, float *dst, void *workspace, int sm_count,
                               cudaStream_t s, bool pdl, const AttnStep *step) {
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
    p.tickets = (unsigned *)workspace;                              // [H], zeroed when the workspace is created
    p.ws = (float *)((uint8_t *)workspace + (((size_t)nheads * 4 + 255) & ~(size_t)255));
    const dim3 grid((unsigned)nheads, (unsigned)attn_splits(nheads, max_seq, sm_count));
    const size_t smem = attn_smem_bytes(head_dim);
    {   // > 48 KB of dynamic shared memory needs the opt-in; the attribute belongs to the (function, device) pair, so it is
        // set on every launch (a cheap call) instead of once per process: a second lk_ctx may sit on another GPU
        cudaError_t e = head_dim == 128
            ? cudaFuncSetAttribute(attn_decode_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)attn_smem_bytes(128))
            : cudaFuncSetAttribute(attn_decode_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)attn_smem_bytes(64));
        if (e != cudaSuccess) return e;
    }
    CUtensorMap tmk, tmv;
    cudaError_t me = make_cache_map(&tmk, kc, head_dim, nheads, max_seq);
    if (me == cudaSuccess) me = make_cache_map(&tmv, vc, head_dim, nheads, max_seq);
    if (me != cudaSuccess) return me;
    if (head_dim == 128) return launch_kernel(attn_decode_kernel<128>, grid, dim3(ATT_THREADS), smem, s, pdl, p, tmk, tmv);
    return launch_kernel(attn_decode_kernel<64>, grid, dim3(ATT_THREADS), smem, s, pdl, p, tmk, tmv);
}

// ---- the variant that allocates nothing (lk_compute_attention_decode_noalloc) ---------------------------------------
// Added for callers that cannot call cudaMalloc (code under CUDA-graph capture,
// tight serving loops, harnesses that forbid it); no counterpart in the reference.  The scratch memory of the decode
// attention (per-head tickets + the split partials the last CTA of a head merges) lives in this file-scope device array
// instead of the context's cudaMalloc'ed workspace: ATTN_STATIC_SCRATCH_BYTES on every GPU that loads the library, held
// until the process exits.  Host code may not pass the array's NAME to a kernel (in host code it is a host-side
// stand-in, not a device address): the device address is fetched once per device, with that device current, by
// cudaGetSymbolAddress below.  Every call on a device shares the array and finds the last call's contents in it: the
// tickets, which must start at zero, are cleared by a stream-ordered memset in front of every launch; the partials are
// written before they are read.  Two calls that overlap on the GPU (two streams, two contexts of one device) race.
constexpr size_t ATTN_STATIC_SCRATCH_BYTES = 512 * 1024;
static __device__ __align__(256) unsigned char g_attn_static_scratch[ATTN_STATIC_SCRATCH_BYTES];
constexpr int ATTN_STATIC_MAX_DEVICES = 64;
static void *g_attn_static_addr[ATTN_STATIC_MAX_DEVICES];     // device address of the array, per device (0 = not fetched yet)

size_t attn_static_scratch_capacity() { return ATTN_STATIC_SCRATCH_BYTES; }
// what one launch really touches: the tickets (rounded to 256 bytes) + nheads * splits partials of (head_dim + 4) floats
size_t attn_scratch_needed(int head_dim, int nheads, int64_t max_seq, int sm_count) {
    if (nheads <= 0 || max_seq <= 0) return 0;
    return (((size_t)nheads * 4 + 255) & ~(size_t)255) + (size_t)nheads * attn_splits(nheads, max_seq, sm_count) * (head_dim + 4) * 4;
}

cudaError_t launch_attn_decode_static(int device, const float *q, const float *kc, const float *vc, const int32_t *pos, int head_dim,
                                      int nheads, int64_t max_seq, float scale, float *dst, int sm_count, cudaStream_t s) {
    if (device < 0 || device >= ATTN_STATIC_MAX_DEVICES) return cudaErrorInvalidDevice;
    if (attn_scratch_needed(head_dim, nheads, max_seq, sm_count) > ATTN_STATIC_SCRATCH_BYTES) return cudaErrorInvalidValue;
    int cur = -1;
    cudaError_t e = cudaGetDevice(&cur);
    if (e != cudaSuccess) return e;
    if (cur != device) return cudaErrorInvalidDevice;          // the symbol's address is per device: that device must be current
    if (!g_attn_static_addr[device]) {
        void *ptr = nullptr;
        e = cudaGetSymbolAddress(&ptr, g_attn_static_scratch);
        if (e != cudaSuccess) return e;
        if (!ptr || ((uintptr_t)ptr & 255)) return cudaErrorInvalidSymbol;
        g_attn_static_addr[device] = ptr;
    }
    void *scratch = g_attn_static_addr[device];
    e = cudaMemsetAsync(scratch, 0, ((size_t)nheads * 4 + 255) & ~(size_t)255, s);      // tickets start at zero whatever the last call left
    if (e != cudaSuccess) return e;
    return launch_attn_decode(q, kc, vc, pos, head_dim, nheads, max_seq, scale, dst, scratch, sm_count, s, false, nullptr);
}

}  // namespace lk

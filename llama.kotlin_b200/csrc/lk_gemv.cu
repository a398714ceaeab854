// lk_gemv.cu — K1: decode GEMV  dst[i] = sum_k deq(a)[i,k] * b[k]   (N == 1)
//
// Replaces the M x 1 case of computeMatMul's Q4_0 / Q4_1 / Q8_0 x F32 branches
// (core/GGMLComputeOps.kt:1448-1527 -> computeDotProductQ40F32/Q41F32/Q80F32,
// :120-145, :70-115, :43-68) on the reference's own block layouts (34/18/20 B,
// Kotlin nibble order core/GGMLTypes.kt:647-651).
//
// HBM-bound: every weight byte is read once, so the design is about keeping
// ~100 KB per SM in flight and spending < 3 issue slots per weight element.
//
//  * persistent grid, one CTA per SM; CTA c owns a contiguous, balanced row
//    range (contiguous rows == one contiguous byte range of the packed matrix);
//  * one producer warp streams that byte range through an NSTAGES-deep shared
//    memory ring with 1-D TMA bulk copies (cp.async.bulk -> UBLKCP) completing
//    on mbarriers; weights use an L2 evict_first policy (read once);
//  * consumer thread (rg, c) owns block-pair column c (64 consecutive k) for
//    every row of row group rg: its 64 activations live in registers for the
//    whole kernel, so shared memory only carries weight bytes (9 / 10 / 17
//    conflict-free word loads per 64 elements);
//  * nibbles are turned into F32 with one PRMT each: byte-isolate
//    (w & 0x0F0F0F0F, (w>>4) & 0x0F0F0F0F), then PRMT the byte into mantissa
//    bits 16..19 of 0x43000000 => exactly 128 + nib; the bias is removed once
//    per block algebraically:  sum (nib-8) x = sum (128+nib) x - 136 * sum x.
//    The multiply-accumulates are packed FFMA2 (fma.rn.f32x2), 2 MACs/issue;
//  * Q8_0 bytes: XOR 0x80, PRMT into the low mantissa byte of 0x4B000000
//    (2^23 + q + 128), exact FADD2 de-bias, FFMA2;
//  * per-row partials: transposed warp-shuffle reduction, then a cross-warp
//    sum through shared memory (double-buffered, one named barrier per stage);
//  * programmatic dependent launch: weight prefetch is issued before
//    griddepcontrol.wait (weights are immutable inside a captured op
//    sequence), the activation vector is read after it, so back-to-back GEMVs
//    overlap the tail of one kernel with the HBM ramp of the next.
//
// Numerics: F32 throughout; summation order differs from the reference's
// sequential k loop and the +128 bias costs ~5 bits of the 24-bit accumulator
// (bound ~3e-6 relative), far inside the 1e-3 matmul tolerance.
#include "lk_common.cuh"
#include "lk_kernels.h"

namespace lk {

constexpr int GEMV_MAX_STAGES = 8;
constexpr int GEMV_RED_STRIDE = 16;   // max column warps
constexpr int GEMV_RED_ROWS = 64;     // max rows per stage

template <int TYPE> struct PairWords;
template <> struct PairWords<T_Q4_0> { static constexpr int value = 9; };
template <> struct PairWords<T_Q4_1> { static constexpr int value = 10; };
template <> struct PairWords<T_Q8_0> { static constexpr int value = 17; };

// 128 + nibble held in byte j of an isolated word (exact)
__device__ __forceinline__ float nibf(uint32_t iso, int j) {
    return __uint_as_float(__byte_perm(iso, 0x43000000u, 0x7044u | (uint32_t)(j << 8)));
}
// 2^23 + (unsigned byte j)
__device__ __forceinline__ float bytef(uint32_t w, int j) {
    return __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7440u | (uint32_t)j));
}

// accumulate one word of 4-bit quants: bytes [B0, B1) of w hold element pairs P0, P0+1, ...
template <int B0, int B1, int P0>
__device__ __forceinline__ void q4_word(uint32_t w, const float2 (&x)[32], float2 &acc_a, float2 &acc_b) {
    const uint32_t lo = w & 0x0F0F0F0Fu, hi = (w >> 4) & 0x0F0F0F0Fu;
#pragma unroll
    for (int j = B0; j < B1; j++) {
        const float2 f = make_float2(nibf(lo, j), nibf(hi, j));
        if ((j & 1) == 0) acc_a = ffma2(f, x[P0 + j - B0], acc_a);
        else              acc_b = ffma2(f, x[P0 + j - B0], acc_b);
    }
}

// accumulate bytes [B0, B1) of an int8 word; element E0 is byte B0 (E0 even, B0 even)
template <int B0, int B1, int P0>
__device__ __forceinline__ void q8_word(uint32_t w, const float2 (&x)[32], float2 &acc_a, float2 &acc_b) {
    const uint32_t t = w ^ 0x80808080u;
    const float2 bias = make_float2(-8388736.0f, -8388736.0f);  // -(2^23 + 128), exact
#pragma unroll
    for (int j = B0; j < B1; j += 2) {
        const float2 f = fadd2(make_float2(bytef(t, j), bytef(t, j + 1)), bias);
        if ((j & 2) == 0) acc_a = ffma2(f, x[P0 + (j - B0) / 2], acc_a);
        else              acc_b = ffma2(f, x[P0 + (j - B0) / 2], acc_b);
    }
}

// dot of one block pair (64 consecutive k of one row) with the thread's 64 activations.
// wp: 4-byte aligned shared-memory address of the pair.  sx0/sx1: sum of x over each block.
template <int TYPE>
__device__ __forceinline__ float pair_dot(const uint32_t *wp, const float2 (&x)[32], float sx0, float sx1) {
    const float2 z = make_float2(0.f, 0.f);
    float2 a0 = z, b0 = z, a1 = z, b1 = z;
    if (TYPE == T_Q4_0) {
        uint32_t w[9];
#pragma unroll
        for (int i = 0; i < 9; i++) w[i] = wp[i];
        q4_word<2, 4, 0>(w[0], x, a0, b0);
        q4_word<0, 4, 2>(w[1], x, a0, b0);
        q4_word<0, 4, 6>(w[2], x, a0, b0);
        q4_word<0, 4, 10>(w[3], x, a0, b0);
        q4_word<0, 2, 14>(w[4], x, a0, b0);
        q4_word<0, 4, 16>(w[5], x, a1, b1);
        q4_word<0, 4, 20>(w[6], x, a1, b1);
        q4_word<0, 4, 24>(w[7], x, a1, b1);
        q4_word<0, 4, 28>(w[8], x, a1, b1);
        const float d0 = h2f((uint16_t)(w[0] & 0xFFFFu)), d1 = h2f((uint16_t)(w[4] >> 16));
        const float s0 = (a0.x + a0.y) + (b0.x + b0.y), s1 = (a1.x + a1.y) + (b1.x + b1.y);
        return d0 * fmaf(-136.0f, sx0, s0) + d1 * fmaf(-136.0f, sx1, s1);
    } else if (TYPE == T_Q4_1) {
        uint32_t w[10];
        const uint2 *wp2 = (const uint2 *)wp;  // pair is 8-byte aligned; LDS.64 is conflict-free at stride 40 B
#pragma unroll
        for (int i = 0; i < 5; i++) { const uint2 v = wp2[i]; w[2 * i] = v.x; w[2 * i + 1] = v.y; }
        q4_word<0, 4, 0>(w[1], x, a0, b0);
        q4_word<0, 4, 4>(w[2], x, a0, b0);
        q4_word<0, 4, 8>(w[3], x, a0, b0);
        q4_word<0, 4, 12>(w[4], x, a0, b0);
        q4_word<0, 4, 16>(w[6], x, a1, b1);
        q4_word<0, 4, 20>(w[7], x, a1, b1);
        q4_word<0, 4, 24>(w[8], x, a1, b1);
        q4_word<0, 4, 28>(w[9], x, a1, b1);
        const float d0 = h2f((uint16_t)(w[0] & 0xFFFFu)), m0 = h2f((uint16_t)(w[0] >> 16));
        const float d1 = h2f((uint16_t)(w[5] & 0xFFFFu)), m1 = h2f((uint16_t)(w[5] >> 16));
        const float s0 = (a0.x + a0.y) + (b0.x + b0.y), s1 = (a1.x + a1.y) + (b1.x + b1.y);
        return (d0 * fmaf(-128.0f, sx0, s0) + m0 * sx0) + (d1 * fmaf(-128.0f, sx1, s1) + m1 * sx1);
    } else {
        uint32_t w[17];
#pragma unroll
        for (int i = 0; i < 17; i++) w[i] = wp[i];
        q8_word<2, 4, 0>(w[0], x, a0, b0);
#pragma unroll
        for (int i = 1; i < 8; i++) {
            // word i holds elements 2 + 4(i-1) .. => pairs 1 + 2(i-1), +1
            const uint32_t t = w[i] ^ 0x80808080u;
            const float2 bias = make_float2(-8388736.0f, -8388736.0f);
            a0 = ffma2(fadd2(make_float2(bytef(t, 0), bytef(t, 1)), bias), x[1 + 2 * (i - 1)], a0);
            b0 = ffma2(fadd2(make_float2(bytef(t, 2), bytef(t, 3)), bias), x[2 + 2 * (i - 1)], b0);
        }
        q8_word<0, 2, 15>(w[8], x, a0, b0);
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const uint32_t t = w[9 + i] ^ 0x80808080u;
            const float2 bias = make_float2(-8388736.0f, -8388736.0f);
            a1 = ffma2(fadd2(make_float2(bytef(t, 0), bytef(t, 1)), bias), x[16 + 2 * i], a1);
            b1 = ffma2(fadd2(make_float2(bytef(t, 2), bytef(t, 3)), bias), x[17 + 2 * i], b1);
        }
        const float d0 = h2f((uint16_t)(w[0] & 0xFFFFu)), d1 = h2f((uint16_t)(w[8] >> 16));
        const float s0 = (a0.x + a0.y) + (b0.x + b0.y), s1 = (a1.x + a1.y) + (b1.x + b1.y);
        return d0 * s0 + d1 * s1;
    }
}

struct GemvParams {
    const uint8_t *a;     // packed rows, 16-byte aligned
    const float *b;       // K activations, 16-byte aligned
    float *dst;           // M_total outputs
    float *peer[8];       // optional peer-mapped copies of dst (fused all-gather)
    int npeers;
    int K, M;             // this launch's rows
    int64_t row_offset;   // dst index of row 0 (sharded weights)
    int row_bytes;        // K/32 * blockBytes
    int npairs;           // K/64
    int ncw;              // column warps = ceil(npairs/32)
    int rg;               // row groups
    int cwarps;           // consumer warps = ncw * rg (producer is warp `cwarps`)
    int stage_rows;       // rg * ACC
    int nstages;
    int unit_rows;        // CTA row ranges are multiples of this (keeps TMA sizes 16-byte multiples)
    int x_off, stage_off; // dynamic smem offsets (bytes)
};

// transposed shuffle reduction of ACC per-lane values over the 32 lanes;
// afterwards lane l holds the full sum of value index l / (32/ACC).
template <int ACC>
__device__ __forceinline__ float warp_reduce_multi(float (&acc)[ACC], int lane) {
    int o = 16;
#pragma unroll
    for (int n = ACC; n > 1; n >>= 1) {
        const int h = n >> 1;
        const bool up = (lane & o) != 0;
#pragma unroll
        for (int i = 0; i < h; i++) {
            const float keep = up ? acc[i + h] : acc[i];
            const float send = up ? acc[i] : acc[i + h];
            acc[i] = keep + __shfl_xor_sync(0xffffffffu, send, o);
        }
        o >>= 1;
    }
    float v = acc[0];
    for (; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

template <int TYPE, int ACC, int MAXW>
__global__ void __launch_bounds__((MAXW + 1) * 32, 1) gemv_kernel(const GemvParams p) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t *full = (uint64_t *)smem;                   // [GEMV_MAX_STAGES]
    uint64_t *empty = full + GEMV_MAX_STAGES;            // [GEMV_MAX_STAGES]
    uint64_t *xbar = empty + GEMV_MAX_STAGES;            // [1]
    float *red = (float *)(smem + 256);                  // [2][GEMV_RED_ROWS][GEMV_RED_STRIDE]
    float *xs = (float *)(smem + p.x_off);               // [K]
    uint8_t *stage0 = smem + p.stage_off;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int stage_bytes = p.stage_rows * p.row_bytes;

    // balanced contiguous row range of this CTA, in units of unit_rows
    const int units = p.M / p.unit_rows;
    const int r_begin = (int)((int64_t)units * blockIdx.x / gridDim.x) * p.unit_rows;
    const int r_end = (int)((int64_t)units * (blockIdx.x + 1) / gridDim.x) * p.unit_rows;
    const int nrows = r_end - r_begin;
    const int niter = (nrows + p.stage_rows - 1) / p.stage_rows;

    if (tid == 0) {
        for (int s = 0; s < p.nstages; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], p.cwarps);
        }
        mbar_init(xbar, 1);
        fence_mbar_init();
    }
    __syncthreads();
    pdl_launch_dependents();  // the next GEMV may start its own weight prefetch now

    if (warp == p.cwarps) {
        // ===== producer warp: one lane streams this CTA's byte range =====
        if (lane == 0) {
            const uint64_t pol_w = l2_policy_evict_first();
            const uint64_t pol_x = l2_policy_evict_last();
            const uint8_t *src = p.a + (int64_t)r_begin * p.row_bytes;
            int it = 0;
            const int pre = niter < p.nstages ? niter : p.nstages;
            for (; it < pre; it++) {  // weight prefetch: independent of the previous kernel
                const int rows = min(p.stage_rows, nrows - it * p.stage_rows);
                const uint32_t bytes = (uint32_t)(rows * p.row_bytes);
                mbar_arrive_expect_tx(&full[it], bytes);
                tma_load_1d(stage0 + (size_t)it * stage_bytes, src + (size_t)it * stage_bytes, bytes, &full[it], pol_w);
            }
            pdl_wait();  // activations may be the previous kernel's output
            mbar_arrive_expect_tx(xbar, (uint32_t)(p.K * 4));
            tma_load_1d(xs, p.b, (uint32_t)(p.K * 4), xbar, pol_x);
            for (; it < niter; it++) {
                const int s = it % p.nstages;
                mbar_wait(&empty[s], ((it / p.nstages) - 1) & 1);
                const int rows = min(p.stage_rows, nrows - it * p.stage_rows);
                const uint32_t bytes = (uint32_t)(rows * p.row_bytes);
                mbar_arrive_expect_tx(&full[s], bytes);
                tma_load_1d(stage0 + (size_t)s * stage_bytes, src + (size_t)it * stage_bytes, bytes, &full[s], pol_w);
            }
        }
        return;
    }

    // ===== consumer warps =====
    pdl_wait();  // orders our dst stores after the previous kernel as well
    const int cw = warp % p.ncw;         // column warp
    const int rgi = warp / p.ncw;        // row group
    const int col = cw * 32 + lane;      // block-pair column
    const bool active = col < p.npairs;
    constexpr int PW = PairWords<TYPE>::value;

    float2 x[32];
    float sx0 = 0.f, sx1 = 0.f;
    mbar_wait(xbar, 0);
    if (active) {
        const float4 *xp = (const float4 *)(xs + (size_t)col * 64);
#pragma unroll
        for (int i = 0; i < 16; i++) {
            const float4 v = xp[i];
            x[2 * i] = make_float2(v.x, v.y);
            x[2 * i + 1] = make_float2(v.z, v.w);
        }
#pragma unroll
        for (int i = 0; i < 16; i++) sx0 += x[i].x + x[i].y;
#pragma unroll
        for (int i = 16; i < 32; i++) sx1 += x[i].x + x[i].y;
    } else {
#pragma unroll
        for (int i = 0; i < 32; i++) x[i] = make_float2(0.f, 0.f);
    }

    const int nct = p.cwarps * 32;
    for (int it = 0; it < niter; it++) {
        const int s = it % p.nstages;
        const int rows_here = min(p.stage_rows, nrows - it * p.stage_rows);
        mbar_wait(&full[s], (it / p.nstages) & 1);
        const uint8_t *st = stage0 + (size_t)s * stage_bytes;
        float acc[ACC];
#pragma unroll
        for (int a = 0; a < ACC; a++) {
            const int row = a * p.rg + rgi;
            acc[a] = 0.f;
            if (active && row < rows_here)
                acc[a] = pair_dot<TYPE>((const uint32_t *)(st + (size_t)row * p.row_bytes) + col * PW, x, sx0, sx1);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);  // this warp is done reading the stage
        float *rb = red + (it & 1) * (GEMV_RED_ROWS * GEMV_RED_STRIDE);
        {
            const float v = warp_reduce_multi<ACC>(acc, lane);
            if ((lane & (32 / ACC - 1)) == 0) {
                const int a = lane / (32 / ACC);
                rb[(a * p.rg + rgi) * GEMV_RED_STRIDE + cw] = v;
            }
        }
        asm volatile("bar.sync 1, %0;" ::"r"(nct) : "memory");
        if (tid < rows_here) {
            float v = rb[tid * GEMV_RED_STRIDE];
            for (int c = 1; c < p.ncw; c++) v += rb[tid * GEMV_RED_STRIDE + c];
            const int64_t o = p.row_offset + r_begin + it * p.stage_rows + tid;
            p.dst[o] = v;
            for (int q = 0; q < p.npeers; q++) p.peer[q][o] = v;
        }
    }
}

// ---------------------------------------------------------------------------
// host side: shape analysis + launch
// ---------------------------------------------------------------------------
template <int TYPE, int ACC, int MAXW>
static cudaError_t launch_one(const GemvParams &p, int grid, int smem, int pdl, cudaStream_t s) {
    auto kern = gemv_kernel<TYPE, ACC, MAXW>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3((unsigned)((p.cwarps + 1) * 32));
    cfg.dynamicSmemBytes = (size_t)smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, p);
}

template <int TYPE, int MAXW>
static cudaError_t launch_acc(int acc, const GemvParams &p, int grid, int smem, int pdl, cudaStream_t s) {
    switch (acc) {
        case 1: return launch_one<TYPE, 1, MAXW>(p, grid, smem, pdl, s);
        case 2: return launch_one<TYPE, 2, MAXW>(p, grid, smem, pdl, s);
        case 4: return launch_one<TYPE, 4, MAXW>(p, grid, smem, pdl, s);
        case 8: return launch_one<TYPE, 8, MAXW>(p, grid, smem, pdl, s);
        default: return cudaErrorInvalidValue;
    }
}

template <int TYPE>
static cudaError_t launch_type(int acc, const GemvParams &p, int grid, int smem, int pdl, cudaStream_t s) {
    if (p.cwarps <= 8) return launch_acc<TYPE, 8>(acc, p, grid, smem, pdl, s);
    return launch_acc<TYPE, 16>(acc, p, grid, smem, pdl, s);
}

cudaError_t launch_gemv(int type, const void *a, int64_t K, int64_t M, const float *b, float *dst,
                        int64_t row_offset, const GemvPeers &peers, const GemvTuning &tune, int sm_count,
                        cudaStream_t s) {
    const int bb = block_bytes(type);
    if (bb == 0 || K <= 0 || M <= 0 || (K % 64) != 0 || K > 64 * 32 * 16 || M > (1 << 30))
        return cudaErrorNotSupported;
    if (((uintptr_t)a & 15) || ((uintptr_t)b & 15)) return cudaErrorNotSupported;

    GemvParams p = {};
    p.a = (const uint8_t *)a; p.b = b; p.dst = dst;
    p.npeers = peers.n;
    for (int i = 0; i < peers.n && i < 8; i++) p.peer[i] = peers.ptr[i];
    p.K = (int)K; p.M = (int)M; p.row_offset = row_offset;
    p.row_bytes = (int)(K / 32) * bb;
    p.npairs = (int)(K / 64);
    p.ncw = (p.npairs + 31) / 32;
    const int target_warps = tune.variant > 0 ? tune.variant : 8;   // consumer warps to aim for
    p.rg = target_warps / p.ncw;
    if (p.rg < 1) p.rg = 1;
    while (p.rg * p.ncw > 16) p.rg--;
    if (p.rg < 1) return cudaErrorNotSupported;
    p.cwarps = p.rg * p.ncw;
    p.unit_rows = 1;
    while ((p.unit_rows * p.row_bytes) % 16) p.unit_rows *= 2;   // <= 8 because row_bytes is even
    if (M % p.unit_rows) return cudaErrorNotSupported;

    const int budget = tune.smem_budget > 0 ? tune.smem_budget : 112 * 1024;
    const int stage_target = tune.stage_bytes > 0 ? tune.stage_bytes : 20 * 1024;
    const int fixed = 256 + 2 * GEMV_RED_ROWS * GEMV_RED_STRIDE * 4;
    p.x_off = fixed;
    p.stage_off = (p.x_off + (int)K * 4 + 127) & ~127;
    int acc = 8;
    while (acc > 1 && (int64_t)p.rg * acc * p.row_bytes > stage_target) acc >>= 1;
    while ((p.rg * acc) % p.unit_rows && acc < 8) acc <<= 1;
    if ((p.rg * acc) % p.unit_rows) return cudaErrorNotSupported;
    p.stage_rows = p.rg * acc;
    const int stage_bytes = p.stage_rows * p.row_bytes;
    int nst = (budget - p.stage_off) / stage_bytes;
    const int max_st = tune.max_stages > 0 ? tune.max_stages : GEMV_MAX_STAGES;
    if (nst > max_st) nst = max_st;
    if (nst > GEMV_MAX_STAGES) nst = GEMV_MAX_STAGES;
    if (nst < 2) {
        nst = 2;  // very long rows: take what two stages need, up to the 227 KB limit
        if (p.stage_off + 2 * stage_bytes > 227 * 1024) return cudaErrorNotSupported;
    }
    p.nstages = nst;
    const int smem = p.stage_off + nst * stage_bytes;

    int grid = sm_count * (tune.ctas_per_sm > 0 ? tune.ctas_per_sm : 1);
    const int units = (int)(M / p.unit_rows);
    if (grid > units) grid = units;

    switch (type) {
        case T_Q4_0: return launch_type<T_Q4_0>(acc, p, grid, smem, tune.pdl, s);
        case T_Q4_1: return launch_type<T_Q4_1>(acc, p, grid, smem, tune.pdl, s);
        case T_Q8_0: return launch_type<T_Q8_0>(acc, p, grid, smem, tune.pdl, s);
        default: return cudaErrorNotSupported;
    }
}

}  // namespace lk

// lk_gemv_mma.cu — K1m: a chain of dependent decode GEMVs as one persistent launch, plane sums on the warp-level int8 MMA
//
// Same contract as K1c (lk_gemv.cu: gemv_chain_kernel; links = lk_compute_matmul_fused calls, in order, of
// computeMatMul's Q4_0 / Q4_1 / Q8_0 x F32 branches, core/GGMLComputeOps.kt:1448-1527, with the layer graph's neighbours
// folded in), same block layouts (18 / 20 / 34 bytes, core/GGMLTypes.kt:543-732) read where they lie, same arithmetic:
// activations as 23-bit block fixed point in three byte planes, exact int32 plane sums, F32 per block.  What changes is
// who takes the plane sums.  The chain timeline (tools/chain_timeline.cu, profiles/r02i_chain_timeline.txt) showed K1c's
// consumers, not the weight stream, pacing every phase: 62 instructions per 32-element block per lane, 24 of them DP4As,
// on 8 working warps = 5.3 TB/s-equivalent, below the 6.5 TB/s HBM delivers, and every microsecond of a hand-off is lost.
// Here one IMMA.16832.U8.S8 (mma.sync.m16n8k32, measured 0.5 per clock per SM, tools/ubench_imma.cu; this kernel needs
// 0.09) replaces 12 warp-level DP4As:
//   * A (16 x 32, u8) = the nibbles of 16 weight rows: lane (g, t) of a warp loads two quant words (8 bytes) of rows g
//     and g + 8 and isolates low / high nibbles as bytes (AND, SHR + AND), exactly the unpacking the DP4A kernel does,
//     but for a 16-row strip at once;
//   * B (32 x 8, s8) = the activation planes: columns 0-2 the three planes of the even block of a block pair, 4-6 those
//     of the odd block (3 and 7 zero); the 32 k-slots of one MMA cover 16 elements of each block, a second MMA into the
//     same accumulators the other 16.  Planes are signed base-256 digits of T (bytes of T + 0x8080, the lower two
//     XOR 0x80), so one operand type serves all three;
//   * C (16 x 8, s32): lane (g, t') holds two adjacent columns of rows g and g + 8: (plane 0, plane 1) or (plane 2, 0)
//     of one block -> I = c0 + (c1 << 8) exactly, one I2F, then F32 with the block's d (read from the row), the
//     activation scale and the -8 s sumT term from a small per-block table: 13 instructions per row pair and block pair.
// About 42 instructions per 1024 weight elements per warp instead of 62, the FMA pipe nearly idle, and all 16 consumer
// warps work in every phase: a stage is one strip of <= 16 rows (rows longer than ~2.6 KB in equal column chunks), warp w takes K-slice w
// of it, parks its two row sums per lane quad in shared memory (no per-stage CTA barrier), one fixed-order reduction
// per phase.  Rows are staged with one 1-D TMA bulk copy each at a padded stride (stride / 4 = 4 mod 8: the eight rows
// a load instruction touches fall into different banks), in a byte ring of variable-size stages whose offsets the
// producer publishes per stage.  Range fix-up, RMSNorm prologue, residual / SwiGLU epilogues, stacked segments, grid
// barrier and time-out reporting as in K1c.  Results agree with K1c / the separate launches to F32 rounding (different
// summation order: 16 K-slices instead of block-column quads), not bit for bit.
#include <algorithm>
#include <vector>

#include "lk_common.cuh"
#include "lk_kernels.h"

namespace lk {

constexpr int MC_MAX = GEMV_CHAIN_MAX;  // phases per launch
constexpr int MC_W = 16;                // consumer warps
constexpr int MC_NCT = MC_W * 32;
constexpr int MC_PW = 4;                // producer warps (a CTA's registers are allocated in units of four warps: the three extra ones are free)
constexpr int MC_NB = 8;                // stages in flight (mbarrier pairs)
constexpr int MC_PS = 25;               // floats per row of the partial-sum buffer: 16 K-slices, 8 fix-up columns, 1 pad (odd stride)
constexpr int MC_HDR = 512;             // header bytes of dynamic shared memory
constexpr int MC_XB = 96;               // bytes of activation planes per block: 3 planes x 2 halves x 16
constexpr int MC_MASK_WORDS = 16;       // range flags: K <= 16384 (Llama-2-13B's down projection: 13824)
constexpr int MC_KMAX = GEMV_MCHAIN_KMAX;   // longest activation vector a mailbox buffer holds

// header layout (bytes): 0 full[8] | 64 empty[8] | 128 stage_off[8] | 160 any[2] | 168 epoch base | 176 zero[4] |
//                        192 mask[2][16] | 320 sum-of-squares partials[2][16]       ([2]: by phase parity, see mc_prep_warp)
constexpr int MC_OFF_STAGE = 128, MC_OFF_ANY = 160, MC_OFF_EPOCH = 168, MC_OFF_ZERO = 176, MC_OFF_MASK = 192, MC_OFF_RMS = 320;
static_assert(MC_OFF_MASK + 2 * MC_MASK_WORDS * 4 <= MC_OFF_RMS && MC_OFF_RMS + 2 * MC_W * 4 <= MC_HDR, "header layout");

struct MPhase {
    const uint8_t *a[3];       // weight segments (identical shape), 16-byte aligned
    float *dst[3];
    const float *b;            // K activations
    const float *pro_w;        // RMSNorm gain (may be NULL with pro_mode 1)
    float *b_out;              // CTA 0 also stores the activations actually used
    const float *residual;     // dst = residual + a x b
    float *dst_raw;            // optional: the bare product
    float *epi_act, *epi_out;  // epi_mode 1: silu(seg 0), silu(seg 0) * seg 1
    float pro_eps, inv_k;      // inv_k = 1 / K
    int pro_mode, epi_mode, nseg;
    int K, M, nblk, row_bytes;
    int stride;                // shared-memory row stride (bytes) of a staged row chunk: >= its bytes, = 16 mod 32
    int sr;                    // rows per stage (strip): 16
    int nchunk, c0;            // a strip is staged as nchunk column chunks of c0 block pairs each (the last one may be shorter)
    int rows_per_cta, rows_rem;   // M = rows_per_cta * grid + rows_rem
    int rows_max;              // rows of the largest CTA share (segment s starts at partial-sum row s * rows_max)
    int fix_gap, nib;
    int from_mbox;             // the activations are the previous phase's output: read them, value + epoch, from the mailbox
    int feed;                  // which output the next phase consumes: 0 none, 1 + s: dst[s] (after the residual add), 4: epi_out
};
struct MChainParams {
    MPhase ph[MC_MAX];
    int nphase;
    int wait_mode;
    int part_off, part_floats, x_off, sc_off, ring_off, ring_bytes;     // two partial-sum buffers of part_floats each (by phase parity)
    int inflight;              // stages the producer keeps in flight (issued, not landed); the ring holds more, landed ones
    unsigned *bar;             // words 0-1: the context's monotonic 64-bit launch / arrival counter (shared with K1c), word 2: time-outs
    uint2 *mbox;               // mailbox: 2 x MC_KMAX (value bits, epoch) pairs
    unsigned long long *tl;    // tools/chain_timeline.cu only
};
static_assert(sizeof(MChainParams) <= 4000, "kernel parameter space");
static_assert(MC_NB == 8, "slot arithmetic in the stage loop");
static_assert(sizeof(MPhase) <= 192, "the parameter warm-up touches four lines");

#ifdef LK_GEMV_TIMELINE
#define MC_TL_STRIDE (8 + 8 * 8)
#define MC_TL(ph, slot) do { if (cp.tl) cp.tl[(size_t)blockIdx.x * MC_TL_STRIDE + 8 + (ph) * 8 + (slot)] = (unsigned long long)clock64(); } while (0)
#else
#define MC_TL(ph, slot) do { } while (0)
#endif

__device__ __forceinline__ uint32_t mc_lds32(uint32_t a) { uint32_t v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ uint32_t mc_lds16(uint32_t a) { uint16_t v; asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(a)); return v; }
__device__ __forceinline__ float2 mc_lds64f(uint32_t a) { float2 v; asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(a)); return v; }
__device__ __forceinline__ uint4 mc_lds128(uint32_t a) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));
    return v;
}
// mbarrier wait / arrive on a 32-bit shared address (no generic -> shared conversion per stage)
__device__ __forceinline__ void mc_mbar_wait(uint32_t addr, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "MC_WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@!p bra MC_WAIT_%=;\n\t}"
        ::"r"(addr), "r"(parity) : "memory");
}
__device__ __forceinline__ void mc_mbar_arrive(uint32_t addr) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(addr) : "memory");
}
// D = A (16 x 32 bytes, row) x B (32 x 8 signed bytes, col) + C; SIGNED_A: Q8_0 quants, else nibbles / unsigned
template <bool SIGNED_A>
__device__ __forceinline__ void mc_imma(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    if (SIGNED_A)
        asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
    else
        asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// the same with C = 0 (the first MMA of a block pair)
template <bool SIGNED_A>
__device__ __forceinline__ void mc_imma0(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    if (SIGNED_A)
        asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%10,%10,%10};"
                     : "=r"(c[0]), "=r"(c[1]), "=r"(c[2]), "=r"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1), "r"(0));
    else
        asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%10,%10,%10};"
                     : "=r"(c[0]), "=r"(c[1]), "=r"(c[2]), "=r"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1), "r"(0));
}

__device__ __forceinline__ unsigned long long mc_ld_acquire_gpu(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void mc_red_release_gpu_add(unsigned long long *p, unsigned long long v) {
    asm volatile("red.release.gpu.global.add.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ float mc_fmax3_nan_abs(float a, float b, float c) {
    float d;
    asm("max.NaN.abs.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}
__device__ __forceinline__ float mc_fmin3_abs(float a, float b, float c) {
    float d;
    asm("min.abs.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}
__device__ __forceinline__ float mc_silu(float x) { return __fmul_rn(x, __fdiv_rn(1.0f, __fadd_rn(1.0f, expf(-x)))); }   // = lk_elementwise.cu silu1

// ---- hand-off between phases without a grid barrier -------------------------------------------------------------------------
// K1c separates phases with a grid-wide barrier: every CTA stores its outputs, fences, arrives on a counter and polls it, then
// loads the activations: three L2 round trips and a fence in a row, 2.2 us per hand-off on the timeline
// (profiles/r02i_mchain_timeline.txt), as long as the phase's own weight stream.  Here the data is its own flag: the epilogue
// stores every output that the next phase multiplies with a second time, into a mailbox, as an 8-byte (value, epoch) pair
// (one 8-byte store: value and flag cannot be seen apart), and the next phase's activation load polls those pairs until they
// carry its epoch.  No fence, no counter, no separate load: one store and one load on the critical path.  Epochs come from the
// context's monotonic counter (every CTA adds 1 when it is done, so the counter is a multiple of the grid size between
// launches; epoch = counter / grid * 8 + phase + 1 is unique per launch and phase, nothing is ever reset), the two mailbox
// buffers alternate by phase: a CTA overwrites buffer b for phase i + 2 only after it has read all of phase i + 1's
// activations, which every CTA stored after it was done reading phase i's.  What the launch writes apart from the mailbox
// (the tensors the graph names) is never read by another CTA of the same launch: the host side checks that and declines
// chains where it could be (launch_gemv_mchain).
__device__ __forceinline__ void mc_st_pair(uint2 *q, float v, uint32_t epoch) {
    asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" ::"l"(q), "r"(__float_as_uint(v)), "r"(epoch) : "memory");
}
__device__ __forceinline__ uint2 mc_ld_pair(const uint2 *q) {
    uint2 v;
    asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(q) : "memory");
    return v;
}
__device__ __forceinline__ uint4 mc_ld_pairs(const uint2 *q) {       // two pairs
    uint4 v;
    asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(q) : "memory");
    return v;
}
// eight consecutive activations of the mailbox, polled until every pair carries `epoch`; bounded (~1 s), a time-out is counted
__device__ __forceinline__ void mc_load8_mbox(const uint2 *q, uint32_t epoch, float (&v)[8], unsigned *timeouts) {
    uint4 a, b, c, d;
    const long long t0 = clock64();
    while (true) {
        a = mc_ld_pairs(q); b = mc_ld_pairs(q + 2); c = mc_ld_pairs(q + 4); d = mc_ld_pairs(q + 6);
        const bool ok = a.y == epoch && a.w == epoch && b.y == epoch && b.w == epoch && c.y == epoch && c.w == epoch && d.y == epoch && d.w == epoch;
        if (ok) break;
        if (clock64() - t0 > 2000000000ll) { atomicAdd(timeouts, 1u); break; }     // never hang the device
    }
    v[0] = __uint_as_float(a.x); v[1] = __uint_as_float(a.z); v[2] = __uint_as_float(b.x); v[3] = __uint_as_float(b.z);
    v[4] = __uint_as_float(c.x); v[5] = __uint_as_float(c.z); v[6] = __uint_as_float(d.x); v[7] = __uint_as_float(d.z);
}
__device__ __forceinline__ void mc_load8(const MPhase &p, const uint2 *mb, uint32_t epoch, unsigned *timeouts, int task, float (&v)[8]) {
    if (p.from_mbox) {
        mc_load8_mbox(mb + (size_t)task * 8, epoch, v, timeouts);
    } else {
        const float4 *src = (const float4 *)(p.b + (size_t)task * 8);
        const float4 f0 = __ldcg(src), f1 = __ldcg(src + 1);
        v[0] = f0.x; v[1] = f0.y; v[2] = f0.z; v[3] = f0.w; v[4] = f1.x; v[5] = f1.y; v[6] = f1.z; v[7] = f1.w;
    }
}

// ---- activations -> signed byte planes + per-block table -----------------------------------------------------------------
// task = (block, quarter): 8 consecutive floats; the four lanes of a block are adjacent.  T_k = rint(x_k * 2^(22-E)) as in
// K1 (lk_gemv.cu prep_x: same scale, same rounding, same range detection), stored as the signed base-256 digits of T.
// Plane bytes of block b, plane p, half h at xs + 96 b + 32 p + 16 h, in the order the A fragments present the elements:
//   Q4, Kotlin order (byte j = elements 2j, 2j+1):  [e0 e2 e4 e6 | e1 e3 e5 e7 | e8 e10 e12 e14 | e9 .. e15] + 16 h
//   Q4, ggml order   (byte j = elements j, j+16):   [e0..e3 | e16..e19 | e4..e7 | e20..e23] + 8 h
//   Q8_0:                                           [e0..e3 | e4..e7 | e8..e11 | e12..e15] + 16 h
// Table entry of block b (16 bytes at sc + 16 b): s, c * s * sumT, 65536 s, 0  (c = -8 for Q4_0, 1 for Q4_1, 0 for Q8_0).
// One round: the eight activations v of task `task` (block task >> 2, quarter task & 3; the four lanes of a block are adjacent)
template <int TYPE>
__device__ __forceinline__ void mc_convert8(const MPhase &p, uint8_t *xs, float4 *sc, uint32_t *mask, uint32_t *any, int lane, int task, bool on, const float (&v)[8]) {
    // max |v| over the block with NaN propagation: a NaN or inf element leaves exponent 255
    float amax = 0.f;
#pragma unroll
    for (int i = 0; i < 8; i += 2) amax = mc_fmax3_nan_abs(amax, v[i], v[i + 1]);
    uint32_t ab = __float_as_uint(amax) & 0x7FFFFFFFu;
    ab = max(ab, __shfl_xor_sync(0xffffffffu, ab, 1));
    ab = max(ab, __shfl_xor_sync(0xffffffffu, ab, 2));
    int e = (int)(ab >> 23);
    const bool nonfinite = e == 255;
    {   // range check (lk_gemv.cu GemvSide): the block's smallest non-zero |x| below 2^-fix_gap of its maximum -> flag
        float amin = __uint_as_float(0x7F800000u);
#pragma unroll
        for (int i = 0; i < 8; i += 2) amin = mc_fmin3_abs(amin, v[i], v[i + 1]);
        uint32_t mb = __float_as_uint(amin);
        mb = min(mb, __shfl_xor_sync(0xffffffffu, mb, 1));
        mb = min(mb, __shfl_xor_sync(0xffffffffu, mb, 2));
        if ((int)(mb >> 23) + p.fix_gap < e && !nonfinite) {     // all four lanes of the block branch alike
            uint32_t mn = 0xFFFFFFFFu;
#pragma unroll
            for (int i = 0; i < 8; i++) mn = min(mn, __float_as_uint(v[i]) * 2u - 1u);    // +-0 wraps to 0xFFFFFFFF
            const unsigned quad = 0xFu << (lane & ~3);
            mn = min(mn, __shfl_xor_sync(quad, mn, 1));
            mn = min(mn, __shfl_xor_sync(quad, mn, 2));
            const uint32_t t = mn + 1u;
            if (on && (task & 3) == 0 && t != 0u && (int)(t >> 24) + p.fix_gap < e) {
                atomicOr(&mask[task >> 7], 1u << ((task >> 2) & 31));
                *(volatile uint32_t *)any = 1u;
            }
        }
    }
    e = e < 24 ? 24 : e;
    const float c = nonfinite ? 0.f : __uint_as_float((uint32_t)(253 + 22 - e) << 23);
    const float s = nonfinite ? __uint_as_float(0x7FC00000u) : __uint_as_float((uint32_t)(e + 1 - 22) << 23);
    int T[8];
    int sumT = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const int t = (int)__float_as_uint(fmaf(v[i], c, 12582912.0f)) - 0x4B400000;
        sumT += t;
        T[i] = t + 0x8080;
    }
    sumT += __shfl_xor_sync(0xffffffffu, sumT, 1);
    sumT += __shfl_xor_sync(0xffffffffu, sumT, 2);
    if (on) {
        const int blk = task >> 2, q = task & 3;
        uint8_t *slot = xs + (size_t)blk * MC_XB;
#pragma unroll
        for (int pl = 0; pl < 3; pl++) {
            const uint32_t s2 = 0x0040u + 0x0011u * pl;       // byte pl of a, byte pl of b
            const uint32_t flip = pl < 2 ? 0x80808080u : 0u;
            uint32_t w0, w1;
            if (TYPE == T_Q8_0 || p.nib) {                      // consecutive elements per word
                w0 = __byte_perm(__byte_perm(T[0], T[1], s2), __byte_perm(T[2], T[3], s2), 0x5410) ^ flip;
                w1 = __byte_perm(__byte_perm(T[4], T[5], s2), __byte_perm(T[6], T[7], s2), 0x5410) ^ flip;
            } else {                                            // even elements, odd elements
                w0 = __byte_perm(__byte_perm(T[0], T[2], s2), __byte_perm(T[4], T[6], s2), 0x5410) ^ flip;
                w1 = __byte_perm(__byte_perm(T[1], T[3], s2), __byte_perm(T[5], T[7], s2), 0x5410) ^ flip;
            }
            if (TYPE != T_Q8_0 && p.nib) {                      // ggml order: elements 8q.. sit in half q & 1, words (q >> 1) and (q >> 1) + 2
                uint32_t *d = (uint32_t *)(slot + pl * 32 + (q & 1) * 16 + (q >> 1) * 4);
                d[0] = w0; d[2] = w1;
            } else {
                *(uint2 *)(slot + pl * 32 + (q >> 1) * 16 + (q & 1) * 8) = make_uint2(w0, w1);
            }
        }
        if (q == 0) sc[blk] = make_float4(s, TYPE == T_Q8_0 ? 0.f : (TYPE == T_Q4_0 ? -8.0f : 1.0f) * s * (float)sumT, 65536.0f * s, 0.f);
    }
}

// Every consumer warp converts the activations of ITS OWN K-slices (its block pairs of every column chunk: at most 128 tasks =
// four rounds of 32 lanes) and nobody else reads them in the stage loop, so a warp starts multiplying as soon as its own slice of
// the mailbox has arrived: no CTA-wide barrier between the hand-off and the first MMA, and a late CTA of the previous phase
// holds up only the warps whose slice it produces.  Everything a phase leaves in the header for its epilogue (range flags,
// sum-of-squares partials) and the partial-sum buffer are double-buffered by phase parity, because a fast warp is already in
// phase i + 1 while the epilogue threads of phase i still read them.
// RMSNorm prologue (pro_mode 1) WITHOUT 1/rms on the hand-off's critical path: the planes are built from y = x * gain, the
// sum of squares is taken on the side, and the positive scalar 1/rms multiplies the finished row sums in the epilogue
// (dst = 1/rms * sum_k w_k y_k).  The first version normalised before converting: load -> sum -> shuffle tree -> barrier ->
// 16 partials -> div, sqrt, div -> 16 multiplies in front of every conversion, 2.2 us of a 5 us hand-off even on an idle
// memory system (profiles/r02i_mchain_small_shapes.txt).  Costs one more rounding per output (<= 2^-24 relative).
template <int TYPE>
__device__ __forceinline__ void mc_prep_warp(const MPhase &p, uint8_t *smem, const MChainParams &cp, int lane, int warp, const uint2 *mb, uint32_t epoch, int ph,
                                             int nbp, int c0, int nchunk) {
    uint8_t *xs = smem + cp.x_off;
    float4 *sc = (float4 *)(smem + cp.sc_off);
    const int par = ph & 1;
    uint32_t *mask = (uint32_t *)(smem + MC_OFF_MASK) + par * MC_MASK_WORDS;
    uint32_t *any = (uint32_t *)(smem + MC_OFF_ANY) + par;
    const bool norm = p.pro_mode == 1;
    const bool gain = norm && p.pro_w != nullptr;
    // this warp's tasks (8 activations each): the eight tasks of every block pair of its slice of every chunk, flattened
    int nt = 0;
#pragma unroll 1
    for (int ch = 0; ch < nchunk; ch++) { const int cl = min(c0, nbp - ch * c0); nt += 8 * (((warp + 1) * cl) / MC_W - (warp * cl) / MC_W); }
    float ssq = 0.f;
#pragma unroll 1
    for (int r = 0; 32 * r < nt; r++) {               // K = 4096: one round; K = 11008: three (one L2 round trip each)
        const int i = 32 * r + lane;
        const bool on = i < nt;
        int task = 0;
        {
            int rem = i;
#pragma unroll 1
            for (int ch = 0; ch < nchunk; ch++) {
                const int cs = ch * c0, cl = min(c0, nbp - cs);
                const int pw = cs + (warp * cl) / MC_W, n = 8 * (cs + ((warp + 1) * cl) / MC_W - pw);
                if (rem < n || ch == nchunk - 1) { task = 8 * pw + rem; break; }
                rem -= n;
            }
        }
        float v[8];
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = 0.f;
        float4 g0 = make_float4(1.f, 1.f, 1.f, 1.f), g1 = g0;
        if (on) {
            if (gain) { const float4 *ws = (const float4 *)(p.pro_w + (size_t)task * 8); g0 = __ldcg(ws); g1 = __ldcg(ws + 1); }
            mc_load8(p, mb, epoch, cp.bar + 2, task, v);
        }
#ifdef LK_GEMV_TIMELINE
        if (threadIdx.x == 0 && r == 0) { if (v[0] + v[3] == 12345.f) xs[0] = 1; MC_TL(ph, 1); }      // first activations loaded
#endif
        if (norm) {
#pragma unroll
            for (int k = 0; k < 8; k++) ssq = fmaf(v[k], v[k], ssq);
            if (gain) {
                v[0] = __fmul_rn(v[0], g0.x); v[1] = __fmul_rn(v[1], g0.y); v[2] = __fmul_rn(v[2], g0.z); v[3] = __fmul_rn(v[3], g0.w);
                v[4] = __fmul_rn(v[4], g1.x); v[5] = __fmul_rn(v[5], g1.y); v[6] = __fmul_rn(v[6], g1.z); v[7] = __fmul_rn(v[7], g1.w);
            }
        }
        mc_convert8<TYPE>(p, xs, sc, mask, any, lane, task, on, v);
    }
    if (norm) {
        ssq = warp_sum(ssq);
        if (lane == 0) ((float *)(smem + MC_OFF_RMS))[par * MC_W + warp] = ssq;     // summed in the epilogue
    }
    __syncwarp();
}

// ---- cold path of the range fix-up: sum_k w(i, k) * r_k, r_k = x_k - T_k s, for the flagged blocks of this CTA's rows,
// into fix-up column j = (rank of the block among the flagged ones) & 7 of the row's partial sums (the arithmetic of
// lk_gemv.cu gemv_fixup_body; the per-block sum is a shuffle tree here instead of a serial chain).  Warp w owns the rows
// w, w + 16, ... of the CTA's share (all segments) for EVERY flagged block, lane k owns element k of the block: one load of
// the residual per lane and block, one byte and one scale per lane and row, so even a single flagged block is spread over the
// whole CTA (the first version looped over the 32 elements inside one thread per row with a dependent L2 load each: ~9 us
// per phase with one or two blocks flagged, twice the phase itself; profiles/r02m_layer_types.txt, Q4_1).
template <int TYPE>
__device__ __noinline__ void mc_fixup(const MPhase *pp, const MChainParams *cpp, uint8_t *smem, int r_begin, int nrows, const uint2 *mb, int par, float *part) {
    const MPhase &p = *pp;
    const float4 *sc = (const float4 *)(smem + cpp->sc_off);
    const uint32_t *mask = (const uint32_t *)(smem + MC_OFF_MASK) + par * MC_MASK_WORDS;
    const int nw = (p.nblk + 31) >> 5;
    constexpr int BB = block_bytes(TYPE);
    constexpr int QOFF = TYPE == T_Q4_1 ? 4 : 2;
    const int warp = (int)threadIdx.x >> 5, lane = (int)threadIdx.x & 31;
    const int total = p.nseg * nrows;
    // this lane's byte and nibble of a block (Q4: Kotlin order byte k / 2, ggml order byte k % 16)
    const int qb = TYPE == T_Q8_0 ? lane : (p.nib ? (lane & 15) : (lane >> 1));
    const bool hi = TYPE != T_Q8_0 && (p.nib ? (lane >> 4) : (lane & 1));
    // my rows' fix-up columns start at zero
    for (int rr = warp + MC_W * (lane >> 3); rr < total; rr += MC_W * 4) {
        const int seg = rr >= 2 * nrows ? 2 : rr >= nrows ? 1 : 0;
        part[(size_t)(seg * p.rows_max + rr - seg * nrows) * MC_PS + MC_W + (lane & 7)] = 0.f;
    }
    __syncwarp();
    int idx = 0;
#pragma unroll 1
    for (int w = 0; w < nw; w++) {
        uint32_t m = mask[w];
#pragma unroll 1
        while (m) {
            const int bit = __ffs((int)m) - 1;
            m &= m - 1;
            const int j = idx++ & 7;
            const int blk = w * 32 + bit;
            const float s = sc[blk].x;
            const float c = __frcp_rn(s);                                 // a power of two: exact
            const int kk = blk * 32 + lane;
            float v = p.from_mbox ? __uint_as_float(mc_ld_pair(mb + kk).x) : __ldcg(p.b + kk);
            if (p.pro_mode == 1 && p.pro_w) v = __fmul_rn(v, __ldcg(p.pro_w + kk));      // y = x * gain, as in mc_prep_warp
            const int T = (int)__float_as_uint(fmaf(v, c, 12582912.0f)) - 0x4B400000;
            const float rk = fmaf(-(float)T, s, v);                       // exact residual of this lane's element
#pragma unroll 2
            for (int rr = warp; rr < total; rr += MC_W) {
                const int seg = rr >= 2 * nrows ? 2 : rr >= nrows ? 1 : 0;
                const int r = rr - seg * nrows;
                const uint8_t *bp = p.a[seg] + (size_t)(r_begin + r) * p.row_bytes + (size_t)blk * BB;
                const uint32_t hdr = __ldg((const unsigned short *)bp) | (TYPE == T_Q4_1 ? (uint32_t)__ldg((const unsigned short *)bp + 1) << 16 : 0u);
                const uint32_t byte = __ldg(bp + QOFF + qb);
                const float d = h2f((uint16_t)(hdr & 0xFFFFu));
                float wv;                                                 // the reference's weight, core/GGMLComputeOps.kt:43-145
                if (TYPE == T_Q8_0) {
                    wv = __fmul_rn(d, (float)(int8_t)byte);
                } else {
                    const float q = (float)(hi ? (byte >> 4) : (byte & 15u));
                    wv = TYPE == T_Q4_0 ? __fmul_rn(d, __fsub_rn(q, 8.0f)) : __fadd_rn(__fmul_rn(d, q), h2f((uint16_t)(hdr >> 16)));
                }
                const float sum = warp_sum(__fmul_rn(wv, rk));
                if (lane == 0) part[(size_t)(seg * p.rows_max + r) * MC_PS + MC_W + j] += sum;
            }
            __syncwarp();
        }
    }
}

// ---- one block pair of a strip: rows g (and g + 8 if TWO) x the pair's 64 activations -------------------------------------
// pa0 / pa1: shared addresses of this lane's first quant word (aligned down) in rows g and g + 8; pd0 / pd1: of the block's
// first word (d, and m for Q4_1); b: this lane's B fragments of the pair's two MMAs; s: this lane's table entry half.
template <int TYPE, bool TWO>
__device__ __forceinline__ void mc_pair(uint32_t pa0, uint32_t pa1, uint32_t pd0, uint32_t pd1, const uint4 &b, const float2 &s, uint32_t sel,
                                        float &acc0, float &acc1) {
    int c[4];
    if (TYPE == T_Q8_0) {
        // 16 signed quant bytes per lane and row: four words, word-aligned for the odd block of a pair only
        const uint32_t u0 = mc_lds32(pa0), u1 = mc_lds32(pa0 + 4), u2 = mc_lds32(pa0 + 8), u3 = mc_lds32(pa0 + 12), u4 = mc_lds32(pa0 + 16);
        uint32_t r[4] = {0, 0, 0, 0};
        if (TWO) {
            const uint32_t v0 = mc_lds32(pa1), v1 = mc_lds32(pa1 + 4), v2 = mc_lds32(pa1 + 8), v3 = mc_lds32(pa1 + 12), v4 = mc_lds32(pa1 + 16);
            r[0] = __byte_perm(v0, v1, sel); r[1] = __byte_perm(v1, v2, sel); r[2] = __byte_perm(v2, v3, sel); r[3] = __byte_perm(v3, v4, sel);
        }
        mc_imma0<true>(c, __byte_perm(u0, u1, sel), r[0], __byte_perm(u1, u2, sel), r[1], b.x, b.y);
        mc_imma<true>(c, __byte_perm(u2, u3, sel), r[2], __byte_perm(u3, u4, sel), r[3], b.z, b.w);
        const float d0 = h2f((uint16_t)mc_lds16(pd0));
        acc0 = fmaf(d0, (float)(c[0] + (c[1] << 8)) * s.x, acc0);
        if (TWO) {
            const float d1 = h2f((uint16_t)mc_lds16(pd1));
            acc1 = fmaf(d1, (float)(c[2] + (c[3] << 8)) * s.x, acc1);
        }
    } else {
        // eight quant bytes (16 nibbles) per lane and row; Q4_0 blocks are word-aligned for the odd block of a pair only
        uint32_t q0, q1, r0 = 0, r1 = 0;
        if (TYPE == T_Q4_0) {
            const uint32_t u0 = mc_lds32(pa0), u1 = mc_lds32(pa0 + 4), u2 = mc_lds32(pa0 + 8);
            q0 = __byte_perm(u0, u1, sel); q1 = __byte_perm(u1, u2, sel);
            if (TWO) {
                const uint32_t v0 = mc_lds32(pa1), v1 = mc_lds32(pa1 + 4), v2 = mc_lds32(pa1 + 8);
                r0 = __byte_perm(v0, v1, sel); r1 = __byte_perm(v1, v2, sel);
            }
        } else {                                  // Q4_1: 20-byte blocks, quants at +4: always aligned
            q0 = mc_lds32(pa0); q1 = mc_lds32(pa0 + 4);
            if (TWO) { r0 = mc_lds32(pa1); r1 = mc_lds32(pa1 + 4); }
        }
        mc_imma0<false>(c, q0 & 0x0F0F0F0Fu, r0 & 0x0F0F0F0Fu, (q0 >> 4) & 0x0F0F0F0Fu, (r0 >> 4) & 0x0F0F0F0Fu, b.x, b.y);
        mc_imma<false>(c, q1 & 0x0F0F0F0Fu, r1 & 0x0F0F0F0Fu, (q1 >> 4) & 0x0F0F0F0Fu, (r1 >> 4) & 0x0F0F0F0Fu, b.z, b.w);
        if (TYPE == T_Q4_0) {                     // d * (s I - 8 s sumT)
            const float d0 = h2f((uint16_t)mc_lds16(pd0));
            acc0 = fmaf(d0, fmaf((float)(c[0] + (c[1] << 8)), s.x, s.y), acc0);
            if (TWO) {
                const float d1 = h2f((uint16_t)mc_lds16(pd1));
                acc1 = fmaf(d1, fmaf((float)(c[2] + (c[3] << 8)), s.x, s.y), acc1);
            }
        } else {                                  // d s I + m s sumT
            const uint32_t w0 = mc_lds32(pd0);
            acc0 = fmaf(h2f((uint16_t)(w0 & 0xFFFFu)), (float)(c[0] + (c[1] << 8)) * s.x, acc0);
            acc0 = fmaf(h2f((uint16_t)(w0 >> 16)), s.y, acc0);
            if (TWO) {
                const uint32_t w1 = mc_lds32(pd1);
                acc1 = fmaf(h2f((uint16_t)(w1 & 0xFFFFu)), (float)(c[2] + (c[3] << 8)) * s.x, acc1);
                acc1 = fmaf(h2f((uint16_t)(w1 >> 16)), s.y, acc1);
            }
        }
    }
}

// ---- one stage: rows [0, n) of a strip x this warp's block pairs [0, npairs) --------------------------------------------
template <int TYPE, bool TWO>
__device__ __forceinline__ void mc_pairs(uint32_t pa0, uint32_t pa1, uint32_t pd0, uint32_t pd1, uint32_t pb, uint32_t bstep, uint32_t psc,
                                         int npairs, uint32_t sel, float &acc0, float &acc1) {
    constexpr uint32_t PAIR_B = 2 * block_bytes(TYPE);
#pragma unroll 2
    for (int i = 0; i < npairs; i++) {
        const uint4 b = mc_lds128(pb);
        const float2 s = mc_lds64f(psc);
        mc_pair<TYPE, TWO>(pa0, pa1, pd0, pd1, b, s, sel, acc0, acc1);
        pa0 += PAIR_B; pa1 += PAIR_B; pd0 += PAIR_B; pd1 += PAIR_B; pb += bstep; psc += 32;
    }
}

// The same for a K-slice of exactly four block pairs (K = 4096 over 16 warps) whose B fragments and table entries the warp
// holds in registers for the whole phase: fully unrolled, 8 shared-memory loads per pair instead of 10 (16 wavefronts -> 10).
template <int TYPE, bool TWO>
__device__ __forceinline__ void mc_pairs4(uint32_t pa0, uint32_t pa1, uint32_t pd0, uint32_t pd1, const uint4 (&breg)[4], const float2 (&sreg)[4],
                                          uint32_t sel, float &acc0, float &acc1) {
    constexpr uint32_t PAIR_B = 2 * block_bytes(TYPE);
#pragma unroll
    for (int i = 0; i < 4; i++)
        mc_pair<TYPE, TWO>(pa0 + PAIR_B * i, pa1 + PAIR_B * i, pd0 + PAIR_B * i, pd1 + PAIR_B * i, breg[i], sreg[i], sel, acc0, acc1);
}

template <int TYPE>
__global__ void __launch_bounds__((MC_W + MC_PW) * 32, 1) gemv_mchain_kernel(const __grid_constant__ MChainParams cp) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t *full = (uint64_t *)smem;
    uint64_t *empty = full + MC_NB;
    volatile uint32_t *stage_off = (volatile uint32_t *)(smem + MC_OFF_STAGE);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int bid = (int)blockIdx.x;
    uint8_t *ring = smem + cp.ring_off;

    if (warp >= MC_W) {
        // ===== producer warps: one bulk copy per row; the four warps run the same ring bookkeeping in registers (it is a
        // deterministic function of the stage sequence and of the consumers' releases) and each issues every fourth row of a
        // stage — per-lane cp.async.bulk is issued one lane at a time (ELECT loop, ~50 cycles each), so one warp issuing all 16
        // rows of a stage capped the stream (measured: profiles/r02i_mchain_timeline3.txt) =====
        const int pw = warp - MC_W;
        if (pw == 0) {
            if (lane == 0) {
#pragma unroll 1
                for (int i = 0; i < MC_NB; i++) { mbar_init(&full[i], 1); mbar_init(&empty[i], MC_W); }
                fence_mbar_init();
            }
            if (lane < 4) ((uint32_t *)(smem + MC_OFF_ZERO))[lane] = 0u;
        }
        __syncwarp();
        __syncthreads();   // consumers may now wait on the barriers
        if (cp.wait_mode == 0) pdl_wait();       // a weight was written by an op that may still run
        const uint64_t pol_w = l2_policy_evict_first();
        uint32_t qs[MC_NB], qe[MC_NB];           // ring bytes of the stage in barrier slot j (valid while it is outstanding)
#pragma unroll
        for (int j = 0; j < MC_NB; j++) { qs[j] = 0; qe[j] = 0; }
        int head = 0, si = 0;
        uint32_t wr = 0;
#pragma unroll 1
        for (int ph = 0; ph < cp.nphase; ph++) {
            const MPhase &p = cp.ph[ph];
            const int r_begin = p.rows_per_cta * bid + min(bid, p.rows_rem);
            const int nrows = p.rows_per_cta + (bid < p.rows_rem ? 1 : 0);
            constexpr int PAIR_B = 2 * block_bytes(TYPE);
            const int stride = p.stride, row_bytes = p.row_bytes, nchunk = p.nchunk, c0 = p.c0, nbp = p.nblk >> 1, sr = p.sr, nseg = p.nseg;
#pragma unroll 1
            for (int seg = 0; seg < nseg; seg++) {
                const uint8_t *src = p.a[seg] + (size_t)r_begin * row_bytes;
#pragma unroll 1
                for (int left = nrows; left > 0;) {
                    const int rows = min(sr, left);
#pragma unroll 1
                    for (int ch = 0; ch < nchunk; ch++) {
                        const uint32_t cbytes = (uint32_t)(min(c0, nbp - ch * c0) * PAIR_B);     // bytes of one row of this chunk
                        const uint8_t *csrc = src + (size_t)ch * c0 * PAIR_B;
                        const uint32_t bytes = (uint32_t)(rows * stride);
                        if (wr + bytes > (uint32_t)cp.ring_bytes) wr = 0;
                        // the slot's barrier pair and the ring bytes must be free: wait for the oldest stages, in order, until they are
                        while (true) {
                            bool need = si - head >= MC_NB;
#pragma unroll
                            for (int j = 0; j < MC_NB; j++) {
                                const bool live = ((j - head) & (MC_NB - 1)) < si - head;
                                need = need || (live && wr < qe[j] && qs[j] < wr + bytes);
                            }
                            if (!need) break;
                            mbar_wait(&empty[head & (MC_NB - 1)], (uint32_t)(head / MC_NB) & 1u);     // (a __nanosleep between polls cost 1 %: the reaction time matters more than the issue slots)
                            head++;
                        }
                        // Pace the stream: at most `inflight` stages issued and not yet landed.  More requests in flight add no
                        // bandwidth (64 KB per SM saturate HBM, tools/ubench_tma.cu), only queueing delay in L2 / HBM, and every
                        // load on the consumers' critical path at a hand-off (activations, parameters, instructions) waits in
                        // that same queue: with the whole ring requested at once a hand-off took 6-7 us even with the dot
                        // products compiled out (profiles/r02i_mchain_stream_only.txt).
                        if (si >= cp.inflight) { const int sj = si - cp.inflight; mbar_wait(&full[sj & (MC_NB - 1)], (uint32_t)(sj / MC_NB) & 1u); }
                        const int slot = si & (MC_NB - 1);
#pragma unroll
                        for (int j = 0; j < MC_NB; j++) if (j == slot) { qs[j] = wr; qe[j] = wr + bytes; }
                        if (pw == 0 && lane == 0) {
                            stage_off[slot] = wr;
                            mbar_arrive_expect_tx(&full[slot], (uint32_t)rows * cbytes);
                        }
                        const int r = pw + 4 * lane;
                        if (lane < 4 && r < rows) tma_load_1d(ring + wr + (size_t)r * stride, csrc + (size_t)r * row_bytes, cbytes, &full[slot], pol_w);
                        wr += bytes; si++;
                    }
                    src += (size_t)rows * row_bytes; left -= rows;
                }
            }
            if (pw == 0 && lane == 0) MC_TL(ph, 7);
        }
        return;
    }

    // ===== consumer warps =====
    if (tid < 2 * MC_MASK_WORDS) ((uint32_t *)(smem + MC_OFF_MASK))[tid] = 0u;
    if (tid < 2) ((uint32_t *)(smem + MC_OFF_ANY))[tid] = 0u;
    __syncthreads();   // barriers initialised by the producer
#ifdef LK_GEMV_TIMELINE
    if (tid == 0 && cp.tl) {
        unsigned long long *q_ = cp.tl + (size_t)blockIdx.x * MC_TL_STRIDE, t_;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t_));
        unsigned sm_; asm volatile("mov.u32 %0, %%smid;" : "=r"(sm_));
        q_[0] = t_; q_[1] = (unsigned long long)clock64(); q_[2] = sm_;
    }
#endif
    pdl_wait();                // phase 0's activations may be the previous kernel's output
    pdl_launch_dependents();
    // epoch base of this launch (see the hand-off notes above): read after the wait, when every earlier launch has added its grid
    unsigned long long *const bar_cnt = (unsigned long long *)cp.bar;
    if (tid == 0) *(volatile uint32_t *)(smem + MC_OFF_EPOCH) = (uint32_t)(mc_ld_acquire_gpu(bar_cnt) / gridDim.x) * 8u;
    asm volatile("bar.sync 1, %0;" ::"r"(MC_NCT) : "memory");
    const uint32_t epoch0 = *(volatile uint32_t *)(smem + MC_OFF_EPOCH);
    const uint32_t smem_base = smem_u32(smem);
    const int g = lane >> 2, t = lane & 3;
    const int bj = t >> 1, bh = t & 1;                     // A side: block of the pair, half of the block
    // B side: column n = lane >> 2 = 4 jc + plane, k-slots of lane t belong to block t >> 1: only matching lanes hold non-zero bytes
    const bool b_nz = (g & 3) < 3 && (g >> 2) == bj;
    constexpr int BB = block_bytes(TYPE);
    constexpr int QOFF = TYPE == T_Q4_1 ? 4 : 2;
    // first quant byte of this lane within the pair, aligned down to a word; PRMT selector for the realignment
    const int qbyte = bj * BB + QOFF + bh * (TYPE == T_Q8_0 ? 16 : 8);
    const uint32_t sel = (qbyte & 2) ? 0x5432u : 0x3210u;
    const uint32_t a_lane = (uint32_t)(qbyte & ~3), d_lane = (uint32_t)(bj * BB);       // + pair offset within the staged chunk
    const uint32_t xb_lane = b_nz ? smem_base + cp.x_off + (uint32_t)(bj * MC_XB + (g & 3) * 32 + bh * 16) : smem_base + MC_OFF_ZERO;
    const uint32_t bstep = b_nz ? 2u * MC_XB : 0u;
    const uint32_t sc_lane = smem_base + cp.sc_off + (uint32_t)(bj * 16 + bh * 8);
    const uint32_t ring_base = smem_base + cp.ring_off;
    int si = 0;
#pragma unroll 1
    for (int ph = 0; ph < cp.nphase; ph++) {
        const MPhase &p = cp.ph[ph];
        if (tid == 0) MC_TL(ph, 0);
        if (tid == 32 && ph + 1 < cp.nphase) {   // warm the constant cache with the next phase's parameter lines
            const int *q = (const int *)&cp.ph[ph + 1];      // sizeof(MPhase) = 184, not line-aligned: up to four 64-byte lines
            const int t0 = q[0], t1 = q[16], t2 = q[32], t3 = q[sizeof(MPhase) / 4 - 1];
            asm volatile("" ::"r"(t0), "r"(t1), "r"(t2), "r"(t3));
        }
        const int par = ph & 1;
        const int r_begin = p.rows_per_cta * bid + min(bid, p.rows_rem);
        const int nrows = p.rows_per_cta + (bid < p.rows_rem ? 1 : 0);
        const uint2 *mb = cp.mbox + (size_t)par * MC_KMAX;          // this phase's activations, if they come by mailbox
        const uint32_t epoch = epoch0 + (uint32_t)ph + 1u;
        const bool norm = p.pro_mode == 1;
        // this warp's K-slice of column chunk ch (block pairs [ch c0, ch c0 + cl), cl = min(c0, nbp - ch c0)): pairs
        // [ch c0 + (warp cl) / 16, ch c0 + ((warp + 1) cl) / 16)
        const int nbp = p.nblk >> 1, c0 = p.c0, nchunk = p.nchunk, stride = p.stride, sr = p.sr, nseg = p.nseg, rows_max = p.rows_max;
        const int pw0 = (warp * min(c0, nbp)) / MC_W;           // first pair of the slice of chunk 0
        // side loads whose latency hides under the stage loop: the residual of this thread's output row, and this CTA's 1/grid
        // slice of the normalised activations (the unfused RMSNorm node's tensor, stored in the epilogue)
        float res_pre = 0.f;
        if (p.residual && tid < nrows) res_pre = __ldcg(p.residual + r_begin + tid);
        float bo_g = 1.f;
        uint2 bo_x = make_uint2(0u, epoch);
        const int bo_per = (p.K + (int)gridDim.x - 1) / (int)gridDim.x;
        const int bo_k = bid * bo_per + tid;
        const bool bo_on = norm && p.b_out && tid < bo_per && bo_k < p.K;
        if (bo_on) {
            if (p.from_mbox) bo_x = mc_ld_pair(mb + bo_k); else bo_x.x = __float_as_uint(__ldcg(p.b + bo_k));
            if (p.pro_w) bo_g = __ldcg(p.pro_w + bo_k);
        }
        // activations of this warp's slices -> planes + table (no CTA barrier: see mc_prep_warp)
        mc_prep_warp<TYPE>(p, smem, cp, lane, warp, mb, epoch, ph, nbp, c0, nchunk);
        if (tid == 0) MC_TL(ph, 3);
        // K = 4096: four pairs per warp, B fragments and table entries stay in registers for the phase
        const bool res4 = nchunk == 1 && nbp == 4 * MC_W;
        uint4 breg[4];
        float2 sreg[4];
        if (res4) {
#pragma unroll
            for (int i = 0; i < 4; i++) {
                breg[i] = mc_lds128(xb_lane + (uint32_t)(pw0 + i) * bstep);
                sreg[i] = mc_lds64f(sc_lane + (uint32_t)(pw0 + i) * 32u);
            }
        }
#ifdef LK_GEMV_TIMELINE
        bool first_stage = true;
#endif
        float *part = (float *)(smem + cp.part_off) + (size_t)par * cp.part_floats;     // by phase parity (see mc_prep_warp)
        float *pr_lane = part + (size_t)g * MC_PS + warp;      // this lane's partial-sum slot of strip row g
#pragma unroll 1
        for (int seg = 0; seg < nseg; seg++) {
            float *pr = pr_lane + (size_t)seg * rows_max * MC_PS;
#pragma unroll 1
            for (int row0 = 0; row0 < nrows; row0 += sr, pr += sr * MC_PS) {
                const int n = min(sr, nrows - row0);
                const uint32_t ro_a = (uint32_t)((g < n ? g : 0) * stride), ro_b = (uint32_t)((g + 8 < n ? g + 8 : 0) * stride);
#pragma unroll 1
                for (int ch = 0; ch < nchunk; ch++) {
                    const uint32_t slot8 = (uint32_t)(si & (MC_NB - 1)) * 8u;
                    mc_mbar_wait(smem_base + slot8, (uint32_t)(si >> 3) & 1u);
#ifdef LK_GEMV_TIMELINE
                    if (first_stage && tid == 0) { MC_TL(ph, 4); first_stage = false; }
#endif
                    const uint32_t sbase = ring_base + mc_lds32(smem_base + MC_OFF_STAGE + (slot8 >> 1));
                    const int cs = ch * c0, cl = min(c0, nbp - cs);
                    const int pw = cs + (warp * cl) / MC_W, npw = cs + ((warp + 1) * cl) / MC_W - pw;
                    const uint32_t coff = (uint32_t)((pw - cs) * 2 * BB);     // this warp's first pair within the staged chunk
                    const uint32_t pa0 = sbase + ro_a + coff + a_lane, pa1 = sbase + ro_b + coff + a_lane;
                    const uint32_t pd0 = sbase + ro_a + coff + d_lane, pd1 = sbase + ro_b + coff + d_lane;
                    float acc0 = 0.f, acc1 = 0.f;
#ifdef MC_SKIP_COMPUTE      // tools only: the stream alone (producer + ring + releases), no dot products
                    if (false) {} else if (n < 0)
#endif
                    if (res4) {
                        if (n > 8) mc_pairs4<TYPE, true>(pa0, pa1, pd0, pd1, breg, sreg, sel, acc0, acc1);
                        else mc_pairs4<TYPE, false>(pa0, pa1, pd0, pd1, breg, sreg, sel, acc0, acc1);
                    } else {
                        const uint32_t pb = xb_lane + (uint32_t)pw * bstep, psc = sc_lane + (uint32_t)pw * 32u;
                        if (n > 8) mc_pairs<TYPE, true>(pa0, pa1, pd0, pd1, pb, bstep, psc, npw, sel, acc0, acc1);
                        else mc_pairs<TYPE, false>(pa0, pa1, pd0, pd1, pb, bstep, psc, npw, sel, acc0, acc1);
                    }
                    acc0 += __shfl_xor_sync(0xffffffffu, acc0, 1);
                    acc1 += __shfl_xor_sync(0xffffffffu, acc1, 1);
                    acc0 += __shfl_xor_sync(0xffffffffu, acc0, 2);
                    acc1 += __shfl_xor_sync(0xffffffffu, acc1, 2);
                    if (t == 0) {        // chunk 1 adds to what this same lane parked for chunk 0
                        if (g < n) pr[0] = ch == 0 ? acc0 : pr[0] + acc0;
                        if (g + 8 < n) pr[8 * MC_PS] = ch == 0 ? acc1 : pr[8 * MC_PS] + acc1;
                    }
                    __syncwarp();
                    if (lane == 0) mc_mbar_arrive(smem_base + 64u + slot8);    // this warp is done reading the stage
                    si++;
                }
            }
        }
        if (bo_on && bo_x.y != epoch) {     // the slice element had not arrived when it was fetched before the loop: poll it now
            const long long t0 = clock64();
            do { bo_x = mc_ld_pair(mb + bo_k); } while (bo_x.y != epoch && clock64() - t0 < 2000000000ll);
        }
        if (tid == 0) MC_TL(ph, 5);
        asm volatile("bar.sync 1, %0;" ::"r"(MC_NCT) : "memory");   // every K-slice of every row is parked (and every warp has converted its activations)
        if (tid == 0) MC_TL(ph, 2);
        const bool flagged = ((volatile uint32_t *)(smem + MC_OFF_ANY))[par] != 0u;
        if (__builtin_expect(flagged, 0)) {
            mc_fixup<TYPE>(&p, &cp, smem, r_begin, nrows, mb, par, part);
            asm volatile("bar.sync 1, %0;" ::"r"(MC_NCT) : "memory");
            // re-arm this parity's flags: the next phase uses the other set, the one after it starts behind another CTA barrier
            if (tid < MC_MASK_WORDS) ((uint32_t *)(smem + MC_OFF_MASK))[par * MC_MASK_WORDS + tid] = 0u;
            if (tid == MC_MASK_WORDS) ((uint32_t *)(smem + MC_OFF_ANY))[par] = 0u;
        }
        // ---- fixed-order reduction over the K-slices + epilogue, one thread per output ----
        float rinv = 1.0f;
        if (norm && (tid < nseg * nrows || bo_on)) {      // 1/rms from the warps' sums of squares (MUFU.RSQ: <= 2 ulp)
            const float *q = (const float *)(smem + MC_OFF_RMS) + par * MC_W;
            float tot = ((q[0] + q[1]) + (q[2] + q[3])) + ((q[4] + q[5]) + (q[6] + q[7]));
            tot += ((q[8] + q[9]) + (q[10] + q[11])) + ((q[12] + q[13]) + (q[14] + q[15]));
            rinv = rsqrtf(fmaf(tot, p.inv_k, p.pro_eps));
        }
        auto rowsum = [&](int prow) {
            const float *q = part + (size_t)prow * MC_PS;
            float v = ((q[0] + q[1]) + (q[2] + q[3])) + ((q[4] + q[5]) + (q[6] + q[7]));
            v += ((q[8] + q[9]) + (q[10] + q[11])) + ((q[12] + q[13]) + (q[14] + q[15]));
            if (flagged) v += ((q[16] + q[17]) + (q[18] + q[19])) + ((q[20] + q[21]) + (q[22] + q[23]));
            return norm ? __fmul_rn(v, rinv) : v;
        };
        const int feed = p.feed;
        uint2 *mb_next = cp.mbox + (size_t)((ph + 1) & 1) * MC_KMAX;
        if (p.epi_mode == 1) {
            // gate/up pair (SwiGLU, LlamaMLP.forward model/LlamaModel.kt:159-176) with the elementwise kernels' roundings
            for (int r = tid; r < nrows; r += MC_NCT) {
                const float gv = rowsum(r), uv = rowsum(rows_max + r);
                const int64_t o = r_begin + r;
                const float act = mc_silu(gv);
                const float hv = __fmul_rn(act, uv);
                if (feed == 4) mc_st_pair(mb_next + o, hv, epoch + 1u);      // first: the next phase waits for this store
                p.epi_out[o] = hv;
                p.dst[0][o] = gv;
                p.dst[1][o] = uv;
                if (p.epi_act) p.epi_act[o] = act;
            }
        } else {
            for (int idx = tid; idx < nseg * nrows; idx += MC_NCT) {
                const int seg = idx >= 2 * nrows ? 2 : idx >= nrows ? 1 : 0;
                const int r = idx - seg * nrows;
                float v = rowsum(seg * rows_max + r);
                const int64_t o = r_begin + r;
                if (p.residual) {              // fused residual ADD (computeAdd's single F32 add); no stacked segments with a residual
                    if (p.dst_raw) p.dst_raw[o] = v;
                    v = __fadd_rn(idx == tid ? res_pre : __ldcg(p.residual + o), v);
                }
                if (feed == 1 + seg) mc_st_pair(mb_next + o, v, epoch + 1u);
                p.dst[seg][o] = v;
            }
        }
        if (bo_on) p.b_out[bo_k] = __fmul_rn(__fmul_rn(__uint_as_float(bo_x.x), rinv), bo_g);
        if (tid == 0) MC_TL(ph, 6);
    }
    if (tid == 0) asm volatile("red.relaxed.gpu.global.add.u64 [%0], %1;" ::"l"(bar_cnt), "l"(1ull) : "memory");   // this CTA is done with the launch
#ifdef LK_GEMV_TIMELINE
    if (tid == 0 && cp.tl) { unsigned long long t_; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t_)); cp.tl[(size_t)blockIdx.x * MC_TL_STRIDE + 3] = t_; }
#endif
}

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
cudaError_t launch_gemv_mchain(int type, int n, const GemvChainLink *links, const GemvTuning &tune, int sm_count, cudaStream_t s, unsigned *bar,
                               void *mbox) {
    if (n < 1 || n > MC_MAX || !links || !bar || !mbox) return cudaErrorNotSupported;
    if (type != T_Q4_0 && type != T_Q4_1 && type != T_Q8_0) return cudaErrorNotSupported;
    if (tune.act_planes == 2) return cudaErrorNotSupported;      // the opt-in 16-bit mode stays on K1c
    const int bb = block_bytes(type);
    const int grid = sm_count;
    MChainParams cp = {};
    int part_rows = 0, nblk_max = 0;
    for (int i = 0; i < n; i++) {
        const GemvChainLink &L = links[i];
        const GemvFusion &f = L.fu;
        MPhase &p = cp.ph[i];
        const int64_t K = L.K, M = L.M;
        if (K <= 0 || M <= 0 || (K % 256) != 0 || K > 32 * 32 * MC_MASK_WORDS || M > (1 << 30)) return cudaErrorNotSupported;
        if (((uintptr_t)L.a & 15) || ((uintptr_t)L.b & 15) || ((uintptr_t)f.pro_w & 15) || ((uintptr_t)f.b_out & 15)) return cudaErrorNotSupported;
        p.a[0] = (const uint8_t *)L.a; p.b = L.b; p.dst[0] = L.dst;
        p.nib = tune.nibble_order ? 1 : 0;
        p.K = (int)K; p.M = (int)M; p.inv_k = (float)(1.0 / (double)K);
        p.pro_mode = f.pro_mode; p.pro_w = f.pro_w; p.pro_eps = f.pro_eps; p.b_out = f.b_out;
        p.residual = f.residual; p.dst_raw = f.dst_raw;
        if (f.nseg_extra < 0 || f.nseg_extra > 2 || (f.nseg_extra > 0 && f.residual)) return cudaErrorNotSupported;
        for (int j = 0; j < f.nseg_extra; j++) {
            if (!f.a_extra[j] || !f.dst_extra[j] || ((uintptr_t)f.a_extra[j] & 15)) return cudaErrorNotSupported;
            p.a[j + 1] = (const uint8_t *)f.a_extra[j];
            p.dst[j + 1] = f.dst_extra[j];
        }
        p.nseg = 1 + f.nseg_extra;
        if (f.epi_mode == 1) {
            if (f.nseg_extra != 1 || !f.epi_out) return cudaErrorNotSupported;
            p.epi_mode = 1; p.epi_act = f.epi_act; p.epi_out = f.epi_out;
        }
        p.fix_gap = tune.fix_gap >= 255 ? 255 : std::max(2, tune.fix_gap);
        p.nblk = (int)(K / 32);
        p.row_bytes = p.nblk * bb;
        if (p.row_bytes % 16) return cudaErrorNotSupported;                  // one bulk copy per row
        // a strip of 16 rows is staged whole if it fits ~42 KB, else as up to 8 equal column chunks (boundaries at multiples of 4
        // block pairs: every row chunk is then a 16-byte multiple at a 16-byte aligned address)
        const int nbp = p.nblk / 2;
        p.sr = 16;
        p.nchunk = 1; p.c0 = nbp;
        while (p.nchunk < 8 && 16 * (p.c0 * 2 * bb + 16) > 42 * 1024) {      // (nbp is a multiple of 4: K % 256 == 0)
            p.nchunk++;
            p.c0 = ((nbp / 4 + p.nchunk - 1) / p.nchunk) * 4;
        }
        p.nchunk = (nbp + p.c0 - 1) / p.c0;
        // Q8_0 rows longer than two chunks: K1c's DP4A on whole bytes is the faster chain there (7B layer graph, same box:
        // 54.9 us per layer on K1c, 58.4 on K1m with five chunks per strip; profiles/r02m_layer_types.txt)
        if (type == T_Q8_0 && p.nchunk > 2) return cudaErrorNotSupported;
        const int cb = p.c0 * 2 * bb;                                        // bytes of the larger chunk of a row
        p.stride = cb + ((cb % 32) == 16 ? 0 : 16);                          // stride / 4 = 4 mod 8
        if (p.sr * p.stride > 56 * 1024) return cudaErrorNotSupported;
        p.rows_per_cta = (int)(M / grid); p.rows_rem = (int)(M % grid);
        p.rows_max = p.rows_per_cta + (p.rows_rem ? 1 : 0);
        part_rows = std::max(part_rows, p.nseg * p.rows_max);
        nblk_max = std::max(nblk_max, p.nblk);
    }
    if ((size_t)part_rows * MC_PS * 4 > 24 * 1024) return cudaErrorNotSupported;
    // Hand-offs.  Phase i > 0 may take its activations from phase i - 1 (by mailbox) or from outside the launch; nothing else
    // the launch writes may be read by another CTA of the same launch, because nothing but the mailbox orders CTAs:
    //   * every output range (dst, extra dst, bare product, silu, silu * up, the stored normalised activations) is written by
    //     exactly one phase;
    //   * activations, gains and residuals lie outside every output range, except: activations that ARE the previous phase's
    //     output (-> mailbox), and a residual that is exactly an earlier phase's dst with the same row count (every CTA then
    //     reads only rows it wrote itself).
    struct Range { uintptr_t lo, hi; int phase; };
    std::vector<Range> outs;
    auto add = [&](const void *q, size_t floats, int ph) { if (q) outs.push_back({(uintptr_t)q, (uintptr_t)q + floats * 4, ph}); };
    for (int i = 0; i < n; i++) {
        const MPhase &p = cp.ph[i];
        for (int sg = 0; sg < p.nseg; sg++) add(p.dst[sg], (size_t)p.M, i);
        add(p.dst_raw, (size_t)p.M, i); add(p.epi_act, (size_t)p.M, i); add(p.epi_out, (size_t)p.M, i); add(p.b_out, (size_t)p.K, i);
    }
    for (size_t x = 0; x < outs.size(); x++)
        for (size_t y = x + 1; y < outs.size(); y++)
            if (outs[x].lo < outs[y].hi && outs[y].lo < outs[x].hi) return cudaErrorNotSupported;
    auto written = [&](const void *q, size_t floats) {
        if (!q) return false;
        const uintptr_t lo = (uintptr_t)q, hi = lo + floats * 4;
        for (const Range &r : outs) if (lo < r.hi && r.lo < hi) return true;
        return false;
    };
    for (int i = 0; i < n; i++) {
        MPhase &p = cp.ph[i];
        if (written(p.pro_w, (size_t)p.K)) return cudaErrorNotSupported;
        if (p.residual && written(p.residual, (size_t)p.M)) {
            bool own = false;
            for (int j = 0; j < i; j++) own = own || (cp.ph[j].dst[0] == p.residual && cp.ph[j].M == p.M);
            if (!own) return cudaErrorNotSupported;
        }
        if (!written(p.b, (size_t)p.K)) continue;            // from outside the launch
        if (i == 0 || p.K > MC_KMAX) return cudaErrorNotSupported;
        MPhase &q = cp.ph[i - 1];
        if (q.M != p.K) return cudaErrorNotSupported;
        int feed = 0;
        for (int sg = 0; sg < q.nseg; sg++) if (q.dst[sg] == p.b) feed = 1 + sg;
        if (q.epi_mode == 1 && q.epi_out == p.b) feed = 4;
        if (!feed) return cudaErrorNotSupported;             // an older phase's output, or a side output: K1c takes the chain
        q.feed = feed;
        p.from_mbox = 1;
    }
    cp.mbox = (uint2 *)mbox;
    cp.inflight = tune.max_stages > 0 ? std::min(tune.max_stages, MC_NB - 1) : 2;
    cp.nphase = n;
    cp.wait_mode = tune.wait_mode == 0 ? 0 : 1;
    cp.bar = bar;
    cp.tl = (unsigned long long *)tune.timeline;
    cp.part_off = MC_HDR;
    cp.part_floats = part_rows * MC_PS;
    cp.x_off = (cp.part_off + 2 * part_rows * MC_PS * 4 + 15) & ~15;
    cp.sc_off = cp.x_off + nblk_max * MC_XB;
    cp.ring_off = (cp.sc_off + nblk_max * 16 + 127) & ~127;
    const int total = 227 * 1024;
    cp.ring_bytes = total - cp.ring_off - 128;     // slack: the realigning loads of a stage's last row read one word past it
    for (int i = 0; i < n; i++)
        if (2 * cp.ph[i].sr * cp.ph[i].stride > cp.ring_bytes) return cudaErrorNotSupported;
    const int smem = total;

    void (*kern)(const MChainParams) = type == T_Q4_0 ? gemv_mchain_kernel<T_Q4_0> : type == T_Q4_1 ? gemv_mchain_kernel<T_Q4_1> : gemv_mchain_kernel<T_Q8_0>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    int per_sm = 0;     // the grid barrier needs every CTA resident at once (see launch_gemv_chain)
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, (MC_W + MC_PW) * 32, (size_t)smem);
    if (e != cudaSuccess) return e;
    if (per_sm < 1 || (int64_t)per_sm * sm_count < grid) return cudaErrorNotSupported;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3((unsigned)((MC_W + MC_PW) * 32));
    cfg.dynamicSmemBytes = (size_t)smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = tune.pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, cp);
}

}  // namespace lk

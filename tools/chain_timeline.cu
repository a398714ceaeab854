// chain_timeline.cu — development harness: one chained launch (lk::launch_gemv_chain) of the four GEMV phases of two
// Llama-2-7B decoder layers ([RMSNorm -> Q/K/V], [O + residual], [RMSNorm -> gate/up + SwiGLU], [down + residual]) with
// per-CTA, per-phase clock stamps compiled in (-DLK_GEMV_TIMELINE) and a table of where every phase spends its time.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -DLK_GEMV_TIMELINE \
//        -I llama.kotlin_b200/csrc tools/chain_timeline.cu -o tools/chain_timeline
//   tools/chain_timeline [type 2|3|6] [layers 1|2] [copies] [E] [F] [mma 0|1] [nostamps 0|1] [pdl 0|1]   (mma 1: K1m, lk_gemv_mma.cu)
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "lk_gemv.cu"
#include "lk_gemv_mma.cu"

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

int main(int argc, char **argv) {
    const int type = argc > 1 ? atoi(argv[1]) : 2;
    const int layers = argc > 2 ? atoi(argv[2]) : 2;
    const int copies = argc > 3 ? atoi(argv[3]) : 16;     // distinct weight sets, so that every launch streams HBM
    const int64_t E = argc > 4 ? atoll(argv[4]) : 4096, F = argc > 5 ? atoll(argv[5]) : 11008;
    const int mma = argc > 6 ? atoi(argv[6]) : 0;
    const int bb = lk::block_bytes(type);
    const int grid = 148;
    auto wbytes = [&](int64_t M, int64_t K) { return ((size_t)M * K / 32 * bb + 255) & ~(size_t)255; };
    const size_t per_layer = 4 * wbytes(E, E) + 2 * wbytes(F, E) + wbytes(E, F);
    uint8_t *w;
    CK(cudaMalloc(&w, per_layer * layers * copies));
    {
        std::vector<uint8_t> h(per_layer);
        for (auto &v : h) v = (uint8_t)(rand() & 0xFF);
        for (size_t b = 0; b < per_layer / bb; b++) { h[b * bb + 1] = 0x1E; if (type == 3) { h[b * bb + 2] = 0; h[b * bb + 3] = 0xA9; } }
        for (int c = 0; c < layers * copies; c++) CK(cudaMemcpy(w + c * per_layer, h.data(), per_layer, cudaMemcpyHostToDevice));
    }
    float *act, *gain; unsigned *bar; unsigned long long *tl;
    const size_t act_floats = 32 * (size_t)F;
    CK(cudaMalloc(&act, act_floats * 4)); CK(cudaMalloc(&gain, E * 4));
    CK(cudaMalloc(&bar, 64)); CK(cudaMemset(bar, 0, 64));
    void *mbox; CK(cudaMalloc(&mbox, lk::GEMV_MCHAIN_MBOX_BYTES)); CK(cudaMemset(mbox, 0, lk::GEMV_MCHAIN_MBOX_BYTES));
    CK(cudaMalloc(&tl, (size_t)grid * LK_CTL_STRIDE * 8));
    {
        std::vector<float> hx(act_floats);
        for (auto &v : hx) v = (float)rand() / RAND_MAX - 0.5f;
        CK(cudaMemcpy(act, hx.data(), hx.size() * 4, cudaMemcpyHostToDevice));
        std::vector<float> g(E, 1.0f);
        CK(cudaMemcpy(gain, g.data(), E * 4, cudaMemcpyHostToDevice));
    }
    // activation slots (F floats each)
    auto slot = [&](int i) { return act + (size_t)i * F; };
    cudaStream_t s; CK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    lk::GemvTuning tune;
    tune.timeline = (argc > 7 && atoi(argv[7])) ? nullptr : tl;     // 8th argument 1: no stamps (the stamps cost time themselves)
    tune.pdl = argc > 8 ? atoi(argv[8]) : 0;
    tune.max_stages = argc > 9 ? atoi(argv[9]) : 0;      // K1m: stages in flight
    auto run = [&](int c, int use_mma) {
        lk::GemvChainLink L[8];
        int n = 0;
        for (int l = 0; l < layers; l++) {
            uint8_t *base = w + ((size_t)c * layers + l) * per_layer;
            uint8_t *wq = base, *wk = wq + wbytes(E, E), *wv = wk + wbytes(E, E), *wo = wv + wbytes(E, E);
            uint8_t *wg = wo + wbytes(E, E), *wu = wg + wbytes(F, E), *wd = wu + wbytes(F, E);
            const int o = 16 * l;     // every layer writes its own slots (K1m declines chains whose phases share an output)
            float *x = slot(l ? 9 : 0), *q = slot(o + 1), *k = slot(o + 2), *v = slot(o + 3), *xn = slot(o + 4), *x2 = slot(o + 5), *xn2 = slot(o + 6), *g = slot(o + 7), *u = slot(o + 8), *h = slot(o + 10), *out = slot(o + 9);
            // [RMSNorm -> Q/K/V]
            L[n] = lk::GemvChainLink(); L[n].a = wq; L[n].K = E; L[n].M = E; L[n].b = x; L[n].dst = q;
            L[n].fu.pro_mode = 1; L[n].fu.pro_w = gain; L[n].fu.pro_eps = 1e-5f; L[n].fu.b_out = xn;
            L[n].fu.nseg_extra = 2; L[n].fu.a_extra[0] = wk; L[n].fu.a_extra[1] = wv; L[n].fu.dst_extra[0] = k; L[n].fu.dst_extra[1] = v; n++;
            // [O + residual]  (the attention core is left out: v stands in for its output)
            L[n] = lk::GemvChainLink(); L[n].a = wo; L[n].K = E; L[n].M = E; L[n].b = v; L[n].dst = x2; L[n].fu.residual = x; n++;
            // [RMSNorm -> gate/up + SwiGLU]
            L[n] = lk::GemvChainLink(); L[n].a = wg; L[n].K = E; L[n].M = F; L[n].b = x2; L[n].dst = g;
            L[n].fu.pro_mode = 1; L[n].fu.pro_w = gain; L[n].fu.pro_eps = 1e-5f; L[n].fu.b_out = xn2;
            L[n].fu.nseg_extra = 1; L[n].fu.a_extra[0] = wu; L[n].fu.dst_extra[0] = u; L[n].fu.epi_mode = 1; L[n].fu.epi_out = h; n++;
            // [down + residual]
            L[n] = lk::GemvChainLink(); L[n].a = wd; L[n].K = F; L[n].M = E; L[n].b = h; L[n].dst = out; L[n].fu.residual = x2; n++;
        }
        if (use_mma) CK(lk::launch_gemv_mchain(type, n, L, tune, grid, s, bar, mbox));
        else CK(lk::launch_gemv_chain(type, n, L, tune, grid, s, bar));
        return n;
    };
    int nph = 0;
    if (mma) {   // K1m against K1c on the same inputs: every activation slot the chain writes
        std::vector<float> r0(act_floats), r1(act_floats), init(act_floats);
        CK(cudaMemcpy(init.data(), act, act_floats * 4, cudaMemcpyDeviceToHost));
        run(0, 0); CK(cudaStreamSynchronize(s));
        CK(cudaMemcpy(r0.data(), act, act_floats * 4, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(act, init.data(), act_floats * 4, cudaMemcpyHostToDevice));
        run(0, 1); CK(cudaStreamSynchronize(s));
        CK(cudaMemcpy(r1.data(), act, act_floats * 4, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(act, init.data(), act_floats * 4, cudaMemcpyHostToDevice));
        for (int sl = 0; sl < 27; sl++) {
            double mx = 0, md = 0; int bad = 0;
            for (int64_t i = 0; i < F; i++) { const double a = r0[sl * F + i], b = r1[sl * F + i]; mx = std::max(mx, fabs(a)); md = std::max(md, fabs(a - b)); if (a != a || b != b) bad++; }
            printf("slot %2d: max |K1c| %.4g, max |K1m - K1c| %.4g (rel %.3g), NaNs %d\n", sl, mx, md, mx > 0 ? md / mx : 0.0, bad);
        }
    }
    for (int r = 0; r < 2; r++) for (int c = 0; c < copies; c++) nph = run(c, mma);
    CK(cudaStreamSynchronize(s));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    CK(cudaEventRecord(e0, s));
    for (int c = 0; c < copies; c++) run(c, mma);
    CK(cudaEventRecord(e1, s));
    CK(cudaStreamSynchronize(s));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    printf("type %d, %d layer(s) per launch (%d phases), %d launches: %.2f us per launch = %.2f us per layer (stream bound %.2f us per layer at 6535 GB/s)\n",
           type, layers, nph, copies, ms * 1e3 / copies, ms * 1e3 / copies / layers, per_layer / 6535e3);
    // the last launch's stamps
    std::vector<unsigned long long> h((size_t)grid * LK_CTL_STRIDE);
    CK(cudaMemcpy(h.data(), tl, h.size() * 8, cudaMemcpyDeviceToHost));
    unsigned bt[4]; CK(cudaMemcpy(bt, bar, 16, cudaMemcpyDeviceToHost));
    printf("barrier time-outs: %u\n", bt[2]);
    if (!tune.timeline) return 0;
    // SM clock from the (globaltimer, clock64) pairs; all stamps as microseconds since the first CTA's entry
    unsigned long long g0 = ~0ull;
    for (int b = 0; b < grid; b++) g0 = std::min(g0, h[(size_t)b * LK_CTL_STRIDE]);
    double mhz = 0; int nm = 0;
    for (int b = 0; b < grid; b++) {
        const unsigned long long *q = h.data() + (size_t)b * LK_CTL_STRIDE;
        const double dg = (double)(q[3] - q[0]), dc = (double)(q[8 + (nph - 1) * 8 + 6] - q[1]);
        if (dg > 0 && dc > 0) { mhz += dc / dg * 1e3; nm++; }
    }
    mhz /= std::max(nm, 1);
    printf("SM clock ~%.0f MHz\n", mhz);
    auto us = [&](int b, int ph, int sl) {
        const unsigned long long *q = h.data() + (size_t)b * LK_CTL_STRIDE;
        const unsigned long long v = q[8 + ph * 8 + sl];
        if (!v) return -1.0;
        return (double)(q[0] - g0) / 1e3 + (double)(v - q[1]) / mhz;
    };
    const char *names[8] = {"phase_start", "bar_arrive|x_loaded", "bar_passed|closing_barrier_passed", "x_prepped", "first_stage_full", "loop_done", "finished", "prod_last_issue"};
    for (int ph = 0; ph < nph; ph++) {
        printf("phase %d:\n", ph);
        for (int sl = 0; sl < 8; sl++) {
            std::vector<double> v;
            for (int b = 0; b < grid; b++) { const double t = us(b, ph, sl); if (t >= 0) v.push_back(t); }
            if (v.empty()) continue;
            std::sort(v.begin(), v.end());
            printf("   %-18s min %7.2f  med %7.2f  max %7.2f us\n", names[sl], v.front(), v[v.size() / 2], v.back());
        }
    }
    if (getenv("LK_TL_DUMP")) {     // per-CTA rows of one phase: smid, rows, stamps
        const int ph = atoi(getenv("LK_TL_DUMP"));
        printf("phase %d per CTA: bid smid start x_loaded x_prepped first_stage loop_done finished prod_last\n", ph);
        for (int b = 0; b < grid; b++) {
            const unsigned long long *q = h.data() + (size_t)b * LK_CTL_STRIDE;
            printf("%3d %3llu %7.2f %7.2f %7.2f %7.2f %7.2f %7.2f %7.2f\n", b, q[2], us(b, ph, 0), us(b, ph, 1), us(b, ph, 3), us(b, ph, 4), us(b, ph, 5), us(b, ph, 6), us(b, ph, 7));
        }
    }
    return 0;
}

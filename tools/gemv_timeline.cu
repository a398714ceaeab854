// gemv_timeline.cu — development harness: replays a PDL-chained CUDA graph of decode GEMVs
// (each on its own weight copy) with per-CTA %globaltimer stamps compiled in
// (-DLK_GEMV_TIMELINE) and prints where each launch spends its time.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -DLK_GEMV_TIMELINE \
//        -I llama.kotlin_b200/csrc tools/gemv_timeline.cu -o tools/gemv_timeline
//   tools/gemv_timeline <type 2|3|6> <M> <K> [copies] [pdl 0|1] [warps] [stage_bytes] [budget] [wait_mode 0|1|2]
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "lk_gemv.cu"

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

int main(int argc, char **argv) {
    const int type = argc > 1 ? atoi(argv[1]) : 2;
    const int64_t M = argc > 2 ? atoll(argv[2]) : 4096, K = argc > 3 ? atoll(argv[3]) : 4096;
    const int copies = argc > 4 ? atoi(argv[4]) : 24;
    lk::GemvTuning tune;
    tune.pdl = argc > 5 ? atoi(argv[5]) : 1;
    tune.variant = argc > 6 ? atoi(argv[6]) : 0;
    tune.stage_bytes = argc > 7 ? atoi(argv[7]) : 0;
    tune.smem_budget = argc > 8 ? atoi(argv[8]) : 0;
    tune.wait_mode = argc > 9 ? atoi(argv[9]) : 1;
    const int bb = lk::block_bytes(type);
    const size_t wb = (size_t)M * K / 32 * bb, wstride = (wb + 255) & ~(size_t)255;
    uint8_t *w; float *x, *dst; unsigned long long *tl;
    CK(cudaMalloc(&w, wstride * copies));
    CK(cudaMalloc(&x, K * 4)); CK(cudaMalloc(&dst, M * 4 * copies));
    const int grid = 148;
    CK(cudaMalloc(&tl, (size_t)copies * grid * 32 * 8));
    {   // plausible bytes: scale 0x1E00.. (~6e-3), random quants
        std::vector<uint8_t> h(wb);
        for (size_t i = 0; i < wb; i++) h[i] = (uint8_t)(rand() & 0xFF);
        for (size_t b = 0; b < wb / bb; b++) { h[b * bb] = (uint8_t)(rand() & 0xFF); h[b * bb + 1] = 0x1E; if (type == 3) { h[b * bb + 2] = 0; h[b * bb + 3] = 0xA9; } }
        for (int c = 0; c < copies; c++) CK(cudaMemcpy(w + c * wstride, h.data(), wb, cudaMemcpyHostToDevice));
        std::vector<float> hx(K);
        for (auto &v : hx) v = (float)rand() / RAND_MAX - 0.5f;
        CK(cudaMemcpy(x, hx.data(), K * 4, cudaMemcpyHostToDevice));
    }
    cudaStream_t s; CK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    lk::GemvPeers none;
    CK(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
    for (int c = 0; c < copies; c++) {
        tune.timeline = tl + (size_t)c * grid * 32;
        CK(lk::launch_gemv(type, w + c * wstride, K, M, x, dst + (size_t)c * M, 0, none, tune, grid, s));
    }
    cudaGraph_t g; CK(cudaStreamEndCapture(s, &g));
    cudaGraphExec_t ex; CK(cudaGraphInstantiate(&ex, g, 0));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    for (int r = 0; r < 3; r++) CK(cudaGraphLaunch(ex, s));
    CK(cudaStreamSynchronize(s));
    CK(cudaMemset(tl, 0, (size_t)copies * grid * 32 * 8));
    CK(cudaEventRecord(e0, s));
    CK(cudaGraphLaunch(ex, s));
    CK(cudaEventRecord(e1, s));
    CK(cudaStreamSynchronize(s));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    std::vector<unsigned long long> h((size_t)copies * grid * 32);
    CK(cudaMemcpy(h.data(), tl, h.size() * 8, cudaMemcpyDeviceToHost));
    printf("type %d M %lld K %lld copies %d pdl %d wait_mode %d: %.3f us/launch, %.1f GB/s\n", type, (long long)M, (long long)K, copies,
           tune.pdl, tune.wait_mode, ms * 1e3 / copies, (wb + 4.0 * K + 4.0 * M) / (ms * 1e-3 / copies) / 1e9);
    const char *ext[8] = {"final_bar_w0", "-", "prep_loaded", "prep_converted", "prep_stored", "pad_zeroed", "-", "-"};
    const char *names[8] = {"entry", "prod_start", "prefetched", "pdl_released", "x_prepped", "x_in_regs", "loop_done", "stored"};
    // per CTA: slot 0/29 = globaltimer entry/exit, slot 31 = clock64 at entry, others clock64
    unsigned long long t0 = ~0ull;
    for (int b = 0; b < grid; b++) t0 = std::min(t0, h[b * 32]);
    unsigned long long prev_start = 0, prev_end = 0;
    for (int c = 0; c < copies; c++) {
        const unsigned long long *k = h.data() + (size_t)c * grid * 32;
        unsigned long long kstart = ~0ull, kend = 0;
        for (int b = 0; b < grid; b++) { if (k[b * 32]) kstart = std::min(kstart, k[b * 32]); kend = std::max(kend, k[b * 32 + 29]); }
        if (c < 2 || c >= copies - 1 || c == copies / 2) {
            double mhz = 0; int nm = 0;
            for (int b = 0; b < grid; b++) { const double dg = (double)(k[b * 32 + 29] - k[b * 32]), dc = (double)(k[b * 32 + 7] - k[b * 32 + 31]); if (dg > 0) { mhz += dc / dg * 1e3; nm++; } }
            printf("launch %2d: first entry at %7.2f us (prev start +%.2f, prev end %+.2f), last exit at %7.2f; dur %.2f us; SM clock ~%.0f MHz\n", c,
                   (kstart - t0) / 1e3, c ? (kstart - prev_start) / 1e3 : 0.0, c ? ((double)kstart - (double)prev_end) / 1e3 : 0.0,
                   (kend - t0) / 1e3, (kend - kstart) / 1e3, mhz / std::max(nm, 1));
            {
                std::vector<double> v;
                for (int b = 0; b < grid; b++) v.push_back((double)(k[b * 32] - kstart) / 1e3);
                std::sort(v.begin(), v.end());
                printf("   %-14s min %6.2f  med %6.2f  max %6.2f us after the first CTA\n", "cta_entry", v.front(), v[v.size() / 2], v.back());
            }
            for (int sl = 1; sl < 28; sl++) {
                std::vector<double> v;   // cycles since this CTA's entry
                for (int b = 0; b < grid; b++) if (k[b * 32 + sl]) v.push_back((double)(k[b * 32 + sl] - k[b * 32 + 31]));
                if (v.empty()) continue;
                std::sort(v.begin(), v.end());
                char nm2[32];
                if (sl < 8) snprintf(nm2, sizeof nm2, "%s", names[sl]); else if (sl >= 20) snprintf(nm2, sizeof nm2, "%s", ext[sl - 20]); else snprintf(nm2, sizeof nm2, "it%d_%s", (sl - 8) / 2, (sl & 1) ? "computed" : "full");
                printf("   %-14s min %7.0f  med %7.0f  max %7.0f cycles since CTA entry\n", nm2, v.front(), v[v.size() / 2], v.back());
            }
        }
        prev_start = kstart; prev_end = kend;
    }
    return 0;
}

// overlap_timeline.cu — evidence for the headline's launch overlap (VERDICT r1 "prove the overlap"): replays the bench
// step's launch pattern — independent Q4_0 decode GEMVs at the three Llama-2-7B shapes, each on its own weight copy, x
// and dst, launched with programmatic dependent launch in END mode — with per-CTA %globaltimer entry / exit stamps and
// %smid compiled in (-DLK_GEMV_TIMELINE), and prints per launch its lifetime and distance to the previous launch, how
// many launches are in flight on average, and per SM how often CTAs of consecutive launches were resident together.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -DLK_GEMV_TIMELINE -I llama.kotlin_b200/csrc \
//        tools/overlap_timeline.cu -o tools/overlap_timeline && tools/overlap_timeline [launches=24] [wait_mode=2]
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "lk_gemv.cu"

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

int main(int argc, char **argv) {
    const int n = argc > 1 ? atoi(argv[1]) : 24;
    const int wait_mode = argc > 2 ? atoi(argv[2]) : 2;
    const int64_t shapes[3][2] = {{4096, 4096}, {11008, 4096}, {4096, 11008}};
    const int grid = 148, bb = 18;
    std::vector<uint8_t *> w(n);
    std::vector<float *> x(n), d(n);
    unsigned long long *tl;
    CK(cudaMalloc(&tl, (size_t)n * grid * 32 * 8));
    srand(7);
    for (int i = 0; i < n; i++) {
        const int64_t M = shapes[i % 3][0], K = shapes[i % 3][1];
        const size_t wb = (size_t)M * K / 32 * bb;
        CK(cudaMalloc(&w[i], wb)); CK(cudaMalloc(&x[i], K * 4)); CK(cudaMalloc(&d[i], M * 4));
        std::vector<uint8_t> h(wb);
        for (size_t k = 0; k < wb; k++) h[k] = (uint8_t)(rand() & 0xFF);
        for (size_t b = 0; b < wb / bb; b++) h[b * bb + 1] = 0x1E;          // scale ~6e-3
        CK(cudaMemcpy(w[i], h.data(), wb, cudaMemcpyHostToDevice));
        std::vector<float> hx(K);
        for (auto &v : hx) v = (float)rand() / RAND_MAX - 0.5f;
        CK(cudaMemcpy(x[i], hx.data(), K * 4, cudaMemcpyHostToDevice));
    }
    cudaStream_t s; CK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    lk::GemvPeers none;
    lk::GemvTuning tune;
    tune.pdl = 1;
    tune.wait_mode = wait_mode;
    CK(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
    for (int i = 0; i < n; i++) {
        tune.timeline = tl + (size_t)i * grid * 32;
        CK(lk::launch_gemv(2, w[i], shapes[i % 3][1], shapes[i % 3][0], x[i], d[i], 0, none, tune, grid, s));
    }
    cudaGraph_t g; CK(cudaStreamEndCapture(s, &g));
    cudaGraphExec_t ex; CK(cudaGraphInstantiate(&ex, g, 0));
    for (int r = 0; r < 3; r++) CK(cudaGraphLaunch(ex, s));
    CK(cudaStreamSynchronize(s));
    CK(cudaMemset(tl, 0, (size_t)n * grid * 32 * 8));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    CK(cudaEventRecord(e0, s));
    CK(cudaGraphLaunch(ex, s));
    CK(cudaEventRecord(e1, s));
    CK(cudaStreamSynchronize(s));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    std::vector<unsigned long long> h((size_t)n * grid * 32);
    CK(cudaMemcpy(h.data(), tl, h.size() * 8, cudaMemcpyDeviceToHost));
    double bytes = 0;
    for (int i = 0; i < n; i++) bytes += (double)shapes[i % 3][0] * shapes[i % 3][1] / 32 * bb + 4.0 * shapes[i % 3][1] + 4.0 * shapes[i % 3][0];
    printf("%d independent Q4_0 GEMV launches (4096x4096, 11008x4096, 4096x11008 in turn), PDL, wait_mode %d: %.2f us in all, %.2f us / launch, %.0f GB/s\n",
           n, wait_mode, ms * 1e3, ms * 1e3 / n, bytes / (ms * 1e-3) / 1e9);
    unsigned long long t0 = ~0ull;
    for (size_t q = 0; q < (size_t)n * grid; q++) if (h[q * 32]) t0 = std::min(t0, h[q * 32]);
    std::vector<double> st(n), en(n);
    for (int i = 0; i < n; i++) {
        unsigned long long a = ~0ull, z = 0;
        for (int b = 0; b < grid; b++) { const unsigned long long *k = &h[((size_t)i * grid + b) * 32]; if (k[0]) a = std::min(a, k[0]); z = std::max(z, k[29]); }
        st[i] = (a - t0) / 1e3; en[i] = (z - t0) / 1e3;
        printf("launch %2d (%5lld x %5lld): first CTA in at %7.2f us, last CTA out at %7.2f us, alive %5.2f us, start %+5.2f us after the previous launch's start, %+6.2f us relative to its end\n",
               i, (long long)shapes[i % 3][0], (long long)shapes[i % 3][1], st[i], en[i], en[i] - st[i], i ? st[i] - st[i - 1] : 0.0, i ? st[i] - en[i - 1] : 0.0);
    }
    double area = 0;
    for (int i = 0; i < n; i++) area += en[i] - st[i];
    printf("window %.2f us; sum of launch lifetimes %.2f us => %.2f launches in flight on average\n", en[n - 1] - st[0], area, area / (en[n - 1] - st[0]));
    int pairs = 0, co = 0;
    double overlap_us = 0;
    for (int i = 1; i < n; i++)
        for (int b = 0; b < grid; b++) {
            const unsigned long long *k = &h[((size_t)i * grid + b) * 32];
            for (int pb = 0; pb < grid; pb++) {       // the CTA of the previous launch that ran on the same SM
                const unsigned long long *q = &h[((size_t)(i - 1) * grid + pb) * 32];
                if ((unsigned)q[30] != (unsigned)k[30]) continue;
                pairs++;
                if (k[0] < q[29]) { co++; overlap_us += (double)(std::min(q[29], k[29]) - k[0]) / 1e3; }
            }
        }
    printf("CTA pairs (launch i, launch i-1) on the same SM: %d; the later CTA entered before the earlier one left in %d of them (%.1f %%), mean co-resident time %.2f us\n",
           pairs, co, 100.0 * co / std::max(pairs, 1), overlap_us / std::max(co, 1));
    return 0;
}

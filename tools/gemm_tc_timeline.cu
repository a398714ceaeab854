// gemm_tc_timeline.cu — development harness for the tcgen05 prefill GEMM (K2): runs launch_gemm_tc with
// per-CTA clock64 counters compiled in (-DLK_TC_TIMELINE) and prints where the roles spend their time.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -DLK_TC_TIMELINE \
//        -I llama.kotlin_b200/csrc tools/gemm_tc_timeline.cu -o tools/gemm_tc_timeline
//   tools/gemm_tc_timeline <type 2|3|6> <M> <K> <N> [cluster]
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "lk_gemm_tc.cu"

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

int main(int argc, char **argv) {
    const int type = argc > 1 ? atoi(argv[1]) : 6;
    const int64_t M = argc > 2 ? atoll(argv[2]) : 11008, K = argc > 3 ? atoll(argv[3]) : 4096, N = argc > 4 ? atoll(argv[4]) : 512;
    const int cluster = argc > 5 ? atoi(argv[5]) : 2;
    const int bb = lk::block_bytes(type);
    const size_t wb = (size_t)M * K / 32 * bb;
    uint8_t *w; float *x, *dst; void *ws; unsigned long long *dbg;
    CK(cudaMalloc(&w, wb)); CK(cudaMalloc(&x, (size_t)K * N * 4)); CK(cudaMalloc(&dst, (size_t)M * N * 4));
    CK(cudaMalloc(&ws, lk::gemm_tc_workspace_bytes(K, N) + 1024));
    const int max_ctas = 4096;
    CK(cudaMalloc(&dbg, (size_t)max_ctas * 16 * 8));
    {
        std::vector<uint8_t> h(wb);
        for (size_t i = 0; i < wb; i++) h[i] = (uint8_t)(rand() & 0xFF);
        for (size_t b = 0; b < wb / bb; b++) { h[b * bb] = (uint8_t)(rand() & 0xFF); h[b * bb + 1] = 0x1E; if (type == 3) { h[b * bb + 2] = 0; h[b * bb + 3] = 0xA9; } }
        CK(cudaMemcpy(w, h.data(), wb, cudaMemcpyHostToDevice));
        std::vector<float> hx((size_t)K * N);
        for (auto &v : hx) v = (float)rand() / RAND_MAX - 0.5f;
        CK(cudaMemcpy(x, hx.data(), hx.size() * 4, cudaMemcpyHostToDevice));
    }
    cudaStream_t s; CK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    lk::g_gemm_tc_debug = dbg;
    for (int r = 0; r < 3; r++) CK(lk::launch_gemm_tc(type, w, K, M, x, N, dst, ws, 148, cluster, s));
    CK(cudaStreamSynchronize(s));
    CK(cudaMemsetAsync(dbg, 0, (size_t)max_ctas * 16 * 8, s));
    CK(cudaEventRecord(e0, s));
    CK(lk::launch_gemm_tc(type, w, K, M, x, N, dst, ws, 148, cluster, s));
    CK(cudaEventRecord(e1, s));
    CK(cudaStreamSynchronize(s));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    std::vector<unsigned long long> h((size_t)max_ctas * 16);
    CK(cudaMemcpy(h.data(), dbg, h.size() * 8, cudaMemcpyDeviceToHost));
    int n = 0;
    while (n < max_ctas && h[(size_t)n * 16 + 12]) n++;
    printf("type %d M %lld K %lld N %lld cluster %d: %.2f us (pack + gemm), %.1f TFLOP/s, %d CTAs, %lld K-steps\n", type, (long long)M, (long long)K,
           (long long)N, cluster, ms * 1e3, 2.0 * M * K * N / (ms * 1e-3) / 1e12, n, (long long)K / 64);
    const char *names[16] = {"", "mma: wait full (sum)", "B producer: wait b_empty (sum)", "raw producer: wait raw_empty (sum)",
                             "dequant warp: wait raw_full (sum)", "dequant warp: wait a_empty (sum)", "", "mma: loop entry", "mma: first full",
                             "dequant warp: mainloop done", "dequant warp: acc_full seen", "dequant warp: epilogue done", "cta: end", "", "", ""};
    for (int slot : {7, 8, 1, 2, 3, 4, 5, 9, 10, 11, 12}) {
        std::vector<unsigned long long> v;
        for (int i = 0; i < n; i++) v.push_back(h[(size_t)i * 16 + slot]);
        std::sort(v.begin(), v.end());
        double mean = 0;
        for (auto u : v) mean += (double)u;
        mean /= std::max(n, 1);
        printf("  %-36s min %8llu  median %8llu  max %8llu  mean %9.0f clk\n", names[slot], v.empty() ? 0 : v.front(), v.empty() ? 0 : v[v.size() / 2],
               v.empty() ? 0 : v.back(), mean);
    }
    return 0;
}

// Standalone numerics + timing harness for dmlb_gemm_bf16 (no torch): every case is checked against a
// double-precision CPU reference built from the same bf16-rounded inputs.
//   build:  python build_native.py --harness        run: tests/native/gemm_harness [--quick]
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "../../include/dmlb.h"

#define CK(x)                                                                          \
    do {                                                                               \
        cudaError_t e = (x);                                                           \
        if (e != cudaSuccess) {                                                        \
            printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); \
            exit(2);                                                                   \
        }                                                                              \
    } while (0)

static uint32_t rng_state = 12345u;
static float frand() {  // small multiples of 1/8 in [-1, 1]: exactly representable in bf16
    rng_state = rng_state * 1664525u + 1013904223u;
    return (float)((int)((rng_state >> 16) % 17) - 8) / 8.0f;
}

struct Case {
    int M, N, K, a_mn, b_mn, split, out_f32, bias, relu, tile_n, perm_cin, perm_khw;
};

static int run_case(const Case& c, bool timing, bool check = true) {
    const int M = c.M, N = c.N, K = c.K;
    auto pad8 = [](int x) { return (x + 7) / 8 * 8; };
    // storage shapes
    const int a_rows = c.a_mn ? K : M, a_cols = c.a_mn ? M : K, lda = pad8(a_cols);
    const int b_rows = c.b_mn ? K : N, b_cols = c.b_mn ? N : K, ldb = pad8(b_cols);
    const int ldo = c.out_f32 ? (N + 3) / 4 * 4 : pad8(N);
    std::vector<float> A((size_t)M * K), B((size_t)N * K), bias(N);
    for (auto& x : A) x = frand();
    for (auto& x : B) x = frand();
    for (auto& x : bias) x = frand();
    std::vector<__nv_bfloat16> hA((size_t)a_rows * lda, __float2bfloat16(7.0f)), hB((size_t)b_rows * ldb, __float2bfloat16(7.0f));
    for (int m = 0; m < M; ++m)
        for (int k = 0; k < K; ++k) {
            size_t idx = c.a_mn ? (size_t)k * lda + m : (size_t)m * lda + k;
            hA[idx] = __float2bfloat16(A[(size_t)m * K + k]);
        }
    for (int n = 0; n < N; ++n)
        for (int k = 0; k < K; ++k) {
            size_t idx = c.b_mn ? (size_t)k * ldb + n : (size_t)n * ldb + k;
            hB[idx] = __float2bfloat16(B[(size_t)n * K + k]);
        }
    __nv_bfloat16 *dA, *dB;
    float* dbias;
    void* dout;
    const size_t out_bytes = (size_t)M * ldo * (c.out_f32 ? 4 : 2);
    CK(cudaMalloc(&dA, hA.size() * 2));
    CK(cudaMalloc(&dB, hB.size() * 2));
    CK(cudaMalloc(&dbias, N * 4));
    CK(cudaMalloc(&dout, out_bytes));
    CK(cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dbias, bias.data(), N * 4, cudaMemcpyHostToDevice));
    CK(cudaMemset(dout, 0, out_bytes));
    const int atomic = c.split > 1;
    int rc = dmlb_gemm_bf16(dA, lda, c.a_mn, dB, ldb, c.b_mn, M, N, K, dout, c.out_f32, ldo, c.bias ? dbias : nullptr, c.relu, c.split,
                            atomic, c.perm_cin, c.perm_khw, c.tile_n, nullptr);
    if (rc != 0) {
        printf("FAIL launch rc=%d: %s\n", rc, dmlb_last_error());
        return 1;
    }
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
        printf("FAIL sync: %s\n", cudaGetErrorString(e));
        exit(3);  // context is dead after a trap
    }
    std::vector<uint8_t> hout(out_bytes);
    CK(cudaMemcpy(hout.data(), dout, out_bytes, cudaMemcpyDeviceToHost));
    double max_err = 0, max_ref = 0;
    int bad = 0;
    for (int m = 0; check && m < M; ++m)
        for (int n = 0; n < N; ++n) {
            double acc = 0;
            for (int k = 0; k < K; ++k) acc += (double)A[(size_t)m * K + k] * (double)B[(size_t)n * K + k];
            if (c.bias) acc += bias[n];
            if (c.relu && acc < 0) acc = 0;
            int col = n;
            if (c.perm_cin > 0) col = (n % c.perm_cin) * c.perm_khw + n / c.perm_cin;
            double got = c.out_f32 ? ((float*)hout.data())[(size_t)m * ldo + col]
                                   : (double)__bfloat162float(((__nv_bfloat16*)hout.data())[(size_t)m * ldo + col]);
            double err = fabs(got - acc);
            double tol = c.out_f32 ? 1e-3 + 1e-5 * fabs(acc) : 1e-2 + 8e-3 * fabs(acc);
            if (err > tol) {
                if (bad < 4) printf("   mismatch m=%d n=%d got=%g ref=%g\n", m, n, got, acc);
                ++bad;
            }
            if (err > max_err) max_err = err;
            if (fabs(acc) > max_ref) max_ref = fabs(acc);
        }
    double ms = 0;
    if (timing && !bad) {
        cudaEvent_t e0, e1;
        CK(cudaEventCreate(&e0));
        CK(cudaEventCreate(&e1));
        for (int i = 0; i < 3; ++i)
            dmlb_gemm_bf16(dA, lda, c.a_mn, dB, ldb, c.b_mn, M, N, K, dout, c.out_f32, ldo, c.bias ? dbias : nullptr, c.relu, c.split,
                           atomic, c.perm_cin, c.perm_khw, c.tile_n, nullptr);
        CK(cudaEventRecord(e0));
        const int iters = 20;
        for (int i = 0; i < iters; ++i)
            dmlb_gemm_bf16(dA, lda, c.a_mn, dB, ldb, c.b_mn, M, N, K, dout, c.out_f32, ldo, c.bias ? dbias : nullptr, c.relu, c.split,
                           atomic, c.perm_cin, c.perm_khw, c.tile_n, nullptr);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float t;
        CK(cudaEventElapsedTime(&t, e0, e1));
        ms = t / iters;
    }
    printf("%s M=%d N=%d K=%d A:%s B:%s split=%d out=%s bias=%d relu=%d tile_n=%d perm=%d/%d  max_err=%.3g (max|ref|=%.3g) bad=%d", bad ? "FAIL" : "ok  ",
           M, N, K, c.a_mn ? "MN" : "K", c.b_mn ? "MN" : "K", c.split, c.out_f32 ? "f32" : "bf16", c.bias, c.relu, c.tile_n, c.perm_cin,
           c.perm_khw, max_err, max_ref, bad);
    if (ms > 0) printf("  %.2f us  %.1f TFLOP/s", ms * 1e3, 2.0 * M * N * K / (ms * 1e-3) / 1e12);
    printf("\n");
    fflush(stdout);
    cudaFree(dA);
    cudaFree(dB);
    cudaFree(dbias);
    cudaFree(dout);
    return bad ? 1 : 0;
}

// Per-kernel cost inside a CUDA graph (no host launch overhead): 64 dependent launches captured once, replayed 20 times.
static void graph_latency(const char* label, int M, int N, int K, int split, bool trivial) {
    auto pad8 = [](int x) { return (x + 7) / 8 * 8; };
    __nv_bfloat16 *dA, *dB;
    float* dout;
    unsigned long long* ctr;
    CK(cudaMalloc(&dA, (size_t)M * pad8(K) * 2));
    CK(cudaMalloc(&dB, (size_t)N * pad8(K) * 2));
    CK(cudaMalloc(&dout, (size_t)M * pad8(N) * 4));
    CK(cudaMalloc(&ctr, 8));
    CK(cudaMemset(dA, 0, (size_t)M * pad8(K) * 2));
    CK(cudaMemset(dB, 0, (size_t)N * pad8(K) * 2));
    CK(cudaMemset(dout, 0, (size_t)M * pad8(N) * 4));
    cudaStream_t st;
    CK(cudaStreamCreate(&st));
    auto one = [&]() {
        if (trivial) dmlb_counter_bump(ctr, st);
        else dmlb_gemm_bf16(dA, pad8(K), 0, dB, pad8(K), 0, M, N, K, dout, 1, pad8(N), nullptr, 0, split, split > 1, 0, 0, 0, st);
    };
    one();
    CK(cudaStreamSynchronize(st));
    cudaGraph_t g;
    cudaGraphExec_t ge;
    const int n = 64;
    CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
    for (int i = 0; i < n; ++i) one();
    CK(cudaStreamEndCapture(st, &g));
    CK(cudaGraphInstantiate(&ge, g, 0));
    for (int i = 0; i < 3; ++i) CK(cudaGraphLaunch(ge, st));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    CK(cudaEventRecord(e0, st));
    for (int i = 0; i < 20; ++i) CK(cudaGraphLaunch(ge, st));
    CK(cudaEventRecord(e1, st));
    CK(cudaEventSynchronize(e1));
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    printf("graph-latency %-34s %7.2f us per kernel (pdl=%d)\n", label, ms * 1e3 / (20.0 * n), dmlb_pdl_enabled());
    fflush(stdout);
    cudaGraphExecDestroy(ge); cudaGraphDestroy(g); cudaStreamDestroy(st);
    cudaFree(dA); cudaFree(dB); cudaFree(dout); cudaFree(ctr);
}

int main(int argc, char** argv) {
    if (argc > 1 && !strcmp(argv[1], "--latency")) {
        for (int pdl = 1; pdl >= 0; --pdl) {
            dmlb_set_pdl(pdl);
            graph_latency("1-thread kernel", 0, 0, 0, 1, true);
            graph_latency("gemm 64x10x256 (1 CTA)", 64, 10, 256, 1, false);
            graph_latency("gemm 256x384x1728 (6 CTAs)", 256, 384, 1728, 1, false);
            graph_latency("gemm 256x384x1728 split 13 (78 CTAs)", 256, 384, 1728, 13, false);
            graph_latency("gemm 4096x64x363 (32 CTAs)", 4096, 64, 363, 1, false);
        }
        return 0;
    }
    dmlb_set_gemm_persistent(1);   // exercise the opt-in persistent kernel on the multi-wave cases below
    bool quick = argc > 1 && !strcmp(argv[1], "--quick");
    bool big_only = argc > 1 && !strcmp(argv[1], "--big");   // one large GEMM, timing only: the command ncu wraps
    std::vector<Case> cases = {
        // M, N, K, a_mn, b_mn, split, out_f32, bias, relu, tile_n, perm_cin, perm_khw
        {128, 128, 64, 0, 0, 1, 1, 0, 0, 0, 0, 0},     // one tile, one k-block
        {128, 128, 256, 0, 0, 1, 1, 0, 0, 0, 0, 0},    // K loop, smem ring reuse
        {128, 128, 1024, 0, 0, 1, 1, 0, 0, 0, 0, 0},   // ring wraps several times
        {256, 384, 1728, 0, 0, 1, 0, 1, 1, 0, 0, 0},   // AlexNet conv3 fprop: bf16 out + bias + relu
        {4096, 64, 363, 0, 0, 1, 0, 1, 1, 0, 0, 0},    // AlexNet conv1 fprop (ragged K)
        {64, 10, 256, 0, 0, 1, 1, 1, 0, 0, 0, 0},      // classifier (tiny, ragged M/N)
        {100, 72, 200, 0, 0, 1, 1, 1, 0, 0, 0, 0},     // everything ragged
        {256, 256, 3456, 0, 0, 8, 1, 1, 0, 0, 0, 0},   // split-K atomics
        {128, 64, 64, 0, 0, 1, 1, 0, 0, 32, 0, 0},     // BN=32 tiles
        {128, 512, 128, 0, 0, 1, 1, 0, 0, 256, 0, 0},  // BN=256 tiles
        {128, 128, 128, 0, 1, 1, 1, 0, 0, 0, 0, 0},    // dgrad form: B MN-major
        {256, 1728, 384, 0, 1, 1, 0, 0, 0, 0, 0, 0},   // AlexNet conv3 dgrad
        {1024, 1600, 192, 0, 1, 1, 0, 0, 0, 64, 0, 0}, // conv2 dgrad, BN=64
        {128, 128, 128, 1, 1, 1, 1, 0, 0, 0, 0, 0},    // wgrad form: both MN-major
        {384, 1728, 256, 1, 1, 2, 1, 0, 0, 0, 0, 0},   // conv3 wgrad split-K
        {64, 363, 4096, 1, 1, 16, 1, 0, 0, 0, 3, 121}, // conv1 wgrad, split-K, (kh,kw,cin)->(cin,kh,kw) scatter
        {10, 256, 64, 1, 1, 1, 1, 0, 0, 0, 0, 0},      // classifier wgrad
        {8192, 1024, 96, 0, 0, 1, 0, 1, 1, 0, 0, 0},   // 512 tiles: persistent kernel, double-buffered TMEM, ragged K
        {8200, 1000, 200, 0, 1, 1, 0, 0, 0, 0, 0, 0},  // persistent, B MN-major, ragged M and N
        {2048, 4096, 100, 1, 1, 1, 1, 0, 0, 0, 0, 0},  // persistent, both MN-major, fp32 out
        {16384, 64, 64, 0, 0, 1, 0, 1, 0, 64, 0, 0},   // persistent, one k-block per tile (ResNet 1x1 conv shape)
    };
    std::vector<Case> perf = {
        {8192, 8192, 8192, 0, 0, 1, 0, 0, 0, 128, 0, 0},
        {8192, 8192, 8192, 0, 0, 1, 0, 0, 0, 256, 0, 0},
        {200704, 64, 576, 0, 0, 1, 0, 1, 1, 64, 0, 0},
    };
    int fails = 0;
    if (argc > 1 && !strcmp(argv[1], "--pair")) {
        // CTA-pair (cta_group::2) kernel: numerics on small / ragged / split-K problems, then 8192^3 with the pair kernel on and off
        std::vector<Case> pc = {
            {256, 256, 64, 0, 0, 1, 1, 0, 0, 256, 0, 0},     // one pair, one k-block
            {512, 512, 512, 0, 0, 1, 1, 0, 0, 256, 0, 0},    // 2 x 2 pairs, ring wraps
            {300, 700, 200, 0, 0, 1, 0, 1, 1, 256, 0, 0},    // ragged M / N / K, bias + ReLU, bf16 out
            {256, 256, 3456, 0, 0, 4, 1, 1, 0, 256, 0, 0},   // split-K atomics
            {1024, 1024, 4096, 0, 0, 1, 0, 0, 0, 256, 0, 0}, // long K: many ring revolutions
        };
        for (auto& c : pc) fails += run_case(c, true);
        Case big = {8192, 8192, 8192, 0, 0, 1, 0, 0, 0, 256, 0, 0};
        Case ragged = {2600, 3600, 100, 0, 0, 1, 1, 1, 0, 256, 0, 0};  // 165 tiles >= 2 per TPC: persistent pair kernel, ragged M / N / K
        Case deep = {2560, 4096, 192, 0, 0, 1, 0, 1, 1, 256, 0, 0};    // persistent, bf16 out + bias + ReLU, accumulator buffers alternate
        fails += run_case(ragged, true);
        fails += run_case(deep, true);
        printf("pair kernel on (persistent, double-buffered TMEM):\n");
        fails += run_case(big, true, false);
        dmlb_set_gemm_pair(2);
        printf("pair kernel on (one tile per cluster launch):\n");
        fails += run_case(big, true, false);
        dmlb_set_gemm_pair(0);
        printf("pair kernel off (one CTA per 128 x 256 tile):\n");
        fails += run_case(big, true, false);
        dmlb_set_gemm_pair(1);
        printf(fails ? "HARNESS FAILED (%d)\n" : "HARNESS PASSED\n", fails);
        return fails ? 1 : 0;
    }
    if (big_only) {
        Case c = {8192, 8192, 8192, 0, 0, 1, 0, 0, 0, 256, 0, 0};
        fails += run_case(c, true, false);
        printf(fails ? "HARNESS FAILED (%d)\n" : "HARNESS PASSED\n", fails);
        return fails ? 1 : 0;
    }
    for (auto& c : cases) fails += run_case(c, true);
    if (!quick) {
        // the perf cases are too large for the O(MNK) CPU check: verify a sub-problem via the same code path instead
        for (auto& c : perf) {
            Case small = c;
            small.M = 256; small.N = c.N > 256 ? 256 : c.N; small.K = c.K > 512 ? 512 : c.K;
            fails += run_case(small, false);
            fails += run_case(c, true, false);  // timing only
        }
    }
    printf(fails ? "HARNESS FAILED (%d)\n" : "HARNESS PASSED\n", fails);
    return fails ? 1 : 0;
}

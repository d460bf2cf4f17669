// bf16 entry points of the tcgen05 GEMM family (kind::f16): see gemm_tcgen05.cuh for the kernels and include/dmlb.h for the contracts.
// Reference work taken over: nn.Conv2d / nn.Linear forward and backward, example/models.py:8-13,28-43 (cuDNN / cuBLAS there).
#include "gemm_tcgen05.cuh"

using namespace dmlb;

// 1 when dmlb_*_fin can fold the 2x2/2 pool for an [H, W] map (otherwise run the pool as its own kernel).
extern "C" int dmlb_fin_pool_supported(int H, int W) { return (H > 0 && W > 0 && H % 2 == 0 && W % 2 == 0 && kBM % (2 * W) == 0) ? 1 : 0; }

extern "C" int dmlb_gemm_bf16(const void* A, int lda, int a_mn_major, const void* B, int ldb, int b_mn_major, int M, int N, int K, void* out,
                              int out_is_f32, int ldo, const float* bias, int relu, int split_k, int atomic, int perm_cin, int perm_khw,
                              int tile_n, void* stream) {
    return gemm_impl<ElemBF16>(nullptr, A, lda, a_mn_major, B, ldb, b_mn_major, M, N, K, out, out_is_f32, ldo, bias, relu, split_k, atomic,
                               perm_cin, perm_khw, tile_n, stream);
}

extern "C" int dmlb_conv_implicit_bf16(int kind, const void* act, int ld_act, int B, int Ha, int Wa, int Ca, const void* other, int ld_other,
                                       int KH, int KW, int SH, int SW, int PH, int PW, int n_out, void* out, int out_is_f32, int ldo,
                                       const float* bias, int relu, int split_k, int atomic, int tile_n, void* stream) {
    return conv_implicit_impl<ElemBF16>(nullptr, kind, act, ld_act, B, Ha, Wa, Ca, other, ld_other, KH, KW, SH, SW, PH, PW, n_out, out, out_is_f32,
                                        ldo, bias, relu, split_k, atomic, tile_n, stream);
}

// Split-K GEMM / implicit convolution with the finishing pass folded into the GEMM kernel (one launch instead of two):
//   y[M,N] (bf16, pitch ldy) = act(A * B^T + bias), optionally also pooled = maxpool2x2(y) for rows that enumerate a [.., pool_h, pool_w] map.
//   ws: fp32 [M, ldw] scratch and counters: int [n_counters >= output tiles] — both all zero on entry and all zero again on exit.
extern "C" int dmlb_gemm_bf16_fin(const void* A, int lda, int a_mn_major, const void* B, int ldb, int b_mn_major, int M, int N, int K, float* ws,
                                  int ldw, int* counters, int n_counters, const float* bias, int relu, void* y, int ldy, int pool_h, int pool_w,
                                  void* pooled, int ldp, int split_k, int tile_n, void* stream) {
    const FinishArgs fin = {counters, n_counters, bias, relu, y, ldy, pool_h, pool_w, pooled, ldp};
    return gemm_impl<ElemBF16>(&fin, A, lda, a_mn_major, B, ldb, b_mn_major, M, N, K, ws, 1, ldw, nullptr, 0, split_k, 1, 0, 0, tile_n, stream);
}

extern "C" int dmlb_conv_implicit_fin(int kind, const void* act, int ld_act, int B, int Ha, int Wa, int Ca, const void* other, int ld_other, int KH,
                                      int KW, int SH, int SW, int PH, int PW, int n_out, float* ws, int ldw, int* counters, int n_counters,
                                      const float* bias, int relu, void* y, int ldy, int pool_h, int pool_w, void* pooled, int ldp, int split_k,
                                      int tile_n, void* stream) {
    const FinishArgs fin = {counters, n_counters, bias, relu, y, ldy, pool_h, pool_w, pooled, ldp};
    return conv_implicit_impl<ElemBF16>(&fin, kind, act, ld_act, B, Ha, Wa, Ca, other, ld_other, KH, KW, SH, SW, PH, PW, n_out, ws, 1, ldw, nullptr,
                                        0, split_k, 1, tile_n, stream);
}

// Small-channel first-layer convolution (Cin <= 8): see conv_smallc_impl in gemm_tcgen05.cuh and include/dmlb.h.
extern "C" int dmlb_conv_smallc_bf16(int kind, const void* act, int B, int Ha, int Wa, int cin, const void* other, int ld_other, int KH, int KW,
                                     int SH, int SW, int PH, int PW, int n_out, void* out, int out_is_f32, int ldo, const float* bias, int relu,
                                     int split_k, int pt_layout, void* pooled, int ldp, int pool_w, int tile_n, int hp, int wp, void* stream) {
    return conv_smallc_impl<ElemBF16>(kind, act, B, Ha, Wa, cin, other, ld_other, KH, KW, SH, SW, PH, PW, n_out, out, out_is_f32, ldo, bias, relu,
                                 split_k, pt_layout, pooled, ldp, pool_w, tile_n, hp, wp, stream);
}

// Implicit convolution, cluster split-K form (kind 0 fprop / 2 dgrad): the K-slices of a tile reduce through distributed shared memory and
// the fused epilogue runs on the sum — see gemm_tcgen05.cuh (kCluster) and include/dmlb.h.
extern "C" int dmlb_conv_implicit_cluster_bf16(int kind, const void* act, int ld_act, int B, int Ha, int Wa, int Ca, const void* other, int ld_other,
                                               int KH, int KW, int SH, int SW, int PH, int PW, int n_out, void* out, int out_is_f32, int ldo,
                                               const float* bias, int relu, int cluster_s, int tile_n, void* pooled, int ldp, int pool_w,
                                               const void* act_y, void* act_dy, int act_ldy, float* act_dbias, int act_pool2, int act_relu,
                                               int act_hx, int act_wx, void* stream) {
    GemmEpilogue x = {};
    x.cluster_s = cluster_s; x.pooled = pooled; x.ldp = ldp; x.pool_w = pool_w;
    x.act_y = act_y; x.act_dy = act_dy; x.act_ldy = act_ldy; x.act_dbias = act_dbias; x.act_pool2 = act_pool2; x.act_relu = act_relu;
    x.act_hx = act_hx; x.act_wx = act_wx;
    return conv_implicit_impl<ElemBF16>(nullptr, kind, act, ld_act, B, Ha, Wa, Ca, other, ld_other, KH, KW, SH, SW, PH, PW, n_out, out, out_is_f32, ldo,
                                   bias, relu, 1, 0, tile_n, stream, &x);
}

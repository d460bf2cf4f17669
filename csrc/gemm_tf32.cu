// tf32 entry points of the tcgen05 GEMM family (kind::tf32): fp32 tensors in memory, read by the tensor cores as tf32, fp32 accumulation
// and fp32 outputs.  This is the reference's own precision — it trains fp32 tensors end to end and its convolutions run through cuDNN's
// TF32 kernels (example/main.py:41,53-54,70-73; asgd/optim/Asynchronous.py:27,55) — so the headline benchmark can be like for like.
// Kernels: gemm_tcgen05.cuh (ElemTF32: 32 fp32 per 128-byte k-block, UMMA_K = 8, MN-major tiles in the 32-byte-granule swizzle).
#include "gemm_tcgen05.cuh"

using namespace dmlb;

extern "C" int dmlb_gemm_tf32(const void* A, int lda, int a_mn_major, const void* B, int ldb, int b_mn_major, int M, int N, int K, void* out,
                              int ldo, const float* bias, int relu, int split_k, int atomic, int perm_cin, int perm_khw, int tile_n,
                              void* stream) {
    return gemm_impl<ElemTF32>(nullptr, A, lda, a_mn_major, B, ldb, b_mn_major, M, N, K, out, 1, ldo, bias, relu, split_k, atomic, perm_cin,
                               perm_khw, tile_n, stream);
}

extern "C" int dmlb_conv_implicit_tf32(int kind, const void* act, int ld_act, int B, int Ha, int Wa, int Ca, const void* other, int ld_other,
                                       int KH, int KW, int SH, int SW, int PH, int PW, int n_out, void* out, int ldo, const float* bias,
                                       int relu, int split_k, int atomic, int tile_n, void* stream) {
    return conv_implicit_impl<ElemTF32>(nullptr, kind, act, ld_act, B, Ha, Wa, Ca, other, ld_other, KH, KW, SH, SW, PH, PW, n_out, out, 1, ldo,
                                        bias, relu, split_k, atomic, tile_n, stream);
}

// Small-channel first-layer convolution (Cin <= 4): see conv_smallc_impl in gemm_tcgen05.cuh and include/dmlb.h.
extern "C" int dmlb_conv_smallc_tf32(int kind, const void* act, int B, int Ha, int Wa, int cin, const void* other, int ld_other, int KH, int KW,
                                     int SH, int SW, int PH, int PW, int n_out, void* out, int out_is_f32, int ldo, const float* bias, int relu,
                                     int split_k, int pt_layout, void* pooled, int ldp, int pool_w, int tile_n, int hp, int wp, void* stream) {
    return conv_smallc_impl<ElemTF32>(kind, act, B, Ha, Wa, cin, other, ld_other, KH, KW, SH, SW, PH, PW, n_out, out, out_is_f32, ldo, bias, relu,
                                 split_k, pt_layout, pooled, ldp, pool_w, tile_n, hp, wp, stream);
}

// Implicit convolution, cluster split-K form (kind 0 fprop / 2 dgrad): the K-slices of a tile reduce through distributed shared memory and
// the fused epilogue runs on the sum — see gemm_tcgen05.cuh (kCluster) and include/dmlb.h.
extern "C" int dmlb_conv_implicit_cluster_tf32(int kind, const void* act, int ld_act, int B, int Ha, int Wa, int Ca, const void* other, int ld_other,
                                               int KH, int KW, int SH, int SW, int PH, int PW, int n_out, void* out, int out_is_f32, int ldo,
                                               const float* bias, int relu, int cluster_s, int tile_n, void* pooled, int ldp, int pool_w,
                                               const void* act_y, void* act_dy, int act_ldy, float* act_dbias, int act_pool2, int act_relu,
                                               int act_hx, int act_wx, void* stream) {
    GemmEpilogue x = {};
    x.cluster_s = cluster_s; x.pooled = pooled; x.ldp = ldp; x.pool_w = pool_w;
    x.act_y = act_y; x.act_dy = act_dy; x.act_ldy = act_ldy; x.act_dbias = act_dbias; x.act_pool2 = act_pool2; x.act_relu = act_relu;
    x.act_hx = act_hx; x.act_wx = act_wx;
    return conv_implicit_impl<ElemTF32>(nullptr, kind, act, ld_act, B, Ha, Wa, Ca, other, ld_other, KH, KW, SH, SW, PH, PW, n_out, out, out_is_f32, ldo,
                                   bias, relu, 1, 0, tile_n, stream, &x);
}

/* C ABI of libdmlb.so — the sm_100a kernel library of distributed_ml_pytorch_b200.
 *
 * Every entry point returns 0 on success and a negative value on failure; dmlb_last_error() then
 * holds a human-readable reason.  Pointers are device pointers unless stated otherwise; `stream` is a
 * cudaStream_t passed as void* (NULL = legacy default stream).  All launches are asynchronous.
 *
 * The reference (bkpcoding/distributed_ML_pytorch) ships no native code; each
 * declaration cites the reference call site whose work the kernel takes over.
 */
#ifndef DMLB_H_
#define DMLB_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- runtime ------------------------------------------------------------------------------- */
int dmlb_abi_version(void);
const char* dmlb_last_error(void);
void dmlb_set_error(const char* msg);
long long dmlb_launch_count(void);       /* kernels launched by this library since the last reset */
void dmlb_reset_launch_count(void);
void dmlb_count_launch(void);
int dmlb_sm_count(void);
int dmlb_pdl_enabled(void);               /* programmatic dependent launch between consecutive kernels (default 1; env DMLB_PDL=0) */
void dmlb_set_pdl(int on);
int dmlb_gemm_persistent_enabled(void);   /* persistent double-buffered GEMM for multi-wave short-K grids (opt-in: default 0; env DMLB_PERSISTENT=1) */
void dmlb_set_gemm_persistent(int on);
/* 256-wide tiles of plain K-major GEMMs run on CTA pairs (tcgen05 cta_group::2, one 256 x 256 tile per two SMs); 0 switches that off. */
void dmlb_set_gemm_pair(int on);
int dmlb_gemm_pair_enabled(void);

/* ---- tcgen05 GEMM (reference: nn.Linear / nn.Conv2d in example/models.py:8-13,28-43) ------------
 * D[M,N] (+)= A[M,K] * B[N,K]^T, bf16 operands, fp32 accumulation in TMEM.
 *   a_mn_major = 0: A is row-major [M,K] with pitch lda;  1: A is stored transposed, row-major [K,M].
 *   b_mn_major = 0: B is row-major [N,K] with pitch ldb;  1: B is stored transposed, row-major [K,N].
 *   (A MN-major requires B MN-major.)  Pitches are in elements and must be multiples of 8.
 *   out: bf16 (out_is_f32=0) or fp32 (=1), row-major [M,N] with pitch ldo.
 *   bias: optional fp32 [N]; relu: fused max(x,0).
 *   split_k > 1 partitions K over gridDim.z and accumulates with fp32 red.add (out must be fp32 and
 *   hold the value to accumulate onto, e.g. zeros); atomic=1 does the same with a single split.
 *   perm_cin > 0 scatters output column c to (c % perm_cin) * perm_khw + c / perm_cin, which turns a
 *   (kh,kw,cin)-ordered conv weight gradient into PyTorch's [Cout,Cin,kh,kw] layout.
 *   tile_n: 0 = auto, else 32 / 64 / 128 / 256.
 */
int dmlb_gemm_bf16(const void* A, int lda, int a_mn_major, const void* B, int ldb, int b_mn_major, int M, int N, int K,
                   void* out, int out_is_f32, int ldo, const float* bias, int relu, int split_k, int atomic, int perm_cin,
                   int perm_khw, int tile_n, void* stream);

/* The same GEMM on fp32 tensors read by the tensor cores as tf32 (tcgen05 kind::tf32), fp32 accumulation, fp32 output: the precision
 * the reference trains at (fp32 tensors, cuDNN TF32 convolutions — example/main.py:41,53-54,70-73).  Pitches in elements, multiples of 4.
 * tile_n: 0 = auto, else 32 / 64 / 128. */
int dmlb_gemm_tf32(const void* A, int lda, int a_mn_major, const void* B, int ldb, int b_mn_major, int M, int N, int K, void* out, int ldo,
                   const float* bias, int relu, int split_k, int atomic, int perm_cin, int perm_khw, int tile_n, void* stream);
/* dmlb_conv_implicit_bf16 on fp32 channels-last tensors (tf32 math): the gathered tensor needs Ca % 32 == 0. */
int dmlb_conv_implicit_tf32(int kind, const void* act, int ld_act, int B, int Ha, int Wa, int Ca, const void* other, int ld_other, int KH, int KW,
                            int SH, int SW, int PH, int PW, int n_out, void* out, int ldo, const float* bias, int relu, int split_k,
                            int atomic, int tile_n, void* stream);

/* Small-channel first-layer convolution as an implicit GEMM (reference: the first nn.Conv2d of example/models.py:8,29 — 3 input channels).
 * act is channels-last with ONE 16-byte granule per pixel (bf16: 8 channels, fp32/tf32: 4; unused channels zero; see dmlb_nchw_to_act_*);
 * a 128-byte k-block gathers eight filter taps by TMA im2col, so the reduction is kh*kw*granule long.
 * kind 0 fprop: other = weights [n_out, kh*kw*granule] (pad channels zero); optional fused 2x2 max-pool into pooled (pool_w = output width).
 * kind 1 wgrad: other = dY [pixels, ld_other]; out fp32 [n_out, kh*kw*cin] (+)=, PyTorch's [Cout,Cin,kh,kw] order when pt_layout = 1
 *   (bf16 only in this form).
 * Row-window form (hp > 0, both dtypes, output extents powers of two): act is the zero-padded buffer [B, hp, wp, granule] written by
 *   dmlb_nchw_to_act_* with padding, the weights are [n_out, kh * kwpad * granule] with kwpad = 8*ceil(kw/8), and a whole k-block
 *   (128 pixels x 8 taps) is ONE tiled TMA load through an overlapping-window tensor map. */
int dmlb_conv_smallc_bf16(int kind, const void* act, int B, int Ha, int Wa, int cin, const void* other, int ld_other, int KH, int KW, int SH,
                          int SW, int PH, int PW, int n_out, void* out, int out_is_f32, int ldo, const float* bias, int relu, int split_k,
                          int pt_layout, void* pooled, int ldp, int pool_w, int tile_n, int hp, int wp, void* stream);
int dmlb_conv_smallc_tf32(int kind, const void* act, int B, int Ha, int Wa, int cin, const void* other, int ld_other, int KH, int KW, int SH,
                          int SW, int PH, int PW, int n_out, void* out, int out_is_f32, int ldo, const float* bias, int relu, int split_k,
                          int pt_layout, void* pooled, int ldp, int pool_w, int tile_n, int hp, int wp, void* stream);

/* Implicit-GEMM convolution: the activation operand is gathered by TMA in im2col mode, no column matrix exists in memory.
 * kind 0 fprop (act = x, other = W shadow [n_out=Cout, kh*kw*Cin]), 1 wgrad (act = x, other = dY, out [n_out=Cout, kh*kw*Cin], fp32 accumulate),
 * 2 dgrad (act = dY with Ca = Cout, other = W shadow, n_out = Cin, stride 1 only).  The gathered tensor needs Ca % 64 == 0.
 * Reference: nn.Conv2d forward / backward of example/models.py:29-39 (cuDNN there). */
int dmlb_conv_implicit_bf16(int kind, const void* act, int ld_act, int B, int Ha, int Wa, int Ca, const void* other, int ld_other, int KH, int KW,
                            int SH, int SW, int PH, int PW, int n_out, void* out, int out_is_f32, int ldo, const float* bias, int relu,
                            int split_k, int atomic, int tile_n, void* stream);

/* dmlb_conv_implicit_* with the K reduction split over a thread-block CLUSTER (cluster_s = 2, 4 or 8 CTAs per output tile) instead of an
 * fp32 workspace + finishing kernel: every CTA scatters its TMEM accumulator through distributed shared memory, CTA r sums and finalises
 * tile_n / cluster_s columns (4..32).  kind 0 fprop: bias, ReLU, optional 2x2 max-pool into pooled (pool_w = map width, power of two <= 16).
 * kind 2 dgrad: plain store of dx, or (act_y != NULL) the producer block's pool / ReLU backward + bias gradient as in
 * dmlb_splitk_finish_actbwd: out is then unused, act_dy receives the producer's pre-activation gradient.
 * Reference: nn.Conv2d forward / backward of example/models.py:32-41 at batch 64 (tiny M, long K). */
int dmlb_conv_implicit_cluster_bf16(int kind, const void* act, int ld_act, int B, int Ha, int Wa, int Ca, const void* other, int ld_other, int KH,
                                    int KW, int SH, int SW, int PH, int PW, int n_out, void* out, int out_is_f32, int ldo, const float* bias,
                                    int relu, int cluster_s, int tile_n, void* pooled, int ldp, int pool_w, const void* act_y, void* act_dy,
                                    int act_ldy, float* act_dbias, int act_pool2, int act_relu, int act_hx, int act_wx, void* stream);
int dmlb_conv_implicit_cluster_tf32(int kind, const void* act, int ld_act, int B, int Ha, int Wa, int Ca, const void* other, int ld_other, int KH,
                                    int KW, int SH, int SW, int PH, int PW, int n_out, void* out, int out_is_f32, int ldo, const float* bias,
                                    int relu, int cluster_s, int tile_n, void* pooled, int ldp, int pool_w, const void* act_y, void* act_dy,
                                    int act_ldy, float* act_dbias, int act_pool2, int act_relu, int act_hx, int act_wx, void* stream);

/* Split-K forward (csrc/splitk.cu): for tiny-M, long-K layers (AlexNet conv3-5 at the reference's batch 64,
 * example/main.py:142).  ws is fp32 scratch [M, round_up(N,4)], all zeros on entry and zeroed again by the finish. */
int dmlb_splitk_finish(float* ws, int ldw, const float* bias, int relu, void* y, int ldy, long long M, int N, int pool2, int B, int H, int W,
                       void* pooled, int ldp, void* stream);
int dmlb_splitk_finish_f32(float* ws, int ldw, const float* bias, int relu, void* y, int ldy, long long M, int N, int pool2, int B, int H, int W,
                           void* pooled, int ldp, void* stream);   /* fp32 y / pooled (tf32 path) */
/* Finish of a split-K data-gradient GEMM fused with the activation backward of the block that produced the GEMM's input (K8): reads the
 * summed dgrad from ws (and zeroes it), applies that block's 2x2 max-pool backward (pool2) and ReLU mask using its saved activation y,
 * writes dy (the operand of the block's own wgrad / dgrad) and accumulates its bias gradient.  Replaces dmlb_splitk_finish + dmlb_act_bwd_*. */
int dmlb_splitk_finish_actbwd(float* ws, int ldw, const void* y, void* dy, int ldy, float* dbias, int B, int Hx, int Wx, int C, int pool2,
                              int relu, int is_f32, void* stream);
int dmlb_gemm_bias_act_splitk(const void* A, int lda, const void* Wt, int ldb, int M, int N, int K, float* ws, const float* bias, int relu,
                              void* y, int ldy, int split_k, void* stream);

/* Same computation as dmlb_gemm_bias_act_splitk, for callers that cannot allocate device memory (a host that links
 * this library without torch and owns no allocator, a service whose own pool already holds the whole GPU, a loop
 * that cannot take the device-wide synchronisation of cudaMalloc/cudaFree).  No counterpart in the reference.
 * The fp32 workspace is a fixed array inside the library: M * round_up(N,4) must not exceed
 * dmlb_noalloc_scratch_capacity() floats (1 Mi floats = 4 MiB per GPU, held until process exit); larger problems are
 * rejected with an error.  Calls no cudaMalloc/cudaFree or other allocating routine.  The first call on a device
 * resolves the array's address (make it during start-up).  All calls on a device share
 * the array (each call clears the region it uses first), so calls whose GPU work overlaps race: keep them on one
 * stream or order them yourself.  See csrc/splitk_static.cu. */
long long dmlb_noalloc_scratch_capacity(void);
int dmlb_gemm_bias_act_splitk_noalloc(const void* A, int lda, const void* Wt, int ldb, int M, int N, int K, const float* bias, int relu,
                                      void* y, int ldy, int split_k, void* stream);

/* ---- conv / pool / activation data movement (reference: example/models.py:16-17,28-41) ----------
 * Activations are bf16 channels-last [B,H,W,ld] with ld >= C; column matrices are [B*Ho*Wo, Kpad] with K
 * ordered (kh,kw,cin). */
int dmlb_im2col_nhwc_bf16(const void* x, void* col, int B, int H, int W, int C, int ldx, int KH, int KW, int SH, int SW, int PH, int PW,
                          int Kpad, void* stream);
/* First layer: gathers straight from the user's fp32 NCHW batch (reference main.py:64-70), fusing cast + re-layout. */
int dmlb_im2col_nchw_f32(const float* x, void* col, int B, int H, int W, int C, int KH, int KW, int SH, int SW, int PH, int PW, int Kpad,
                         void* stream);
int dmlb_col2im_nhwc_bf16(const void* dcol, void* dx, int B, int H, int W, int C, int ldx, int KH, int KW, int SH, int SW, int PH, int PW,
                          int Kpad, void* stream);
int dmlb_maxpool_fwd_bf16(const void* x, void* y, int B, int H, int W, int C, int ldx, int ldo, int k, int s, int p, void* stream);
/* General windows (ResNet stem 3x3/2/1): forward records the arg-max tap per output (uint8 [B,Ho,Wo,C]); backward gathers with it. */
int dmlb_maxpool_fwd_idx_bf16(const void* x, void* y, void* idx, int B, int H, int W, int C, int ldx, int ldo, int k, int s, int p, void* stream);
int dmlb_maxpool_bwd_idx_bf16(const void* dout, const void* idx, void* dx, int B, int H, int W, int C, int lddx, int ldd, int k, int s, int p,
                              void* stream);
/* Fused (max-pool backward) o (ReLU mask) + bias gradient; see csrc/conv_ops.cu. dbias (fp32 [C], red.add) may be NULL. */
int dmlb_act_bwd_bf16(const void* y, const void* dout, void* dy, float* dbias, int B, int H, int W, int C, int ldy, int ldd, int has_pool,
                      int k, int s, int p, int relu, void* stream);
int dmlb_relu_fwd_bf16(const void* x, void* y, long long rows, int C, int ldx, int ldy, void* stream);
int dmlb_relu_bwd_bf16(const void* y, const void* dy, void* dx, long long rows, int C, int ldy, int lddy, int lddx, void* stream);
/* The same kernels on fp32 channels-last activations [B,H,W,ld] (ld % 4 == 0) for the tf32 path — the reference keeps fp32 activations
 * (example/main.py:64-73).  dmlb_im2col_nchw_f32_f32 writes an fp32 column matrix. */
int dmlb_im2col_nhwc_f32(const void* x, void* col, int B, int H, int W, int C, int ldx, int KH, int KW, int SH, int SW, int PH, int PW, int Kpad,
                         void* stream);
int dmlb_im2col_nchw_f32_f32(const float* x, void* col, int B, int H, int W, int C, int KH, int KW, int SH, int SW, int PH, int PW, int Kpad,
                             void* stream);
int dmlb_col2im_nhwc_f32(const void* dcol, void* dx, int B, int H, int W, int C, int ldx, int KH, int KW, int SH, int SW, int PH, int PW, int Kpad,
                         void* stream);
int dmlb_maxpool_fwd_f32(const void* x, void* y, int B, int H, int W, int C, int ldx, int ldo, int k, int s, int p, void* stream);
int dmlb_act_bwd_f32(const void* y, const void* dout, void* dy, float* dbias, int B, int H, int W, int C, int ldy, int ldd, int has_pool, int k,
                     int s, int p, int relu, void* stream);
int dmlb_relu_fwd_f32(const void* x, void* y, long long rows, int C, int ldx, int ldy, void* stream);
int dmlb_relu_bwd_f32(const void* y, const void* dy, void* dx, long long rows, int C, int ldy, int lddy, int lddx, void* stream);

/* ---- loss / optimizer / dropout (reference: example/main.py:71,75; asgd/optim/Asynchronous.py:55,63-68) ---- */
/* metrics[0] = mean loss, metrics[1] = #rows with argmax == label; dlogits (dtype of logits, [B,ldd], may be NULL) = (softmax-onehot)/B. */
int dmlb_softmax_xent(const void* logits, int logits_is_bf16, const long long* labels, float* metrics, void* dlogits, int B, int C,
                      int ldl, int ldd, void* stream);
/* Same, and the kernel also posts {loss, hits, launch number + 1 (mod 2^24)} into slot (launch number mod ring_slots) of a ring of 16-byte
 * slots in pinned, device-visible host memory (sequence word last): the host reads a step's result from memory, with no D2H copy and no
 * event between two replays of the training step (the reference blocks on loss.item() every step, example/main.py:79). */
int dmlb_softmax_xent_ring(const void* logits, int logits_is_bf16, const long long* labels, float* metrics, void* dlogits, int B, int C,
                           int ldl, int ldd, float* host_ring, void* ring_ctr, int ring_slots, void* stream);
int dmlb_scale_rows_bf16(void* x, const float* scale, long long rows, int C, int ld, void* stream);
int dmlb_scale_rows_f32(void* x, const float* scale, long long rows, int C, int ld, void* stream);
/* theta -= lr*g over the flat arena (n % 4 == 0); optional Downpour accumulator acc -= lr*g, momentum buffer,
 * weight decay, and a flat bf16 mirror of theta (same offsets) for the tensor-core GEMMs; zero_grads=1 clears the
 * gradients as they are consumed (replaces the next step's zero_grad memset). */
int dmlb_sgd_flat(float* params, float* grads, float* acc, float* mom, void* shadow_bf16, long long n, const float* lr_ptr, float mu,
                  float wd, int zero_grads, void* stream);
/* dmlb_sgd_flat plus the table-backed weight shadows (absolute-address table of dmlb_refresh_shadows) refreshed by the same launch;
 * [table_lo, table_hi) = element range of params that holds those weights. */
int dmlb_sgd_flat_ex(float* params, float* grads, float* acc, float* mom, void* shadow_bf16, long long n, const float* lr_ptr, float mu,
                     float wd, int zero_grads, const void* table, int n_entries, int table_is_f32, long long table_lo, long long table_hi,
                     void* stream);
/* fp32 NCHW batch (the reference's input, example/main.py:64-70) -> channels-last activation with one 16-byte granule per pixel
 * (C <= 8 bf16 / 4 fp32 channels, the rest zero): the input layout of dmlb_conv_smallc_*. */
int dmlb_nchw_to_act_bf16(const float* x, void* y, int B, int C, int H, int W, int hp, int wp, int ph, int pw, void* stream);
int dmlb_nchw_to_act_f32(const float* x, void* y, int B, int C, int H, int W, int hp, int wp, int ph, int pw, void* stream);
/* (hp > 0: y is a zero-bordered buffer [B, hp, wp, granule] and the image lands at rows ph.., columns pw.. — the row-window form) */
/* table: device array of { int64 src_off, int64 dst_off, int32 rows, cin, khw, ld, cpad, src_gemm_order, kw, kwpad } — see csrc/train_ops.cu. */
int dmlb_refresh_shadows(const float* params, void* shadow, const void* table, int n_entries, void* stream);
int dmlb_refresh_shadows_f32(const float* params, void* shadow, const void* table, int n_entries, void* stream);   /* fp32 copies (tf32 path) */
int dmlb_cast_f32_bf16(const float* x, void* y, long long rows, int C, long long ldx, int ldy, void* stream);
int dmlb_cast_bf16_f32(const void* x, float* y, long long rows, int C, int ldx, long long ldy, void* stream);
int dmlb_dropout_fwd_bf16(const void* x, void* y, void* mask, long long rows, int C, int ldx, int ldy, int rows_per_sample, int channel_mode,
                          float p, unsigned long long seed, const void* step, void* stream);
int dmlb_dropout_bwd_bf16(const void* dy, const void* mask, void* dx, long long rows, int C, int lddy, int lddx, float p, void* stream);
int dmlb_dropout_fwd_f32(const void* x, void* y, void* mask, long long rows, int C, int ldx, int ldy, int rows_per_sample, int channel_mode,
                         float p, unsigned long long seed, const void* step, void* stream);
int dmlb_dropout_bwd_f32(const void* dy, const void* mask, void* dx, long long rows, int C, int lddy, int lddx, float p, void* stream);
int dmlb_counter_bump(void* counter, void* stream);

/* ---- batch norm / residual / global average pool (NORTH-STAR ResNets; csrc/norm_ops.cu) ----
 * bf16 channels-last rows [M, ld]; C % 8 == 0, C <= 2048.  sums / dsum are fp32 [2C] and must be zero on entry. */
int dmlb_bn_stats(const void* x, int ldx, float* sums, long long M, int C, void* stream);
/* y = act((x - mean) * rstd * gamma + beta [+ res]); training: statistics from `sums` (saved to save_mean_rstd [2C], folded
 * into running_mean / running_var); inference: sums = NULL and mean_rstd_in = {running_mean, 1/sqrt(running_var + eps)}.
 * zero_2c (may be NULL): an fp32 [2C] scratch this launch resets to zero — bn_apply cleans the backward scratch, bn_bwd_apply the
 * forward one, so a training step needs no memset launches for them. */
int dmlb_bn_apply(const void* x, int ldx, void* y, int ldy, const void* res, int ldr, const float* sums, const float* mean_rstd_in,
                  const float* gamma, const float* beta, float* save_mean_rstd, float* running_mean, float* running_var, long long M, int C,
                  float eps, float momentum, int relu, float* zero_2c, void* stream);
int dmlb_bn_bwd_reduce(const void* x, int ldx, const void* dy, int lddy, const void* yact, int ldya, void* g_out, int ldg,
                       const float* mean_rstd, float* dsum, long long M, int C, int relu, void* stream);
int dmlb_bn_bwd_apply(const void* x, int ldx, const void* g, int ldg, void* dx, int lddx, const float* mean_rstd, const float* gamma,
                      const float* dsum, float* dgamma, float* dbeta, long long M, int C, float* zero_2c, void* stream);
int dmlb_avgpool_fwd_bf16(const void* x, int ldx, void* y, int ldy, int B, int P, int C, void* stream);
int dmlb_avgpool_bwd_bf16(const void* dy, int lddy, void* dx, int lddx, int B, int P, int C, void* stream);

/* ---- split-K with the finishing pass folded into the GEMM kernel (one launch; measured slower than GEMM + dmlb_splitk_finish on the
 * AlexNet shapes, see DESIGN.md, hence opt-in).  y[M,N] bf16 = act(A*B^T + bias) and optionally pooled = maxpool2x2(y) when the rows
 * enumerate a [.., pool_h, pool_w] map (dmlb_fin_pool_supported).  ws (fp32 [M,ldw]) and counters (int [n_counters >= output tiles])
 * must be all zero on entry and are all zero again on exit.  kind: 0 fprop, 2 dgrad. ---- */
int dmlb_gemm_bf16_fin(const void* A, int lda, int a_mn_major, const void* B, int ldb, int b_mn_major, int M, int N, int K, float* ws, int ldw,
                       int* counters, int n_counters, const float* bias, int relu, void* y, int ldy, int pool_h, int pool_w, void* pooled,
                       int ldp, int split_k, int tile_n, void* stream);
int dmlb_conv_implicit_fin(int kind, const void* act, int ld_act, int B, int Ha, int Wa, int Ca, const void* other, int ld_other, int KH, int KW,
                           int SH, int SW, int PH, int PW, int n_out, float* ws, int ldw, int* counters, int n_counters, const float* bias,
                           int relu, void* y, int ldy, int pool_h, int pool_w, void* pooled, int ldp, int split_k, int tile_n, void* stream);
int dmlb_fin_pool_supported(int H, int W);

/* ---- GPU-resident input pipeline (reference: torchvision ToTensor + Normalize in a host DataLoader, example/main.py:23-40,64-65) ----
 * images uint8 [N,H,W,C] and labels int64 [N] stay in HBM; one launch gathers index[B], converts to CHW fp32 and normalises. */
int dmlb_gather_normalize_u8(const void* images, const void* labels, const void* index, float* out_x, void* out_y, int B, int H, int W, int C,
                             const float* mean3, const float* std3, void* stream);

/* ---- peer memory (reference transport: dist.init_process_group('gloo') + isend/recv, example/main.py:165) ---- */
int dmlb_peer_alloc(size_t bytes, void** out);            /* zeroed cudaMalloc, shareable with CUDA IPC */
int dmlb_peer_free(void* p);
int dmlb_ipc_get_handle(void* p, void* handle_out64);     /* 64-byte cudaIpcMemHandle_t */
int dmlb_ipc_open_handle(const void* handle64, void** out);
int dmlb_ipc_close(void* p);
int dmlb_enable_peer_access(int peer_dev);

/* ---- fused communication + update kernels (csrc/comm.cu).  Pointer tables are HOST arrays of `world` device
 * pointers, index = rank; the flat vector is cut into `world` shards of shard_size elements, GPU g owns shard g. ---- */
void dmlb_set_comm_max_blocks(int n);                      /* grid cap of dmlb_push_add / dmlb_pull (0 = default): small when overlapped with training */
int dmlb_comm_chunk_elems(void);                          /* mailbox chunk (shard sizes must be multiples of it) */
int dmlb_comm_flag_words(void);                           /* uint32 words of the per-GPU flag block of the sync kernel */
/* K1 (Asynchronous.py:55,58-60 + server add): master[owner][i] += acc[i]; acc[i] = 0 — vector red.add over NVLink. */
int dmlb_push_add(float* acc, void* const* master_ptrs, int world, long long shard_size, long long n, void* stream);
/* Delta forms of K1 / K2: the worker keeps a SNAPSHOT of its replica as of the last push instead of an accumulator; the update sent is
 * live - snapshot (exact), the snapshot advances to the value read.  Saves the accumulator's read + write in every training step and may
 * run concurrently with training.  dmlb_land_pull applies a fetched pull: params = pulled + (params - snapshot), snapshot = pulled. */
/* Mover used by dmlb_push_add / dmlb_push_delta / dmlb_pull: 0 = by size, 1 = per-thread ld / red.add, 2 = TMA bulk
 * (cp.async.bulk through a shared-memory ring, cp.reduce.async.bulk.add.f32 into the owner). */
void dmlb_set_comm_mover(int mover);
int dmlb_push_delta(const float* live, float* snapshot, void* const* master_ptrs, int world, long long shard_size, long long n, void* stream);
int dmlb_push_mailbox_delta(const float* live, float* snapshot, void* busy, const void* pulled_ver, void* const* mailbox_ptrs,
                            void* const* full_ptrs, int rank, int world, long long shard_size, void* stream);
int dmlb_land_pull(float* params, const float* pulled, float* snapshot, void* shadow_bf16, long long n, void* stream);
/* K3 (Asynchronous.py:17-18,48-49): params[i] = master[owner][i] by peer loads, fused with the cast into the worker's flat bf16
 * weight mirror (shadow_bf16, may be NULL); optional per-chunk version snapshot. */
int dmlb_pull(float* params, void* shadow_bf16, void* const* master_ptrs, void* const* version_ptrs, void* pulled_ver, int world,
              long long shard_size, long long n, void* stream);
/* K2 (NORTH-STAR momentum / staleness): chunked mailbox push and owner-side apply. */
int dmlb_push_mailbox(float* acc, void* busy, const void* pulled_ver, void* const* mailbox_ptrs, void* const* full_ptrs, int rank, int world,
                      long long shard_size, void* stream);
/* owner: v = mu * v + gain * s(tau) * delta; theta += v   (gain = 1: the classic update; (1 - mu) / workers keeps the step size of
 * plain SGD when several workers feed one momentum buffer) */
int dmlb_owner_apply(float* master, float* mom, void* version, void* full, const float* mailbox, void* const* busy_ptrs, void* stats, int rank,
                     int world, long long shard_size, float mu, int stale_mode, float stale_param, float gain, void* stream);
/* K4 (NORTH-STAR synchronous mode): barrier + reduce-scatter + SGD/momentum + all-gather + barrier in one kernel. */
int dmlb_sync_rs_update_ag(void* const* grad_ptrs, void* const* param_ptrs, void* const* flag_ptrs, float* mom, int rank, int world,
                           long long shard_size, const float* lr_ptr, float mu, float wd, float grad_scale, void* stream);

/* Range form for bucketed overlap with the backward pass: this GPU owns [base, base+n) in this launch. */
int dmlb_sync_rs_update_ag_range(void* const* grad_ptrs, void* const* param_ptrs, void* const* flag_ptrs, float* mom, int rank, int world,
                                 long long base, long long n, const float* lr_ptr, float mu, float wd, float grad_scale, int max_blocks,
                                 void* stream);

/* The same step through NVLS: multimem.ld_reduce (reduction inside the NVSwitch) + multimem.st (multicast) on multicast addresses. */
int dmlb_sync_nvls_range(const float* mc_grads, float* mc_params, const float* params_local, void* const* flag_ptrs, float* mom, int rank,
                         int world, long long base, long long n, const float* lr_ptr, float mu, float wd, float grad_scale, int max_blocks,
                         void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DMLB_H_ */

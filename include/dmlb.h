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

#ifdef __cplusplus
}
#endif
#endif /* DMLB_H_ */

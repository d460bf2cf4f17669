// tcgen05 / TMEM / TMA GEMM for sm_100a:  D[M,N] (+)= A[M,K] * B[N,K]^T   (bf16 or tf32 operands, fp32 accumulate)
//
// Templates + host launchers shared by gemm_tcgen05.cu (bf16 operands, kind::f16) and gemm_tf32.cu (fp32 operands read as tf32,
// kind::tf32 — the reference's own precision: fp32 tensors through cuDNN's TF32 convolutions, example/main.py:41,53-54,70-73).
//
// This single kernel family carries every GEMM-shaped op of the framework — Linear and the
// implicit-GEMM view of Conv2d, forward, data-gradient and weight-gradient — i.e. the work the
// reference delegates to cuDNN/cuBLAS through nn.Conv2d / nn.Linear
// (reference example/models.py:8-13,28-43; SURVEY.md §2.5, work-list rows K6-K9).
//
// Operand storage ("major"):
//   K-major  A: row-major [M, K] (lda)      MN-major A: row-major [K, M] (lda)   (i.e. A^T is stored)
//   K-major  B: row-major [N, K] (ldb)      MN-major B: row-major [K, N] (ldb)
// so that, with X the activation/column matrix, W the [Cout, K] weight and dY the output gradient:
//   fprop  Y  = X  * W^T      : A = X  (K-major),   B = W  (K-major)
//   dgrad  dX = dY * W        : A = dY (K-major),   B = W  (MN-major)     -- no transposed weight copy
//   wgrad  dW = dY^T * X      : A = dY (MN-major),  B = X  (MN-major)     -- no transposes at all
//
// Structure (one CTA per 128 x BN output tile and K-split):
//   warp 0   : TMA producer  (cp.async.bulk.tensor.2d, 128B swizzle, kStages-deep mbarrier ring)
//   warp 1   : allocates TMEM, one lane issues tcgen05.mma (M=128, N=BN, K=16) and tcgen05.commit
//   warps 2-5: epilogue — tcgen05.ld the fp32 accumulator (one TMEM lane = one output row per thread),
//              fused bias / ReLU / bf16 cast, or vector red.add for split-K and gradient accumulation,
//              optionally scattering columns (kh,kw,cin)->(cin,kh,kw) so conv weight gradients land
//              directly in the PyTorch-layout fp32 gradient arena.
#pragma once
#include "common.cuh"

namespace dmlb {

constexpr int kBM = 128;
constexpr int kBK = 64;             // 64 bf16 = one 128-byte swizzle row   (bf16-only CTA-pair / persistent kernels)
constexpr int kUmmaK = 16;
constexpr int kGemmThreads = 192;   // 6 warps

// Operand element traits.  A k-block is always 128 bytes of K per row (K-major: one swizzle row) or kBK K-rows of 128 bytes of MN
// (MN-major: one TMA box), so tiles have the same byte size for both element types and only the counts differ.
//   bf16: 64 elements per 128 B, UMMA_K = 16, MN-major tiles in the ordinary 128-byte swizzle (8-row atoms, 16-byte granules).
//   tf32: 32 elements per 128 B, UMMA_K = 8.  MN-major 32-bit operands exist in ONE shared-memory layout only: the 128-byte swizzle
//         with 32-byte granules (4-row atoms; TMA CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, descriptor layout type 1).
struct ElemBF16 {
    static constexpr int kBytes = 2, kBK = 64, kUmmaK = 16;
    static constexpr int kBoxBytes = kBK * 128;        // one MN-major TMA box (kBK K-rows x 128 B of MN)
    static constexpr int kMnStep = kUmmaK * 128;       // descriptor advance per UMMA_K inside an MN-major tile
    static constexpr uint32_t kMnSbo = 1024, kMnLayout = 2;
    static constexpr bool kTf32 = false;
    static constexpr CUtensorMapDataType kTmaType = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    static constexpr CUtensorMapSwizzle kTmaMnSwizzle = CU_TENSOR_MAP_SWIZZLE_128B;
};
struct ElemTF32 {
    static constexpr int kBytes = 4, kBK = 32, kUmmaK = 8;
    static constexpr int kBoxBytes = kBK * 128;
    static constexpr int kMnStep = kUmmaK * 128;
    static constexpr uint32_t kMnSbo = 512, kMnLayout = 1;
    static constexpr bool kTf32 = true;
    // TFLOAT32: the copy engine rounds the fp32 source to tf32 on its way into shared memory (the MMA alone would truncate)
    static constexpr CUtensorMapDataType kTmaType = CU_TENSOR_MAP_DATA_TYPE_TFLOAT32;
    static constexpr CUtensorMapSwizzle kTmaMnSwizzle = CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B;
};

struct GemmEpilogue {
    void* out;            // bf16 or fp32
    const float* bias;    // [N] or nullptr
    int ldo;              // elements
    int out_is_f32;
    int relu;
    int atomic;           // red.add into fp32 out
    int perm_cin;         // >0: out column c -> (c % perm_cin) * perm_khw + c / perm_cin
    int perm_khw;
    // Small-channel convolutions (first layers): the GEMM's columns are (tap, padded channel) with `cpad` channels per tap of which only
    // `cin` are real.  cpad > 0 (accumulating epilogue only): column c = tap * cpad + ch is dropped when ch >= cin, else lands at
    // tap * cin + ch (GEMM order) or, with perm_khw > 0, at ch * perm_khw + tap (PyTorch's [Cout, Cin, kh, kw]).
    int cpad, cin;
    int kw, kwpad;        // kwpad > 0: taps are (kh, kw padded to kwpad); columns with kw >= the real width are dropped too
    // 2x2 / stride-2 max-pool of the stored activation computed in the same epilogue (warp shuffles across the rows = pixels of the
    // tile): pool_w = W of the [B, H, W] map the rows enumerate (power of two <= 16, H even), 0 = off; pooled is [M/4, ldp].
    void* pooled;
    int ldp, pool_w;
    int use_fin;          // the fin_* member of the union below is live (folded split-K finish, kFin kernels)
    int cluster_s;        // host side only: > 1 = launch the cluster split-K form with this many K-slices per tile
    // The two rarely used argument groups share storage (kernel parameters are copied at every launch, and this step is launch-bound).
    union {
        // Split-K with the finishing pass folded in ("last CTA of a tile finishes"): every z-slice red.adds its partial tile into the
        // fp32 workspace `out`, then bumps the tile's arrival counter; the slice that arrives last reads the summed tile back from L2,
        // applies bias / ReLU, rounds to bf16 into fin_y (optionally also the 2x2/2 max-pool into fin_pooled), and writes zeros back,
        // so workspace and counters are clean for the next layer and no separate finishing launch is needed.
        struct {
            int* fin_counters;    // [tiles_n * tiles_m] ints, all zero on entry and on exit
            void* fin_y;          // bf16 [M, fin_ldy]
            const float* fin_bias;
            void* fin_pooled;     // bf16 [M/4, fin_ldp] when fin_pool_w > 0
            int fin_ldy, fin_ldp, fin_relu;
            int fin_pool_w;       // W of the [B,H,W] map the rows enumerate (H, W even, 128 % (2W) == 0); 0 = no pooling
        };
        // Cluster split-K epilogue only: activation backward of the block that produced this dgrad's input (see dmlb_splitk_finish_actbwd):
        // rows are pixels of that block's OUTPUT [B, act_hx, act_wx]; act_y / act_dy are its activation / pre-activation gradient
        // [B, H, W, act_ldy] with (H, W) = (act_hx, act_wx) or twice that when act_pool2; act_dbias (fp32 [N], may be null) += column sums.
        struct {
            const void* act_y;
            void* act_dy;
            float* act_dbias;
            int act_ldy, act_pool2, act_relu, act_hx, act_wx;
        };
    };
};
static_assert(sizeof(GemmEpilogue) == 128, "keep the epilogue arguments compact");

// Operand sources.  A: 0 = K-major matrix (2-D TMA), 1 = MN-major matrix, 2 = im2col gather of an NHWC activation
// (implicit-GEMM convolution: fprop reads x, dgrad reads dY).  B: 0 = K-major matrix, 1 = MN-major matrix,
// 2 = im2col gather, MN-major (wgrad: the column matrix is never materialised), 3 = conv weight [Cout, kh*kw, Cin] read
// through a 3-D map as the MN-major operand of dgrad (no transposed / flipped weight copy).
// Small-channel convolutions (first layers: Cin <= one 16-byte granule, i.e. 4 fp32 / 8 bf16 channels after padding): A = 3 gathers
// EIGHT filter taps per 128-byte k-block, each tap one TMA im2col box of 128 pixels x 16 bytes written unswizzled (the K-major
// no-swizzle core-matrix layout: 8 pixels x 16 B = 128 B, next 8 pixels +128 B, next tap +2048 B); B = 4 is the same gather as the
// MN-major operand of wgrad (one box of kBK pixels x 16 B per tap).  K order = (tap, padded channel).
// Row-window forms (A = 4, B = 5), for the same layers when the input sits in a zero-PADDED granule buffer [B, Hp, Wp, granule] and
// the output extents are powers of two: eight horizontally adjacent taps of one output pixel are 128 CONTIGUOUS bytes of that
// buffer, and neighbouring output pixels start stride granules further — an overlapping-window tensor map (dims: bytes inside an image,
// wo, ho, n) turns a whole k-block (128 or kBK pixels x 128 B) into ONE ordinary tiled TMA load with the 128-byte swizzle.  No
// im2col mode, no padding logic (the zeros are in the buffer), fully coalesced; K order = (kh, kw padded to 8, padded channel).
struct ConvK {
    int C;            // channels of the gathered activation (Cin for fprop / wgrad, Cout for dgrad)
    int KH, KW;
    int P, Q;         // spatial extent of the pixel index space the gather walks (conv output for fprop/wgrad, dX for dgrad)
    int sh, sw;       // traversal stride
    int low_h, low_w; // bounding-box lower corner (= -padding for fprop/wgrad, -(k-1-padding) for dgrad)
    int flip;         // dgrad: weight tap = kh*kw-1 - gather tap
    int wp, kwb;      // row-window forms (A = 4 / B = 5): row pitch of the zero-padded input in pixels, k-blocks per filter row
};

template <int BN, bool kDual>
struct GemmSmem {
    static constexpr int kABytes = kBM * 128;      // 16 KiB: 128 rows x one 128-byte k-block (either element type)
    static constexpr int kBBytes = BN * 128;
    static constexpr int kStageBytes = kABytes + kBBytes;
    // BN <= 128: stages sized so that two CTAs fit in one SM (~100 KB each): one CTA's epilogue overlaps the other's main loop,
    // which is what the skinny-N, large-M convolution GEMMs (memory-bound) need.  BN = 256 (large square GEMMs): one CTA, deep ring.
    // Deep variant (small grids, at most one wave): long ring, one CTA per SM, so every CTA has an SM's L2 port to itself.
    static constexpr int kStagesDual = (BN <= 32) ? 5 : (BN <= 64 ? 4 : (BN <= 128 ? 3 : 4));
    static constexpr int kStagesDeep = (BN <= 64) ? 6 : (BN <= 128 ? 5 : 4);
    static constexpr int kStages = kDual ? kStagesDual : kStagesDeep;
    static constexpr int kBarrierBytes = 256;
    static constexpr int kTotal = kStages * kStageBytes + kBarrierBytes + 1024;  // +1024: manual 1 KiB alignment
};

// Store 32 consecutive output columns of one row: vector stores when the row segment is aligned, scalar otherwise.
__device__ __forceinline__ void store_row32(void* out, int out_is_f32, size_t row, int ldo, int col0, int ncols, const float (&f)[32]) {
    if (out_is_f32) {
        float* o = static_cast<float*>(out) + row * ldo;
        if (ncols == 32 && (ldo & 3) == 0) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) st_f32x4(o + col0 + j, make_float4(f[j], f[j + 1], f[j + 2], f[j + 3]));
        } else {
#pragma unroll
            for (int j = 0; j < 32; ++j)
                if (j < ncols) o[col0 + j] = f[j];
        }
    } else {
        __nv_bfloat16* o = static_cast<__nv_bfloat16*>(out) + row * ldo;
        if (ncols == 32 && (ldo & 7) == 0) {
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
                uint4 pk;
                pk.x = pack_bf16x2(f[j], f[j + 1]);
                pk.y = pack_bf16x2(f[j + 2], f[j + 3]);
                pk.z = pack_bf16x2(f[j + 4], f[j + 5]);
                pk.w = pack_bf16x2(f[j + 6], f[j + 7]);
                *reinterpret_cast<uint4*>(o + col0 + j) = pk;
            }
        } else {
#pragma unroll
            for (int j = 0; j < 32; ++j)
                if (j < ncols) o[col0 + j] = __float2bfloat16_rn(f[j]);
        }
    }
}

// Epilogue of one 128 x BN accumulator tile: every thread owns one TMEM lane = one output row, 32 columns at a time.
template <int BN>
__device__ __forceinline__ void gemm_epilogue_tile(uint32_t tmem_d, int q, int lane, int m0, int n0, int M, int N, const GemmEpilogue& ep,
                                                   bool add_bias) {
    const int row = m0 + q * 32 + lane;
    const bool row_ok = row < M;
#pragma unroll 1
    for (int c0 = 0; c0 < BN; c0 += 32) {
        if (n0 + c0 >= N) break;  // warp-uniform
        uint32_t v[32];
        tmem_ld_32x32(tmem_d + (static_cast<uint32_t>(q * 32) << 16) + c0, v);
        tmem_ld_wait();
        if (!row_ok && ep.pool_w == 0) continue;     // (with the fused pool every lane takes part in the shuffles)
        const int col0 = n0 + c0;
        const int ncols = min(32, N - col0);
        float f[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            float x = __uint_as_float(v[j]);
            if (add_bias && j < ncols) x += __ldg(ep.bias + col0 + j);
            if (ep.relu) x = fmaxf(x, 0.f);
            f[j] = x;
        }
        if (ep.atomic) {
            float* o = static_cast<float*>(ep.out) + static_cast<size_t>(row) * ep.ldo;
            if (ep.cpad > 0) {
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const int c = col0 + j;
                    int tap = c / ep.cpad;
                    const int ch = c - tap * ep.cpad;
                    bool ok = j < ncols && ch < ep.cin;
                    if (ep.kwpad > 0) {
                        const int kh = tap / ep.kwpad, kwp = tap - kh * ep.kwpad;
                        ok = ok && kwp < ep.kw;
                        tap = kh * ep.kw + kwp;
                    }
                    if (ok) red_add_f32(o + (ep.perm_khw > 0 ? ch * ep.perm_khw + tap : tap * ep.cin + ch), f[j]);
                }
            } else if (ep.perm_cin > 0) {
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const int c = col0 + j;
                    if (j < ncols) red_add_f32(o + (c % ep.perm_cin) * ep.perm_khw + c / ep.perm_cin, f[j]);
                }
            } else if (ncols == 32 && (ep.ldo & 3) == 0) {
#pragma unroll
                for (int j = 0; j < 32; j += 4) red_add_f32x4(o + col0 + j, f[j], f[j + 1], f[j + 2], f[j + 3]);
            } else {
#pragma unroll
                for (int j = 0; j < 32; ++j)
                    if (j < ncols) red_add_f32(o + col0 + j, f[j]);
            }
        } else if (ep.out_is_f32 && ep.perm_cin > 0) {
            float* o = static_cast<float*>(ep.out) + static_cast<size_t>(row) * ep.ldo;
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                const int c = col0 + j;
                if (j < ncols) o[(c % ep.perm_cin) * ep.perm_khw + c / ep.perm_cin] = f[j];
            }
        } else {
            if (!ep.out_is_f32) {
                // the stored activation is bf16: pool (and everything downstream) must see the rounded values
#pragma unroll
                for (int j = 0; j < 32; ++j) f[j] = __bfloat162float(__float2bfloat16_rn(f[j]));
            }
            if (row_ok) store_row32(ep.out, ep.out_is_f32, static_cast<size_t>(row), ep.ldo, col0, ncols, f);
            if (ep.pool_w > 0) {
                // rows of a warp are 32 consecutive pixels of a [.., H, W] map with W | 16: lane bit 0 is the column parity and lane bit
                // log2(W) the row parity, so the 2x2 window lives in lanes {l, l^1, l^W, l^W^1} of this warp
                const int W = ep.pool_w;
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    float m = fmaxf(f[j], __shfl_xor_sync(0xffffffffu, f[j], 1));
                    f[j] = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, W));
                }
                if (row_ok && (lane & 1) == 0 && (lane & W) == 0) {
                    const int irow = row / W;                                     // image row index over the whole batch (H is even)
                    const size_t mp = static_cast<size_t>(irow >> 1) * (W >> 1) + ((row - irow * W) >> 1);
                    store_row32(ep.pooled, ep.out_is_f32, mp, ep.ldp, col0, ncols, f);
                }
            }
        }
    }
}

// bf16 store of 4 finished values (bias / ReLU applied by the caller); returns them as rounded floats for the pool.
__device__ __forceinline__ void fin_store4(__nv_bfloat16* dst, float (&x)[4], int nvalid) {
    if (nvalid == 4) {
        uint2 pk;
        pk.x = pack_bf16x2(x[0], x[1]);
        pk.y = pack_bf16x2(x[2], x[3]);
        *reinterpret_cast<uint2*>(dst) = pk;
    } else {
        for (int j = 0; j < nvalid; ++j) dst[j] = __float2bfloat16_rn(x[j]);
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) x[j] = __bfloat162float(__float2bfloat16_rn(x[j]));
}

// Finishing pass of one 128 x BN tile by the 128 epilogue threads of the CTA that arrived last (see GemmEpilogue::fin_counters).
// A warp walks rows, its lanes the columns (4 floats each), so every access is a coalesced 512-byte row segment out of L2.
template <int BN>
__device__ __forceinline__ void splitk_finish_tile(const GemmEpilogue& ep, int ew, int lane, int m0, int n0, int M, int N) {
    float* __restrict__ ws = static_cast<float*>(ep.out);
    const int ldw = ep.ldo;
    __nv_bfloat16* __restrict__ y = static_cast<__nv_bfloat16*>(ep.fin_y);
    const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 1
    for (int c = lane * 4; c < BN; c += 128) {
        const int col = n0 + c;
        if (col >= N) break;
        const int nvalid = min(4, N - col);
        float b[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) b[j] = (ep.fin_bias != nullptr && j < nvalid) ? __ldg(ep.fin_bias + col + j) : 0.f;
        if (ep.fin_pool_w == 0) {
            constexpr int kU = 16;
#pragma unroll 1
            for (int r0 = ew * 32; r0 < ew * 32 + 32; r0 += kU) {
                float4 v[kU];
#pragma unroll
                for (int u = 0; u < kU; ++u) {
                    const int row = m0 + r0 + u;
                    v[u] = row < M ? __ldcg(reinterpret_cast<const float4*>(ws + static_cast<size_t>(row) * ldw + col)) : zero4;
                }
#pragma unroll
                for (int u = 0; u < kU; ++u) {
                    const int row = m0 + r0 + u;
                    if (row >= M) continue;
                    __stcg(reinterpret_cast<float4*>(ws + static_cast<size_t>(row) * ldw + col), zero4);
                    float x[4] = {v[u].x + b[0], v[u].y + b[1], v[u].z + b[2], v[u].w + b[3]};
                    if (ep.fin_relu) {
#pragma unroll
                        for (int j = 0; j < 4; ++j) x[j] = fmaxf(x[j], 0.f);
                    }
                    fin_store4(y + static_cast<size_t>(row) * ep.fin_ldy + col, x, nvalid);
                }
            }
        } else {
            // rows of the tile are NHWC pixels; m0 is a multiple of 2W, so the tile holds 128/(2W) row pairs = 32 whole windows
            __nv_bfloat16* __restrict__ pooled = static_cast<__nv_bfloat16*>(ep.fin_pooled);
            const int W = ep.fin_pool_w, Wp = W >> 1;
            const int pair0 = m0 / (2 * W);
            constexpr int kU = 4;
#pragma unroll 1
            for (int j0 = ew * 8; j0 < ew * 8 + 8; j0 += kU) {
                float4 v[kU][4];
                int base[kU];
#pragma unroll
                for (int u = 0; u < kU; ++u) {
                    const int j = j0 + u;
                    const int pr = j / Wp, wp = j - pr * Wp;
                    base[u] = m0 + pr * 2 * W + 2 * wp;
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const int row = base[u] + (q >> 1) * W + (q & 1);
                        v[u][q] = base[u] < M ? __ldcg(reinterpret_cast<const float4*>(ws + static_cast<size_t>(row) * ldw + col)) : zero4;
                    }
                }
#pragma unroll
                for (int u = 0; u < kU; ++u) {
                    if (base[u] >= M) continue;
                    const int j = j0 + u;
                    const int pr = j / Wp, wp = j - pr * Wp;
                    float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const int row = base[u] + (q >> 1) * W + (q & 1);
                        __stcg(reinterpret_cast<float4*>(ws + static_cast<size_t>(row) * ldw + col), zero4);
                        float x[4] = {v[u][q].x + b[0], v[u][q].y + b[1], v[u][q].z + b[2], v[u][q].w + b[3]};
                        if (ep.fin_relu) {
#pragma unroll
                            for (int jj = 0; jj < 4; ++jj) x[jj] = fmaxf(x[jj], 0.f);
                        }
                        fin_store4(y + static_cast<size_t>(row) * ep.fin_ldy + col, x, nvalid);
#pragma unroll
                        for (int jj = 0; jj < 4; ++jj) mx[jj] = fmaxf(mx[jj], x[jj]);
                    }
                    const size_t mp = static_cast<size_t>(pair0 + pr) * Wp + wp;
                    fin_store4(pooled + mp * ep.fin_ldp + col, mx, nvalid);
                }
            }
        }
    }
}

// Final stage of the cluster split-K kernel: one thread = one output row, `ncols` (<= 32, multiple of 4 except at the right edge) summed
// columns starting at col0.  Forward: bias / ReLU / store (+ 2x2 pool by shuffles).  Backward: either a plain store of dx, or the
// producer block's pool / ReLU backward and bias gradient (ep.act_y != nullptr) — what dmlb_splitk_finish_actbwd does as a kernel.
template <typename T>
__device__ __forceinline__ void actbwd_rows(const GemmEpilogue& ep, int row, bool row_ok, int lane, int col0, int ncols, const float (&g)[32]) {
    const T* __restrict__ y = static_cast<const T*>(ep.act_y);
    T* __restrict__ dy = static_cast<T*>(ep.act_dy);
    const int Hx = ep.act_hx, Wx = ep.act_wx;
    const int b = row / (Hx * Wx), rem = row - b * (Hx * Wx);
    const int hx = rem / Wx, wx = rem - hx * Wx;
    float sum[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) sum[j] = 0.f;
    if (row_ok) {
        if (!ep.act_pool2) {
            const size_t yi = (static_cast<size_t>(b * Hx + hx) * Wx + wx) * ep.act_ldy + col0;
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                if (j < ncols) {
                    const float yv = to_f32<T>(y[yi + j]);
                    const float o = (ep.act_relu && !(yv > 0.f)) ? 0.f : g[j];
                    dy[yi + j] = from_f32<T>(o);
                    sum[j] = o;
                }
            }
        } else {
            const int H = 2 * Hx, W = 2 * Wx;
            size_t yi[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) yi[q] = (static_cast<size_t>(b * H + 2 * hx + (q >> 1)) * W + 2 * wx + (q & 1)) * ep.act_ldy + col0;
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                if (j < ncols) {
                    float yv[4];
#pragma unroll
                    for (int q = 0; q < 4; ++q) yv[q] = to_f32<T>(y[yi[q] + j]);
                    int best = 0;
#pragma unroll
                    for (int q = 1; q < 4; ++q)
                        if (yv[q] > yv[best]) best = q;                   // strict: the first maximum wins (PyTorch's rule)
                    const float o = (ep.act_relu && !(yv[best] > 0.f)) ? 0.f : g[j];
#pragma unroll
                    for (int q = 0; q < 4; ++q) dy[yi[q] + j] = from_f32<T>(q == best ? o : 0.f);
                    sum[j] = o;
                }
            }
        }
    }
    if (ep.act_dbias != nullptr) {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            if (j < ncols) {                                                  // warp-uniform
                const float t = warp_sum(sum[j]);
                if (lane == 0 && t != 0.f) red_add_f32(ep.act_dbias + col0 + j, t);
            }
        }
    }
}

__device__ __forceinline__ void cluster_final_rows(const GemmEpilogue& ep, int row, bool row_ok, int lane, int col0, int ncols, float (&f)[32],
                                                   bool add_bias) {
    if (!ep.use_fin && ep.act_y != nullptr) {
        if (ep.out_is_f32) actbwd_rows<float>(ep, row, row_ok, lane, col0, ncols, f);
        else               actbwd_rows<__nv_bfloat16>(ep, row, row_ok, lane, col0, ncols, f);
        return;
    }
#pragma unroll
    for (int j = 0; j < 32; ++j) {
        float x = f[j];
        if (add_bias && j < ncols) x += __ldg(ep.bias + col0 + j);
        if (ep.relu) x = fmaxf(x, 0.f);
        if (!ep.out_is_f32) x = __bfloat162float(__float2bfloat16_rn(x));
        f[j] = x;
    }
    if (row_ok) store_row32(ep.out, ep.out_is_f32, static_cast<size_t>(row), ep.ldo, col0, ncols, f);
    if (ep.pool_w > 0) {
        const int W = ep.pool_w;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            float m = fmaxf(f[j], __shfl_xor_sync(0xffffffffu, f[j], 1));
            f[j] = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, W));
        }
        if (row_ok && (lane & 1) == 0 && (lane & W) == 0) {
            const int irow = row / W;
            const size_t mp = static_cast<size_t>(irow >> 1) * (W >> 1) + ((row - irow * W) >> 1);
            store_row32(ep.pooled, ep.out_is_f32, mp, ep.ldp, col0, ncols, f);
        }
    }
}

// kCluster: the gridDim.z K-slices of one output tile form a thread-block cluster (launch attribute, size <= 8).  Instead of red.adding
// partial tiles into an fp32 workspace that a second kernel finishes, the slices reduce through DISTRIBUTED SHARED MEMORY: after its
// main loop every CTA scatters its TMEM accumulator — column slice d to CTA d (st.shared::cluster into the operand stages, which are
// free by then) — and, after a cluster barrier, CTA r sums the S partial copies of its BN/S columns and runs the fused epilogue on them.
// No workspace, no atomics, no finishing launch.
template <typename E, int BN, int AM, int BMODE, bool kDual, bool kFin = false, bool kCluster = false>
__global__ void __launch_bounds__(kGemmThreads, ((kDual && BN <= 128) ? 2 : 1))
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                         GemmEpilogue ep, ConvK cv, int M, int N, int K, int kb_per_split) {
    constexpr bool A_MN = AM == 1;
    constexpr bool B_MN = BMODE >= 1;
    using S = GemmSmem<BN, kDual>;
    constexpr int kStages = S::kStages;
    constexpr int kTmemCols = BN < 32 ? 32 : BN;
    static_assert(BN % 16 == 0 && BN >= 32 && BN <= 256, "UMMA M=128 needs N%16==0, 16<=N<=256");
    static_assert(!B_MN || BMODE == 4 || BN % E::kBK == 0, "MN-major B tiles are built from kBK-wide TMA boxes");
    constexpr int kBK = E::kBK;        // shadows the bf16-only namespace constant inside this kernel

    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_base = smem_base + kStages * S::kStageBytes;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (kStages + s); };
    const uint32_t tmem_full_bar = bar_base + 8u * (2 * kStages);
    const uint32_t tmem_holder = bar_base + 8u * (2 * kStages + 1);
    uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
    volatile uint32_t* tmem_holder_ptr = reinterpret_cast<volatile uint32_t*>(smem_gen + kStages * S::kStageBytes + 8 * (2 * kStages + 1));

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int m0 = blockIdx.x * kBM;
    const int n0 = blockIdx.y * BN;
    const int nkb_total = (K + kBK - 1) / kBK;
    // cluster form: an even partition (every slice non-empty because the host keeps gridDim.z <= nkb_total)
    const int kb_begin = kCluster ? static_cast<int>(1ll * blockIdx.z * nkb_total / gridDim.z) : blockIdx.z * kb_per_split;
    const int kb_end = kCluster ? static_cast<int>(1ll * (blockIdx.z + 1) * nkb_total / gridDim.z) : min(nkb_total, kb_begin + kb_per_split);
    const int nkb = kb_end - kb_begin;  // host guarantees >= 1

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tma_a);
        tma_prefetch_desc(&tma_b);
        for (int s = 0; s < kStages; ++s) {
            mbar_init(full_bar(s), 1);
            mbar_init(empty_bar(s), 1);
        }
        mbar_init(tmem_full_bar, 1);
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc<kTmemCols>(tmem_holder);
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_d = *tmem_holder_ptr;
    // Everything above touched only shared memory / TMEM / kernel parameters: it overlapped the previous kernel.
    pdl_launch_dependents();
    pdl_wait();

    if (warp == 0) {
        // ------------------------------ TMA producer ------------------------------
        if (lane == 0) {
            // im2col A: the tile's first pixel in (n, p, q) and its input-space coordinates, computed once
            int a_n = 0, a_h = 0, a_w = 0;
            if constexpr (AM >= 2) {
                const int pq = cv.P * cv.Q;
                a_n = m0 / pq;
                const int rem = m0 - a_n * pq;
                const int pp = rem / cv.Q;
                a_h = cv.low_h + pp * cv.sh;
                a_w = cv.low_w + (rem - pp * cv.Q) * cv.sw;
            }
            for (int it = 0; it < nkb; ++it) {
                const int s = it % kStages;
                const uint32_t ph = (it / kStages) & 1u;
                mbar_wait(empty_bar(s), ph ^ 1u);
                const uint32_t a_dst = smem_base + s * S::kStageBytes;
                const uint32_t b_dst = a_dst + S::kABytes;
                const int k0 = (kb_begin + it) * kBK;
                mbar_arrive_expect_tx(full_bar(s), S::kStageBytes);
                if constexpr (AM == 0) {
                    tma_load_2d(a_dst, &tma_a, k0, m0, full_bar(s));
                } else if constexpr (AM == 1) {
#pragma unroll
                    for (int j = 0; j < kBM / kBK; ++j) tma_load_2d(a_dst + j * E::kBoxBytes, &tma_a, m0 + kBK * j, k0, full_bar(s));
                } else if constexpr (AM == 2) {
                    // k-block = kBK channels of one filter tap (C % kBK == 0): 128 pixels x 128 bytes, one instruction
                    const int tap = k0 / cv.C;
                    const int r = tap / cv.KW;
                    tma_load_im2col_4d(a_dst, &tma_a, k0 - tap * cv.C, a_w, a_h, a_n, tap - r * cv.KW, r, full_bar(s));
                } else if constexpr (AM == 4) {
                    // row-window: k-block (kh, kwb) = 8 taps of filter row kh starting at kw = 8*kwb, for 128 consecutive output pixels
                    const int kb = kb_begin + it;
                    const int kh = kb / cv.kwb;
                    const int pq = cv.P * cv.Q;
                    const int n_ = m0 / pq, rem = m0 - n_ * pq;
                    const int ho = rem / cv.Q;
                    tma_load_4d(a_dst, &tma_a, (kh * cv.wp + (kb - kh * cv.kwb) * 8) * (16 / E::kBytes), rem - ho * cv.Q, ho, n_, full_bar(s));
                } else {
                    // small-channel conv: k-block = 8 taps x one 16-byte channel granule; taps past the filter repeat the last one
                    // (their weights are zero-filled by B's map, the products vanish, the byte count stays that of a full stage)
                    const int ntaps = cv.KH * cv.KW;
                    const int tap0 = (kb_begin + it) * 8;
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const int tap = min(tap0 + j, ntaps - 1);
                        const int r = tap / cv.KW;
                        tma_load_im2col_4d(a_dst + j * 2048, &tma_a, 0, a_w, a_h, a_n, tap - r * cv.KW, r, full_bar(s));
                    }
                }
                if constexpr (BMODE == 0) {
                    tma_load_2d(b_dst, &tma_b, k0, n0, full_bar(s));
                } else if constexpr (BMODE == 1) {
#pragma unroll
                    for (int j = 0; j < BN / kBK; ++j) tma_load_2d(b_dst + j * E::kBoxBytes, &tma_b, n0 + kBK * j, k0, full_bar(s));
                } else if constexpr (BMODE == 2) {
                    // wgrad: reduction rows = kBK output pixels starting at k0, columns = kBK channels of the tap that owns them
                    const int pq = cv.P * cv.Q;
                    const int bn_ = k0 / pq;
                    const int rem = k0 - bn_ * pq;
                    const int pp = rem / cv.Q;
                    const int bh = cv.low_h + pp * cv.sh;
                    const int bw = cv.low_w + (rem - pp * cv.Q) * cv.sw;
#pragma unroll
                    for (int j = 0; j < BN / kBK; ++j) {
                        int kcol = n0 + kBK * j;
                        if (kcol >= N) kcol = 0;   // tile overhang: any legal box, the epilogue masks those columns
                        const int tap = kcol / cv.C;
                        const int r = tap / cv.KW;
                        tma_load_im2col_4d(b_dst + j * E::kBoxBytes, &tma_b, kcol - tap * cv.C, bw, bh, bn_, tap - r * cv.KW, r, full_bar(s));
                    }
                } else if constexpr (BMODE == 3) {
                    // dgrad: reduction rows = kBK output channels of one (flipped) tap, columns = kBK input channels
                    const int tap = k0 / cv.C;
                    const int wtap = cv.flip ? (cv.KH * cv.KW - 1 - tap) : tap;
#pragma unroll
                    for (int j = 0; j < BN / kBK; ++j) {
                        int cin0 = n0 + kBK * j;
                        if (cin0 >= N) cin0 = 0;
                        tma_load_3d(b_dst + j * E::kBoxBytes, &tma_b, cin0, wtap, k0 - tap * cv.C, full_bar(s));
                    }
                } else if constexpr (BMODE == 5) {
                    // row-window wgrad: reduction rows = kBK output pixels starting at k0; every kBK columns are one (kh, kwb) block
                    const int pq = cv.P * cv.Q;
                    const int n_ = k0 / pq, rem = k0 - n_ * pq;
                    const int ho = rem / cv.Q;
                    const int wo = rem - ho * cv.Q;
#pragma unroll
                    for (int j = 0; j < BN / kBK; ++j) {
                        int cb = n0 / kBK + j;
                        if (cb >= cv.KH * cv.kwb) cb = 0;          // tile overhang: any legal block, the epilogue masks those columns
                        const int kh = cb / cv.kwb;
                        tma_load_4d(b_dst + j * E::kBoxBytes, &tma_b, (kh * cv.wp + (cb - kh * cv.kwb) * 8) * (16 / E::kBytes), wo, ho, n_,
                                    full_bar(s));
                    }
                } else {
                    // small-channel wgrad: reduction rows = kBK output pixels starting at k0, columns = (tap, padded channel):
                    // one box of kBK pixels x 16 B per tap, unswizzled (MN-major core matrices: 8 pixels x 16 B)
                    constexpr int kGran = 16 / E::kBytes;                  // channels per 16-byte granule
                    const int ntaps = cv.KH * cv.KW;
                    const int pq = cv.P * cv.Q;
                    const int bn_ = k0 / pq;
                    const int rem = k0 - bn_ * pq;
                    const int pp = rem / cv.Q;
                    const int bh = cv.low_h + pp * cv.sh;
                    const int bw = cv.low_w + (rem - pp * cv.Q) * cv.sw;
                    const int tapb = n0 / kGran;
#pragma unroll 4
                    for (int j = 0; j < BN / kGran; ++j) {
                        const int tap = min(tapb + j, ntaps - 1);
                        const int r = tap / cv.KW;
                        tma_load_im2col_4d(b_dst + j * (kBK * 16), &tma_b, 0, bw, bh, bn_, tap - r * cv.KW, r, full_bar(s));
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------ MMA issuer ------------------------------
        if (lane == 0) {
            constexpr uint32_t idesc = E::kTf32 ? umma_idesc_tf32(kBM, BN, A_MN, B_MN) : umma_idesc_bf16(kBM, BN, A_MN, B_MN);
            for (int it = 0; it < nkb; ++it) {
                const int s = it % kStages;
                const uint32_t ph = (it / kStages) & 1u;
                mbar_wait(full_bar(s), ph);
                tcgen05_fence_after();
                const uint32_t a_src = smem_base + s * S::kStageBytes;
                const uint32_t b_src = a_src + S::kABytes;
#pragma unroll
                for (int k = 0; k < kBK / E::kUmmaK; ++k) {
                    // K-major: one UMMA_K = 32 B inside the 128 B swizzle row (16 bf16 / 8 tf32).
                    // MN-major: one UMMA_K = UMMA_K K-rows of 128 B; the next 128 B of MN start one TMA box (kBoxBytes) further.
                    // small-channel forms (unswizzled): A — one UMMA_K = 2 taps = 2 x 2048 B, taps 2048 B apart (LBO), 8-pixel groups
                    // 128 B apart (SBO); B — one UMMA_K = UMMA_K pixel rows of 16 B, 8-row groups 128 B apart (LBO), taps kBK*16 B apart (SBO)
                    const uint64_t adesc = AM == 3 ? umma_smem_desc_l(a_src + k * 4096, 2048, 128, 0)
                                           : A_MN  ? umma_smem_desc_l(a_src + k * E::kMnStep, E::kBoxBytes, E::kMnSbo, E::kMnLayout)
                                                   : umma_smem_desc(a_src + k * 32, 16, 1024);
                    const uint64_t bdesc = BMODE == 4 ? umma_smem_desc_l(b_src + k * (E::kUmmaK * 16), 128, kBK * 16, 0)
                                           : B_MN     ? umma_smem_desc_l(b_src + k * E::kMnStep, E::kBoxBytes, E::kMnSbo, E::kMnLayout)
                                                      : umma_smem_desc(b_src + k * 32, 16, 1024);
                    if constexpr (E::kTf32) umma_tf32(tmem_d, adesc, bdesc, idesc, (it > 0 || k > 0) ? 1u : 0u);
                    else                    umma_bf16(tmem_d, adesc, bdesc, idesc, (it > 0 || k > 0) ? 1u : 0u);
                }
                umma_commit(empty_bar(s));  // smem stage reusable once these MMAs retire
            }
            umma_commit(tmem_full_bar);     // accumulator complete
        }
    } else {
        // ------------------------------ epilogue ------------------------------
        const int q = warp & 3;                 // TMEM lane quadrant this warp may access
        mbar_wait(tmem_full_bar, 0);
        tcgen05_fence_after();
        if constexpr (!kCluster)
            gemm_epilogue_tile<BN>(tmem_d, q, lane, m0, n0, M, N, ep, ep.bias != nullptr && (!ep.atomic || blockIdx.z == 0));
        if constexpr (kFin) {     // separate instantiation: the finishing code costs ~100 registers the plain kernels should not carry
            // release: this thread's red.adds are ordered before the counter bump; acquire on the other side before reading back
            __threadfence();
            asm volatile("bar.sync 1, 128;" ::: "memory");
            if (warp == 2 && lane == 0) {
                int* ctr = ep.fin_counters + blockIdx.y * gridDim.x + blockIdx.x;
                const bool last = atomicAdd(ctr, 1) == static_cast<int>(gridDim.z) - 1;
                if (last) *ctr = 0;       // nobody else touches this counter before the next kernel
                tmem_holder_ptr[1] = last ? 1u : 0u;
            }
            asm volatile("bar.sync 1, 128;" ::: "memory");
            if (tmem_holder_ptr[1] != 0u) {
                __threadfence();
                splitk_finish_tile<BN>(ep, warp - 2, lane, m0, n0, M, N);
            }
        }
    }

    if constexpr (kCluster) {
        const uint32_t S = gridDim.z, rank = blockIdx.z;        // cluster = (1, 1, S): the rank is the K-slice index
        const uint32_t cs = BN / S;                              // columns this CTA finalises (host: multiple of 4, <= 32)
        // #1: every CTA of the cluster has retired its MMAs, i.e. nobody reads its operand stages any more: they become the exchange area
        //     [source slice][128 rows][cs columns] fp32 (S * 128 * cs * 4 = 128 * BN * 4 bytes <= the stage ring)
        cluster_sync_all();
        if (warp >= 2) {
            const int q = warp & 3;
            const uint32_t row = q * 32 + lane;
#pragma unroll 1
            for (int c0 = 0; c0 < BN; c0 += 32) {
                uint32_t v[32];
                tmem_ld_32x32(tmem_d + (static_cast<uint32_t>(q * 32) << 16) + c0, v);
                tmem_ld_wait();
#pragma unroll
                for (int g = 0; g < 8; ++g) {
                    const uint32_t col = c0 + 4 * g;
                    const uint32_t d = col / cs, off = col - d * cs;
                    const uint32_t laddr = smem_base + ((rank * 128u + row) * cs + off) * 4u;
                    st_cluster_f32x4(mapa_shared(laddr, d), __uint_as_float(v[4 * g]), __uint_as_float(v[4 * g + 1]), __uint_as_float(v[4 * g + 2]),
                                     __uint_as_float(v[4 * g + 3]));
                }
            }
        }
        cluster_sync_all();                                      // #2: all partial slices have landed (release / acquire at cluster scope)
        if (warp >= 2) {
            const int q = warp & 3;
            const uint32_t row = q * 32 + lane;
            float f[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = 0.f;
            for (uint32_t src = 0; src < S; ++src) {
#pragma unroll
                for (int g = 0; g < 8; ++g) {
                    if (4u * g < cs) {
                        const float4 t = ld_shared_f32x4(smem_base + ((src * 128u + row) * cs + 4u * g) * 4u);
                        f[4 * g] += t.x; f[4 * g + 1] += t.y; f[4 * g + 2] += t.z; f[4 * g + 3] += t.w;
                    }
                }
            }
            const int col0 = n0 + static_cast<int>(rank * cs);
            const int grow = m0 + static_cast<int>(row);
            if (col0 < N)       // warp-uniform
                cluster_final_rows(ep, grow, grow < M, lane, col0, min(static_cast<int>(cs), N - col0), f, ep.bias != nullptr);
        }
    }

    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) {
        tcgen05_fence_after();
        tmem_dealloc<kTmemCols>(tmem_d);
    }
}

// ----------------------------------------------------------------------------------------------
// CTA-pair variant (tcgen05 cta_group::2) for large GEMMs with both operands K-major: a cluster of two CTAs on the two SMs of a
// TPC computes one 256 x 256 output tile.  Each CTA stages its own 128 rows of A and only HALF of the B tile (128 of the 256
// columns); the pair's MMA (M = 256, issued by one thread of the leader CTA) reads both halves, so shared-memory traffic for B is
// halved per SM and a stage is 32 KB instead of 48 KB (6-deep ring).  Each CTA keeps its 128 x 256 fp32 accumulator in its own
// TMEM and runs its own epilogue.  Synchronisation: both producers' TMA bytes are counted on the LEADER's full barrier; the
// leader's tcgen05.commit multicasts the "stage free" / "accumulator ready" arrivals to the barriers of both CTAs.
// ----------------------------------------------------------------------------------------------
struct PairSmem {
    static constexpr int kBN = 256;
    static constexpr int kABytes = kBM * kBK * 2;          // 128 x 64 bf16 = 16 KiB (this CTA's rows of A)
    static constexpr int kBBytes = (kBN / 2) * kBK * 2;    // 128 x 64 bf16 = 16 KiB (this CTA's half of the B tile)
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kStages = 7;             // 7 x 32 KiB + barriers = 225.3 KiB of the 227 KiB a CTA may own
    static constexpr int kTotal = kStages * kStageBytes + 256 + 1024;
};

static __global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kGemmThreads, 1)
gemm_bf16_tcgen05_pair_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b, GemmEpilogue ep, int M,
                              int N, int K, int kb_per_split) {
    using S = PairSmem;
    constexpr int kStages = S::kStages;
    constexpr int BN = S::kBN;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_base = smem_base + kStages * S::kStageBytes;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (kStages + s); };
    const uint32_t tmem_full_bar = bar_base + 8u * (2 * kStages);
    const uint32_t tmem_holder = bar_base + 8u * (2 * kStages + 1);
    uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
    volatile uint32_t* tmem_holder_ptr = reinterpret_cast<volatile uint32_t*>(smem_gen + kStages * S::kStageBytes + 8 * (2 * kStages + 1));

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();                  // 0 = leader
    const int m0 = (blockIdx.x >> 1) * (2 * kBM) + static_cast<int>(rank) * kBM;   // this CTA's 128 rows
    const int n0 = blockIdx.y * BN;
    const int nkb_total = (K + kBK - 1) / kBK;
    const int kb_begin = blockIdx.z * kb_per_split;
    const int kb_end = min(nkb_total, kb_begin + kb_per_split);
    const int nkb = kb_end - kb_begin;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tma_a);
        tma_prefetch_desc(&tma_b);
        for (int s = 0; s < kStages; ++s) {
            mbar_init(full_bar(s), 1);      // leader's: one arrive.expect_tx covering the bytes of both CTAs
            mbar_init(empty_bar(s), 1);     // one multicast commit per use
        }
        mbar_init(tmem_full_bar, 1);
        fence_mbar_init();
    }
    cluster_sync_all();                     // both CTAs' barriers exist before anything can arrive on them remotely
    if (warp == 1) tmem_alloc_2cta<BN>(tmem_holder);
    tcgen05_fence_before();
    __syncthreads();
    cluster_sync_all();                     // both halves of the pair own their TMEM columns before the leader's first MMA
    tcgen05_fence_after();
    const uint32_t tmem_d = *tmem_holder_ptr;

    if (warp == 0) {
        if (lane == 0) {
            for (int it = 0; it < nkb; ++it) {
                const int s = it % kStages;
                const uint32_t ph = (it / kStages) & 1;
                mbar_wait(empty_bar(s), ph ^ 1);
                if (rank == 0) mbar_arrive_expect_tx(full_bar(s), 2u * S::kStageBytes);
                const uint32_t a_dst = smem_base + s * S::kStageBytes;
                const int kc = (kb_begin + it) * kBK;
                tma_load_2d_2cta(a_dst, &tma_a, kc, m0, full_bar(s));
                tma_load_2d_2cta(a_dst + S::kABytes, &tma_b, kc, n0 + static_cast<int>(rank) * (BN / 2), full_bar(s));
            }
        }
    } else if (warp == 1) {
        if (rank == 0 && lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(2 * kBM, BN, false, false);
            for (int it = 0; it < nkb; ++it) {
                const int s = it % kStages;
                const uint32_t ph = (it / kStages) & 1;
                mbar_wait(full_bar(s), ph);
                tcgen05_fence_after();
                const uint32_t a_src = smem_base + s * S::kStageBytes;
                const uint32_t b_src = a_src + S::kABytes;
#pragma unroll
                for (int k = 0; k < kBK / kUmmaK; ++k)
                    umma_bf16_2cta(tmem_d, umma_smem_desc(a_src + k * 32, 16, 1024), umma_smem_desc(b_src + k * 32, 16, 1024), idesc,
                                   (it > 0 || k > 0) ? 1u : 0u);
                umma_commit_2cta(empty_bar(s));       // frees stage s in both CTAs
            }
            umma_commit_2cta(tmem_full_bar);          // both epilogues may start
        }
    } else {
        const int q = warp & 3;
        mbar_wait(tmem_full_bar, 0);
        tcgen05_fence_after();
        gemm_epilogue_tile<BN>(tmem_d, q, lane, m0, n0, M, N, ep, ep.bias != nullptr && (!ep.atomic || blockIdx.z == 0));
    }

    tcgen05_fence_before();
    __syncthreads();
    cluster_sync_all();                     // neither CTA may leave (or free TMEM) while the pair's MMAs / multicast arrivals can still touch it
    if (warp == 1) {
        tcgen05_fence_after();
        tmem_dealloc_2cta<BN>(tmem_d);
    }
}

// Persistent form of the CTA-pair kernel: one cluster per TPC walks a strided list of 256 x 256 tiles.  The accumulator is
// double-buffered in TMEM (2 x 256 of the 512 columns of each SM), so the epilogue of tile t (tcgen05.ld, bias / ReLU, stores —
// 128 KB of output per CTA) runs while the pair's tensor cores already work on tile t+1, and the smem ring never drains between
// tiles.  "Accumulator drained" is signalled by the 4 epilogue warps of BOTH CTAs on the leader's barrier (8 arrivals).
static __global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kGemmThreads, 1)
gemm_bf16_tcgen05_pair_persistent_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b, GemmEpilogue ep,
                                         int M, int N, int K, int tiles_m2, int tiles_n, int group_m) {
    using S = PairSmem;
    constexpr int kStages = S::kStages;
    constexpr int BN = S::kBN;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_base = smem_base + kStages * S::kStageBytes;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (kStages + s); };
    auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * kStages + a); };
    auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * kStages + 2 + a); };
    const uint32_t tmem_holder = bar_base + 8u * (2 * kStages + 4);
    uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
    volatile uint32_t* tmem_holder_ptr = reinterpret_cast<volatile uint32_t*>(smem_gen + kStages * S::kStageBytes + 8 * (2 * kStages + 4));

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const int nkb = (K + kBK - 1) / kBK;
    const int num_tiles = tiles_m2 * tiles_n;
    const int first = blockIdx.x >> 1, stride = gridDim.x >> 1;
    // Tile order: bands of group_m row pairs, column by column inside a band, so the ~74 tiles in flight cover a near-square patch of C
    // and share A / B panels through L2 (grouped rasterisation).
    auto tile_mn = [&](int tile, int& mp, int& nt) {
        const int band = group_m * tiles_n;
        const int g = tile / band, r = tile - g * band;
        const int gm = min(group_m, tiles_m2 - g * group_m);
        nt = r / gm;
        mp = g * group_m + (r - nt * gm);
    };

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tma_a);
        tma_prefetch_desc(&tma_b);
        for (int s = 0; s < kStages; ++s) {
            mbar_init(full_bar(s), 1);
            mbar_init(empty_bar(s), 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(tfull_bar(a), 1);
            mbar_init(tempty_bar(a), 8);     // 4 epilogue warps x 2 CTAs (only the leader's copy is used)
        }
        fence_mbar_init();
    }
    cluster_sync_all();
    if (warp == 1) tmem_alloc_2cta<512>(tmem_holder);
    tcgen05_fence_before();
    __syncthreads();
    cluster_sync_all();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_holder_ptr;

    if (warp == 0) {
        if (lane == 0) {
            int it = 0;
            for (int tile = first; tile < num_tiles; tile += stride) {
                int mp, nt;
                tile_mn(tile, mp, nt);
                const int m0 = mp * (2 * kBM) + static_cast<int>(rank) * kBM;
                const int nb = nt * BN + static_cast<int>(rank) * (BN / 2);
                for (int kb = 0; kb < nkb; ++kb, ++it) {
                    const int s = it % kStages;
                    const uint32_t ph = (it / kStages) & 1u;
                    mbar_wait(empty_bar(s), ph ^ 1u);
                    if (rank == 0) mbar_arrive_expect_tx(full_bar(s), 2u * S::kStageBytes);
                    const uint32_t a_dst = smem_base + s * S::kStageBytes;
                    tma_load_2d_2cta(a_dst, &tma_a, kb * kBK, m0, full_bar(s));
                    tma_load_2d_2cta(a_dst + S::kABytes, &tma_b, kb * kBK, nb, full_bar(s));
                }
            }
        }
    } else if (warp == 1) {
        if (rank == 0 && lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(2 * kBM, BN, false, false);
            int it = 0, t = 0;
            for (int tile = first; tile < num_tiles; tile += stride, ++t) {
                const int acc = t & 1;
                mbar_wait(tempty_bar(acc), ((t >> 1) & 1u) ^ 1u);     // both epilogues have drained this accumulator
                tcgen05_fence_after();
                const uint32_t tmem_d = tmem_base + acc * BN;
                for (int kb = 0; kb < nkb; ++kb, ++it) {
                    const int s = it % kStages;
                    const uint32_t ph = (it / kStages) & 1u;
                    mbar_wait(full_bar(s), ph);
                    tcgen05_fence_after();
                    const uint32_t a_src = smem_base + s * S::kStageBytes;
                    const uint32_t b_src = a_src + S::kABytes;
#pragma unroll
                    for (int k = 0; k < kBK / kUmmaK; ++k)
                        umma_bf16_2cta(tmem_d, umma_smem_desc(a_src + k * 32, 16, 1024), umma_smem_desc(b_src + k * 32, 16, 1024), idesc,
                                       (kb > 0 || k > 0) ? 1u : 0u);
                    umma_commit_2cta(empty_bar(s));
                }
                umma_commit_2cta(tfull_bar(acc));
            }
        }
    } else {
        const int q = warp & 3;
        int t = 0;
        for (int tile = first; tile < num_tiles; tile += stride, ++t) {
            const int acc = t & 1;
            int mp, nt;
            tile_mn(tile, mp, nt);
            const int m0 = mp * (2 * kBM) + static_cast<int>(rank) * kBM;
            const int n0 = nt * BN;
            mbar_wait(tfull_bar(acc), (t >> 1) & 1u);
            tcgen05_fence_after();
            gemm_epilogue_tile<BN>(tmem_base + acc * BN, q, lane, m0, n0, M, N, ep, ep.bias != nullptr);
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_leader(tempty_bar(acc));
        }
    }

    tcgen05_fence_before();
    __syncthreads();
    cluster_sync_all();
    if (warp == 1) {
        tcgen05_fence_after();
        tmem_dealloc_2cta<512>(tmem_base);
    }
}

// ----------------------------------------------------------------------------------------------
// Persistent variant for grids of many tiles (ResNet-sized convolutions): one CTA per SM walks a strided list of output
// tiles.  Barrier set-up, TMEM allocation and descriptor prefetch are paid once per SM instead of once per tile, the smem
// ring keeps streaming across tile boundaries, and the accumulator is double-buffered in TMEM (2 x BN columns) so the
// epilogue of tile t (tcgen05.ld, bias/ReLU, stores) overlaps the TMA + MMA main loop of tile t+1.  No split-K here.
// ----------------------------------------------------------------------------------------------
template <int BN>
struct PersistSmem {
    static constexpr int kABytes = kBM * kBK * 2;
    static constexpr int kBBytes = BN * kBK * 2;
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kStages = (BN <= 64) ? 7 : (BN <= 128 ? 5 : 4);
    static constexpr int kTotal = kStages * kStageBytes + 256 + 1024;
};

template <int BN, int AM, int BMODE>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_bf16_tcgen05_persistent_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b, GemmEpilogue ep,
                                    ConvK cv, int M, int N, int K, int tiles_m, int tiles_n) {
    constexpr bool A_MN = AM == 1;
    constexpr bool B_MN = BMODE >= 1;
    using S = PersistSmem<BN>;
    constexpr int kStages = S::kStages;
    constexpr int kTmemCols = (2 * BN <= 32) ? 32 : (2 * BN <= 64 ? 64 : (2 * BN <= 128 ? 128 : (2 * BN <= 256 ? 256 : 512)));
    static_assert(2 * BN <= 512, "two accumulators must fit in TMEM");

    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_base = smem_base + kStages * S::kStageBytes;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (kStages + s); };
    auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * kStages + a); };
    auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * kStages + 2 + a); };
    const uint32_t tmem_holder = bar_base + 8u * (2 * kStages + 4);
    uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
    volatile uint32_t* tmem_holder_ptr = reinterpret_cast<volatile uint32_t*>(smem_gen + kStages * S::kStageBytes + 8 * (2 * kStages + 4));

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int nkb = (K + kBK - 1) / kBK;
    const int num_tiles = tiles_m * tiles_n;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tma_a);
        tma_prefetch_desc(&tma_b);
        for (int s = 0; s < kStages; ++s) {
            mbar_init(full_bar(s), 1);
            mbar_init(empty_bar(s), 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(tfull_bar(a), 1);
            mbar_init(tempty_bar(a), 4);     // one arrival per epilogue warp
        }
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc<kTmemCols>(tmem_holder);
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_holder_ptr;
    pdl_launch_dependents();
    pdl_wait();

    if (warp == 0) {
        if (lane == 0) {
            int it = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                const int m0 = (tile / tiles_n) * kBM;
                const int n0 = (tile % tiles_n) * BN;
                int a_n = 0, a_h = 0, a_w = 0;
                if constexpr (AM == 2) {
                    const int pq = cv.P * cv.Q;
                    a_n = m0 / pq;
                    const int rem = m0 - a_n * pq;
                    const int pp = rem / cv.Q;
                    a_h = cv.low_h + pp * cv.sh;
                    a_w = cv.low_w + (rem - pp * cv.Q) * cv.sw;
                }
                for (int kb = 0; kb < nkb; ++kb, ++it) {
                    const int s = it % kStages;
                    const uint32_t ph = (it / kStages) & 1u;
                    mbar_wait(empty_bar(s), ph ^ 1u);
                    const uint32_t a_dst = smem_base + s * S::kStageBytes;
                    const uint32_t b_dst = a_dst + S::kABytes;
                    const int k0 = kb * kBK;
                    mbar_arrive_expect_tx(full_bar(s), S::kStageBytes);
                    if constexpr (AM == 0) {
                        tma_load_2d(a_dst, &tma_a, k0, m0, full_bar(s));
                    } else if constexpr (AM == 1) {
                        tma_load_2d(a_dst, &tma_a, m0, k0, full_bar(s));
                        tma_load_2d(a_dst + 8192, &tma_a, m0 + 64, k0, full_bar(s));
                    } else {
                        const int tap = k0 / cv.C;
                        const int r = tap / cv.KW;
                        tma_load_im2col_4d(a_dst, &tma_a, k0 - tap * cv.C, a_w, a_h, a_n, tap - r * cv.KW, r, full_bar(s));
                    }
                    if constexpr (BMODE == 0) {
                        tma_load_2d(b_dst, &tma_b, k0, n0, full_bar(s));
                    } else if constexpr (BMODE == 1) {
#pragma unroll
                        for (int j = 0; j < BN / 64; ++j) tma_load_2d(b_dst + j * 8192, &tma_b, n0 + 64 * j, k0, full_bar(s));
                    } else if constexpr (BMODE == 2) {
                        const int pq = cv.P * cv.Q;
                        const int bn_ = k0 / pq;
                        const int rem = k0 - bn_ * pq;
                        const int pp = rem / cv.Q;
                        const int bh = cv.low_h + pp * cv.sh;
                        const int bw = cv.low_w + (rem - pp * cv.Q) * cv.sw;
#pragma unroll
                        for (int j = 0; j < BN / 64; ++j) {
                            int kcol = n0 + 64 * j;
                            if (kcol >= N) kcol = 0;
                            const int tap = kcol / cv.C;
                            const int r = tap / cv.KW;
                            tma_load_im2col_4d(b_dst + j * 8192, &tma_b, kcol - tap * cv.C, bw, bh, bn_, tap - r * cv.KW, r, full_bar(s));
                        }
                    } else {
                        const int tap = k0 / cv.C;
                        const int wtap = cv.flip ? (cv.KH * cv.KW - 1 - tap) : tap;
#pragma unroll
                        for (int j = 0; j < BN / 64; ++j) {
                            int cin0 = n0 + 64 * j;
                            if (cin0 >= N) cin0 = 0;
                            tma_load_3d(b_dst + j * 8192, &tma_b, cin0, wtap, k0 - tap * cv.C, full_bar(s));
                        }
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(kBM, BN, A_MN, B_MN);
            int it = 0, t = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++t) {
                const int acc = t & 1;
                mbar_wait(tempty_bar(acc), ((t >> 1) & 1u) ^ 1u);     // the epilogue has drained this accumulator
                tcgen05_fence_after();
                const uint32_t tmem_d = tmem_base + acc * BN;
                for (int kb = 0; kb < nkb; ++kb, ++it) {
                    const int s = it % kStages;
                    const uint32_t ph = (it / kStages) & 1u;
                    mbar_wait(full_bar(s), ph);
                    tcgen05_fence_after();
                    const uint32_t a_src = smem_base + s * S::kStageBytes;
                    const uint32_t b_src = a_src + S::kABytes;
#pragma unroll
                    for (int k = 0; k < kBK / kUmmaK; ++k) {
                        const uint64_t adesc = A_MN ? umma_smem_desc(a_src + k * 2048, 8192, 1024) : umma_smem_desc(a_src + k * 32, 16, 1024);
                        const uint64_t bdesc = B_MN ? umma_smem_desc(b_src + k * 2048, 8192, 1024) : umma_smem_desc(b_src + k * 32, 16, 1024);
                        umma_bf16(tmem_d, adesc, bdesc, idesc, (kb > 0 || k > 0) ? 1u : 0u);
                    }
                    umma_commit(empty_bar(s));
                }
                umma_commit(tfull_bar(acc));
            }
        }
    } else {
        const int q = warp & 3;
        int t = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++t) {
            const int acc = t & 1;
            const int m0 = (tile / tiles_n) * kBM;
            const int n0 = (tile % tiles_n) * BN;
            mbar_wait(tfull_bar(acc), (t >> 1) & 1u);
            tcgen05_fence_after();
            gemm_epilogue_tile<BN>(tmem_base + acc * BN, q, lane, m0, n0, M, N, ep, ep.bias != nullptr);
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty_bar(acc));
        }
    }

    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) {
        tcgen05_fence_after();
        tmem_dealloc<kTmemCols>(tmem_base);
    }
}


// ----------------------------------------------------------------------------------------------
// host side
// ----------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

// fp32 operands: TFLOAT32 (round to tf32 in the copy engine) unless DMLB_TMA_F32=1 asks for plain FLOAT32 copies (the MMA truncates).
template <typename E>
static CUtensorMapDataType tma_dtype() {
    if constexpr (E::kTf32) {
        static int plain = -1;
        if (plain < 0) {
            const char* e = getenv("DMLB_TMA_F32");
            plain = (e != nullptr && e[0] == '1') ? 1 : 0;
        }
        return plain ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : E::kTmaType;
    }
    return E::kTmaType;
}

// 2-D tensor map over a row-major [outer, inner] matrix with row pitch ld (elements), 128B swizzle.  mn_major: the box is an
// MN-major operand tile (tf32 needs the 32-byte-granule form of the swizzle there).
template <typename E>
static int make_tmap_2d(CUtensorMap* map, const void* base, uint64_t inner, uint64_t outer, uint64_t ld, uint32_t box_inner,
                        uint32_t box_outer, bool mn_major) {
    EncodeTiledFn fn = get_encode_fn();
    DMLB_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled entry point not available");
    DMLB_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "TMA base must be 16-byte aligned");
    DMLB_REQUIRE((ld * E::kBytes) % 16 == 0, "TMA row pitch must be a multiple of 16 bytes (ld % 8 == 0 for bf16, ld % 4 == 0 for fp32)");
    cuuint64_t dims[2] = {inner, outer};
    cuuint64_t strides[1] = {ld * E::kBytes};
    cuuint32_t box[2] = {box_inner, box_outer};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(map, tma_dtype<E>(), 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    mn_major ? E::kTmaMnSwizzle : CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        char buf[256];
        snprintf(buf, sizeof(buf), "cuTensorMapEncodeTiled failed: CUresult %d (inner=%llu outer=%llu ld=%llu box=%ux%u elem=%dB)", (int)r,
                 (unsigned long long)inner, (unsigned long long)outer, (unsigned long long)ld, box_inner, box_outer, E::kBytes);
        dmlb_set_error(buf);
        return -3;
    }
    return 0;
}

typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const int*,
                                   const int*, cuuint32_t, cuuint32_t, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeIm2colFn get_encode_im2col_fn() {
    static EncodeIm2colFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeIm2colFn>(p);
    }
    return fn;
}

// im2col tensor map over an NHWC activation [N, H, W, C] (channel pitch ld): boxes of `pixels` output pixels x kBK channels (128 B).
// Bounding box per spatial dim: [low, dim - 1 + up]; the instruction adds the filter-tap offset to every pixel.
// small_c: one 16-byte channel granule per pixel and tap, written UNSWIZZLED (core-matrix layout of the small-channel operand forms).
template <typename E>
static int make_tmap_im2col(CUtensorMap* map, const void* base, int C, int W, int H, int N, int ld, int low_w, int low_h, int up_w, int up_h,
                            int sw, int sh, int pixels, bool mn_major, bool small_c = false) {
    EncodeIm2colFn fn = get_encode_im2col_fn();
    DMLB_REQUIRE(fn != nullptr, "cuTensorMapEncodeIm2col entry point not available");
    DMLB_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0 && (ld * E::kBytes) % 16 == 0, "im2col TMA needs a 16-byte aligned NHWC tensor");
    const cuuint64_t eb = E::kBytes;
    cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
    cuuint64_t strides[3] = {(cuuint64_t)ld * eb, (cuuint64_t)W * ld * eb, (cuuint64_t)H * W * ld * eb};
    int lower[2] = {low_w, low_h};
    int upper[2] = {up_w, up_h};
    cuuint32_t trav[4] = {1, (cuuint32_t)sw, (cuuint32_t)sh, 1};
    const cuuint32_t box_c = small_c ? (cuuint32_t)(16 / E::kBytes) : (cuuint32_t)E::kBK;
    const CUtensorMapSwizzle swz = small_c ? CU_TENSOR_MAP_SWIZZLE_NONE : (mn_major ? E::kTmaMnSwizzle : CU_TENSOR_MAP_SWIZZLE_128B);
    CUresult r = fn(map, tma_dtype<E>(), 4, const_cast<void*>(base), dims, strides, lower, upper, box_c, (cuuint32_t)pixels, trav,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        char buf[256];
        snprintf(buf, sizeof(buf), "cuTensorMapEncodeIm2col failed: CUresult %d (C=%d W=%d H=%d N=%d ld=%d low=%d,%d up=%d,%d stride=%d,%d pixels=%d elem=%dB)",
                 (int)r, C, W, H, N, ld, low_w, low_h, up_w, up_h, sw, sh, pixels, E::kBytes);
        dmlb_set_error(buf);
        return -3;
    }
    // Same adjustment CUTLASS applies (cute/atom/copy_traits_sm90_im2col.hpp) for tensors smaller than 128 KiB on current drivers.
    int drv = 0;
    if (cudaDriverGetVersion(&drv) == cudaSuccess && drv <= 13010 && (unsigned long long)N * H * W * ld * eb < 131072ull)
        reinterpret_cast<uint64_t*>(map)[1] &= ~(1ull << 21);
    return 0;
}

// 3-D map over a conv weight [Cout][kh*kw][Cin] (row pitch ldrow elements per Cout): box = kBK Cin x 1 tap x kBK Cout, the MN-major
// B operand of dgrad.
template <typename E>
static int make_tmap_w3d(CUtensorMap* map, const void* base, int Cin, int KHW, int Cout, int ldrow) {
    EncodeTiledFn fn = get_encode_fn();
    DMLB_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled entry point not available");
    DMLB_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0 && (1ll * Cin * E::kBytes) % 16 == 0 && (1ll * ldrow * E::kBytes) % 16 == 0,
                 "weight map needs 16-byte aligned taps and rows");
    cuuint64_t dims[3] = {(cuuint64_t)Cin, (cuuint64_t)KHW, (cuuint64_t)Cout};
    cuuint64_t strides[2] = {(cuuint64_t)Cin * E::kBytes, (cuuint64_t)ldrow * E::kBytes};
    cuuint32_t box[3] = {(cuuint32_t)E::kBK, 1, (cuuint32_t)E::kBK};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = fn(map, tma_dtype<E>(), 3, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, E::kTmaMnSwizzle,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        char buf[200];
        snprintf(buf, sizeof(buf), "cuTensorMapEncodeTiled (3-D weight map) failed: CUresult %d (Cin=%d taps=%d Cout=%d ld=%d)", (int)r, Cin, KHW, Cout,
                 ldrow);
        dmlb_set_error(buf);
        return -3;
    }
    return 0;
}

template <typename E, int BN, int AM, int BMODE, bool kDual, bool kFin = false>
static int launch_gemm_v(const CUtensorMap& ta, const CUtensorMap& tb, const GemmEpilogue& ep, const ConvK& cv, int M, int N, int K, int splits,
                         cudaStream_t stream) {
    using S = GemmSmem<BN, kDual>;
    auto kern = gemm_tcgen05_kernel<E, BN, AM, BMODE, kDual, kFin>;
    // the opt-in to > 48 KB of dynamic shared memory is a per-device function attribute: remember it per device, not per process
    static bool attr_set[64] = {};
    int dev = 0;
    DMLB_CHECK_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        DMLB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, S::kTotal));
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    const int nkb = (K + E::kBK - 1) / E::kBK;
    int kb_per = (nkb + splits - 1) / splits;
    splits = (nkb + kb_per - 1) / kb_per;  // every z-slice owns >= 1 k-block
    dim3 grid((M + kBM - 1) / kBM, (N + BN - 1) / BN, splits);
    DMLB_CHECK_CUDA(launch_k(kern, grid, kGemmThreads, S::kTotal, stream, ta, tb, ep, cv, M, N, K, kb_per));
    DMLB_LAUNCH_CHECK();
    return 0;
}

template <int BN, int AM, int BMODE>
static int launch_gemm_persistent(const CUtensorMap& ta, const CUtensorMap& tb, const GemmEpilogue& ep, const ConvK& cv, int M, int N, int K,
                                  cudaStream_t stream) {
    using S = PersistSmem<BN>;
    auto kern = gemm_bf16_tcgen05_persistent_kernel<BN, AM, BMODE>;
    static bool attr_set[64] = {};
    int dev = 0;
    DMLB_CHECK_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        DMLB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, S::kTotal));
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    const int tiles_m = (M + kBM - 1) / kBM, tiles_n = (N + BN - 1) / BN;
    int grid = tiles_m * tiles_n;
    if (grid > dmlb_sm_count()) grid = dmlb_sm_count();
    DMLB_CHECK_CUDA(launch_k(kern, grid, kGemmThreads, S::kTotal, stream, ta, tb, ep, cv, M, N, K, tiles_m, tiles_n));
    DMLB_LAUNCH_CHECK();
    return 0;
}

extern "C" int dmlb_gemm_pair_enabled();
extern "C" int dmlb_gemm_persistent_enabled();

// 256 x 256 tiles on CTA pairs (see gemm_bf16_tcgen05_pair_kernel).  B's map has a 128-row box: each CTA loads half of the tile.
static int launch_gemm_pair(const CUtensorMap& ta, const CUtensorMap& tb_half, const GemmEpilogue& ep, int M, int N, int K, int splits,
                            cudaStream_t stream) {
    using S = PairSmem;
    auto kern = gemm_bf16_tcgen05_pair_kernel;
    static bool attr_set[64] = {};
    int dev = 0;
    DMLB_CHECK_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        DMLB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, S::kTotal));
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    const int nkb = (K + kBK - 1) / kBK;
    int kb_per = (nkb + splits - 1) / splits;
    splits = (nkb + kb_per - 1) / kb_per;
    const int tiles_m2 = (M + 2 * kBM - 1) / (2 * kBM), tiles_n = (N + S::kBN - 1) / S::kBN;
    const int clusters = dmlb_sm_count() / 2;
    if (splits == 1 && tiles_m2 * tiles_n >= 2 * clusters && dmlb_gemm_pair_enabled() != 2) {
        // at least two tiles per TPC: persistent form, epilogue of tile t overlaps the main loop of tile t+1
        auto pk = gemm_bf16_tcgen05_pair_persistent_kernel;
        static bool pattr[64] = {};
        if (dev < 0 || dev >= 64 || !pattr[dev]) {
            DMLB_CHECK_CUDA(cudaFuncSetAttribute(pk, cudaFuncAttributeMaxDynamicSharedMemorySize, S::kTotal));
            if (dev >= 0 && dev < 64) pattr[dev] = true;
        }
        cudaLaunchConfig_t pc = {};
        pc.gridDim = dim3(2 * clusters);
        pc.blockDim = dim3(kGemmThreads);
        pc.dynamicSmemBytes = S::kTotal;
        pc.stream = stream;
        static int group_m = -1;
        if (group_m < 0) {
            const char* e = getenv("DMLB_PAIR_GROUP_M");
            group_m = e != nullptr ? atoi(e) : 8;
            if (group_m < 1) group_m = 1 << 20;      // 0 = column-major over the whole matrix (no banding)
        }
        DMLB_CHECK_CUDA(cudaLaunchKernelEx(&pc, pk, ta, tb_half, ep, M, N, K, tiles_m2, tiles_n, group_m < tiles_m2 ? group_m : tiles_m2));
        DMLB_LAUNCH_CHECK();
        return 0;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * tiles_m2, tiles_n, splits);
    cfg.blockDim = dim3(kGemmThreads);
    cfg.dynamicSmemBytes = S::kTotal;
    cfg.stream = stream;
    DMLB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, ta, tb_half, ep, M, N, K, kb_per));   // cluster shape comes from __cluster_dims__
    DMLB_LAUNCH_CHECK();
    return 0;
}

// Cluster split-K: grid (tiles_m, tiles_n, S), the S K-slices of a tile form a cluster and reduce through distributed shared memory.
template <typename E, int BN, int AM, int BMODE>
static int launch_gemm_cluster(const CUtensorMap& ta, const CUtensorMap& tb, const GemmEpilogue& ep, const ConvK& cv, int M, int N, int K, int S,
                               cudaStream_t stream) {
    if constexpr (BN > 128) {
        dmlb_set_error("cluster split-K is instantiated for tiles up to 128 columns");
        return -4;
    } else {
        using Sm = GemmSmem<BN, false>;
        auto kern = gemm_tcgen05_kernel<E, BN, AM, BMODE, false, false, true>;
        const int nkb = (K + E::kBK - 1) / E::kBK;
        DMLB_REQUIRE(S == 2 || S == 4 || S == 8, "cluster split-K: 2, 4 or 8 K-slices per tile");
        DMLB_REQUIRE(S <= nkb, "cluster split-K: more slices than k-blocks");
        DMLB_REQUIRE(BN % S == 0 && (BN / S) % 4 == 0 && BN / S <= 32, "cluster split-K: each slice finalises 4..32 columns (tile_n / slices)");
        DMLB_REQUIRE(!ep.atomic && ep.perm_cin == 0 && ep.cpad == 0, "cluster split-K writes final values (no accumulating epilogue)");
        DMLB_REQUIRE(128 * BN * 4 <= Sm::kStages * Sm::kStageBytes, "exchange area does not fit the operand stages");
        static bool attr_set[64] = {};
        int dev = 0;
        DMLB_CHECK_CUDA(cudaGetDevice(&dev));
        if (dev < 0 || dev >= 64 || !attr_set[dev]) {
            DMLB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Sm::kTotal));
            if (dev >= 0 && dev < 64) attr_set[dev] = true;
        }
        dim3 grid((M + kBM - 1) / kBM, (N + BN - 1) / BN, S);
        DMLB_CHECK_CUDA(launch_k_cluster(kern, grid, kGemmThreads, Sm::kTotal, stream, (unsigned)S, ta, tb, ep, cv, M, N, K, 0));
        DMLB_LAUNCH_CHECK();
        return 0;
    }
}

// Several waves of tiles and no split-K -> the persistent kernel (bf16); more than one wave -> two CTAs per SM; otherwise the deep ring.
template <typename E, int BN, int AM, int BMODE>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const GemmEpilogue& ep, const ConvK& cv, int M, int N, int K, int splits,
                       cudaStream_t stream) {
    if constexpr ((AM == 0 || AM == 2 || AM == 4) && (BMODE == 0 || BMODE == 3)) {
        if (ep.cluster_s > 1) return launch_gemm_cluster<E, BN, AM, BMODE>(ta, tb, ep, cv, M, N, K, ep.cluster_s, stream);
    } else {
        if (ep.cluster_s > 1) {
            dmlb_set_error("cluster split-K is instantiated for the forward / data-gradient operand forms");
            return -4;
        }
    }
    const int nkb = (K + E::kBK - 1) / E::kBK;
    const int kb_per = (nkb + splits - 1) / splits;
    const long long ctas = 1ll * ((M + kBM - 1) / kBM) * ((N + BN - 1) / BN) * ((nkb + kb_per - 1) / kb_per);
    if constexpr (!E::kTf32 && BN <= 128 && BMODE == 0 && AM != 3) {
        // forward form only (both operands K-major): the backward GEMMs share the GPU with the side-stream wgrad branch and a grid
        // that pins every SM would serialise against it.  Short reductions only: with a long K loop the per-tile set-up is already amortised and the 2-CTA variant rasterises better
        if (splits == 1 && !ep.use_fin && nkb <= 24 && ctas >= 3ll * dmlb_sm_count() && dmlb_gemm_persistent_enabled())
            return launch_gemm_persistent<BN, AM, BMODE>(ta, tb, ep, cv, M, N, K, stream);
    }
    if (ep.use_fin) {
        // folded split-K finish: forward / dgrad operand forms only, small grids by construction -> deep-ring variant (bf16 outputs)
        if constexpr (!E::kTf32 && (AM == 0 || AM == 2) && (BMODE == 0 || BMODE == 3) && BN <= 128)
            return launch_gemm_v<E, BN, AM, BMODE, false, true>(ta, tb, ep, cv, M, N, K, splits, stream);
        dmlb_set_error("fused split-K finish is instantiated for bf16, K-major / im2col A with K-major / weight-map B and tiles up to 128 wide");
        return -4;
    }
    if (BN <= 128 && ctas > dmlb_sm_count()) return launch_gemm_v<E, BN, AM, BMODE, true>(ta, tb, ep, cv, M, N, K, splits, stream);
    return launch_gemm_v<E, BN, AM, BMODE, false>(ta, tb, ep, cv, M, N, K, splits, stream);
}

// Arguments of the folded finishing pass (GemmEpilogue::fin_*), validated here.
struct FinishArgs {
    int* counters;
    int n_counters;
    const float* bias;
    int relu;
    void* y;
    int ldy;
    int pool_h, pool_w;   // 0, 0 = no pooling
    void* pooled;
    int ldp;
};

static int apply_finish(GemmEpilogue& ep, const FinishArgs& f, int M, int N, int BN) {
    DMLB_REQUIRE(ep.atomic && ep.out_is_f32 && ep.bias == nullptr && ep.perm_cin == 0, "fused finish needs a plain accumulating fp32 workspace");
    DMLB_REQUIRE(ep.ldo % 4 == 0 && ep.ldo >= N && (reinterpret_cast<uintptr_t>(ep.out) & 15) == 0, "workspace pitch must be a multiple of 4 covering N, base 16-byte aligned");
    DMLB_REQUIRE(f.counters != nullptr && f.y != nullptr && f.ldy % 4 == 0 && (reinterpret_cast<uintptr_t>(f.y) & 7) == 0, "fused finish: bad output");
    const long long tiles = 1ll * ((M + kBM - 1) / kBM) * ((N + BN - 1) / BN);
    DMLB_REQUIRE(tiles <= f.n_counters, "fused finish: more output tiles than arrival counters");
    ep.use_fin = 1;
    ep.fin_counters = f.counters; ep.fin_y = f.y; ep.fin_bias = f.bias; ep.fin_ldy = f.ldy; ep.fin_relu = f.relu;
    ep.fin_pooled = nullptr; ep.fin_ldp = 0; ep.fin_pool_w = 0;
    if (f.pool_w > 0) {
        DMLB_REQUIRE(f.pool_h % 2 == 0 && f.pool_w % 2 == 0 && kBM % (2 * f.pool_w) == 0 && M % (f.pool_h * f.pool_w) == 0,
                     "fused 2x2 pool needs an even [H,W] map whose row pairs tile 128 rows (W = 2..64, power of two)");
        DMLB_REQUIRE(f.pooled != nullptr && f.ldp % 4 == 0 && (reinterpret_cast<uintptr_t>(f.pooled) & 7) == 0, "fused pool: bad pooled output");
        ep.fin_pooled = f.pooled; ep.fin_ldp = f.ldp; ep.fin_pool_w = f.pool_w;
    }
    return 0;
}

// See include/dmlb.h for the contract (dmlb_gemm_bf16 / dmlb_gemm_tf32).
template <typename E>
static int gemm_impl(const FinishArgs* fin, const void* A, int lda, int a_mn_major, const void* B, int ldb, int b_mn_major, int M, int N, int K,
                     void* out, int out_is_f32, int ldo, const float* bias, int relu, int split_k, int atomic, int perm_cin, int perm_khw,
                     int tile_n, void* stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    constexpr int BK = E::kBK;
    DMLB_REQUIRE(M > 0 && N > 0 && K > 0, "empty GEMM");
    if (split_k < 1) split_k = 1;
    if (split_k > 1) atomic = 1;
    DMLB_REQUIRE(!atomic || out_is_f32, "atomic / split-K output must be fp32");
    DMLB_REQUIRE(!(atomic && relu), "ReLU cannot be fused into an accumulating epilogue");
    DMLB_REQUIRE(!a_mn_major || b_mn_major, "A MN-major with B K-major is not instantiated");
    DMLB_REQUIRE(!E::kTf32 || out_is_f32, "the tf32 GEMM writes fp32");

    int BN = tile_n;
    if (BN == 0) BN = (N > 64) ? 128 : (N > 32 ? 64 : 32);
    if (b_mn_major && BN < 64) BN = 64;
    DMLB_REQUIRE(BN == 32 || BN == 64 || BN == 128 || BN == 256, "tile_n must be 32, 64, 128 or 256");

    CUtensorMap ta, tb;
    int rc;
    if (!a_mn_major) rc = make_tmap_2d<E>(&ta, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda, BK, kBM, false);
    else             rc = make_tmap_2d<E>(&ta, A, (uint64_t)M, (uint64_t)K, (uint64_t)lda, BK, BK, true);
    if (rc) return rc;
    // 256-wide tiles of a plain K-major bf16 GEMM with at least one full pair of row tiles: CTA pairs (cta_group::2), half a B tile per CTA
    const bool pair = !E::kTf32 && BN == 256 && !a_mn_major && !b_mn_major && fin == nullptr && M > kBM && dmlb_gemm_pair_enabled();
    if (!b_mn_major) rc = make_tmap_2d<E>(&tb, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb, BK, (uint32_t)(pair ? BN / 2 : BN), false);
    else             rc = make_tmap_2d<E>(&tb, B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb, BK, BK, true);
    if (rc) return rc;

    GemmEpilogue ep = {};
    ep.out = out; ep.bias = bias; ep.ldo = ldo; ep.out_is_f32 = out_is_f32; ep.relu = relu; ep.atomic = atomic;
    ep.perm_cin = perm_cin; ep.perm_khw = perm_khw;
    if (fin != nullptr) {
        rc = apply_finish(ep, *fin, M, N, BN);
        if (rc) return rc;
    }

    if constexpr (!E::kTf32) {
        if (pair) return launch_gemm_pair(ta, tb, ep, M, N, K, split_k, stream);
    }
    ConvK cv = {};
#define DMLB_GEMM_CASE(bn, amn, bmn) \
    if (BN == bn && (bool)a_mn_major == amn && (bool)b_mn_major == bmn) return launch_gemm<E, bn, amn ? 1 : 0, bmn ? 1 : 0>(ta, tb, ep, cv, M, N, K, split_k, stream);
    DMLB_GEMM_CASE(32, false, false)
    DMLB_GEMM_CASE(64, false, false)
    DMLB_GEMM_CASE(128, false, false)
    if constexpr (!E::kTf32) { DMLB_GEMM_CASE(256, false, false) }
    DMLB_GEMM_CASE(64, false, true)
    DMLB_GEMM_CASE(128, false, true)
    DMLB_GEMM_CASE(64, true, true)
    DMLB_GEMM_CASE(128, true, true)
#undef DMLB_GEMM_CASE
    dmlb_set_error("dmlb_gemm: no kernel instantiated for this (tile_n, major) combination");
    return -4;
}

// Implicit-GEMM convolution (no column matrix in memory): the activation operand is gathered by TMA in im2col mode.
//   kind 0  fprop : out[B*Ho*Wo, Cout] = conv(x, W)           act = x [B,Ha,Wa,Ca=Cin],   other = W [Cout, kh*kw*Cin]
//   kind 1  wgrad : out[Cout, kh*kw*Cin] (+)= dY^T * im2col(x) act = x,                     other = dY [B*Ho*Wo, ld_other]
//   kind 2  dgrad : out[B*Ha'*Wa', Cin]  = conv^T(dY, W)      act = dY [B,Ha,Wa,Ca=Cout],  other = W; stride 1 only;
//                   n_out = Cin;  (Ha, Wa) are the conv OUTPUT extents, the result has the conv INPUT extents.
// Needs Ca % kBK == 0 (a 128-byte k-block must stay inside one filter tap: 64 bf16 / 32 fp32 channels); dgrad also needs 16-byte taps.
// Epilogue arguments as in dmlb_gemm_*.  Reference: nn.Conv2d forward/backward in example/models.py:29-39 (cuDNN there).
// `extra` (may be null): cluster_s / pooled / pool_w / act_* fields to copy into the epilogue (cluster split-K form).
template <typename E>
static int conv_implicit_impl(const FinishArgs* fin, int kind, const void* act, int ld_act, int B, int Ha, int Wa, int Ca, const void* other, int ld_other,
                              int KH, int KW, int SH, int SW, int PH, int PW, int n_out, void* out, int out_is_f32, int ldo,
                              const float* bias, int relu, int split_k, int atomic, int tile_n, void* stream_, const GemmEpilogue* extra = nullptr) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    constexpr int BK = E::kBK;
    DMLB_REQUIRE(kind >= 0 && kind <= 2, "kind must be 0 (fprop), 1 (wgrad) or 2 (dgrad)");
    DMLB_REQUIRE(Ca % BK == 0, "implicit-GEMM convolution needs the gathered tensor's channels to fill whole 128-byte k-blocks (64 bf16 / 32 fp32)");
    if (split_k < 1) split_k = 1;
    if (split_k > 1) atomic = 1;
    DMLB_REQUIRE(!atomic || out_is_f32, "atomic / split-K output must be fp32");
    DMLB_REQUIRE(!(atomic && relu), "ReLU cannot be fused into an accumulating epilogue");
    DMLB_REQUIRE(!E::kTf32 || out_is_f32, "the tf32 convolution writes fp32");
    GemmEpilogue ep = {};
    ep.out = out; ep.bias = bias; ep.ldo = ldo; ep.out_is_f32 = out_is_f32; ep.relu = relu; ep.atomic = atomic; ep.perm_cin = 0; ep.perm_khw = 0;
    if (extra != nullptr) {
        DMLB_REQUIRE(kind != 1 && !atomic && split_k == 1, "the cluster form replaces split-K for fprop / dgrad");
        ep.cluster_s = extra->cluster_s; ep.pooled = extra->pooled; ep.ldp = extra->ldp; ep.pool_w = extra->pool_w;
        ep.act_y = extra->act_y; ep.act_dy = extra->act_dy; ep.act_dbias = extra->act_dbias; ep.act_ldy = extra->act_ldy;
        ep.act_pool2 = extra->act_pool2; ep.act_relu = extra->act_relu; ep.act_hx = extra->act_hx; ep.act_wx = extra->act_wx;
        if (ep.pool_w > 0)
            DMLB_REQUIRE((ep.pool_w & (ep.pool_w - 1)) == 0 && ep.pool_w >= 2 && ep.pool_w <= 16 && ep.pooled != nullptr,
                         "fused 2x2 pool needs a power-of-two map width <= 16");
    }
    ConvK cv = {};
    cv.C = Ca; cv.KH = KH; cv.KW = KW; cv.sh = SH; cv.sw = SW;
    CUtensorMap ta, tb;
    int rc;
    if (kind == 0 || kind == 1) {
        const int Ho = (Ha + 2 * PH - KH) / SH + 1, Wo = (Wa + 2 * PW - KW) / SW + 1;
        DMLB_REQUIRE(Ho > 0 && Wo > 0, "empty conv output");
        cv.P = Ho; cv.Q = Wo; cv.low_h = -PH; cv.low_w = -PW; cv.flip = 0;
        const int Kd = KH * KW * Ca;
        const long long Mpix = 1ll * B * Ho * Wo;
        DMLB_REQUIRE(Mpix < (1ll << 31), "too many output pixels");
        if (kind == 0) {
            const int N = n_out;
            int BN = tile_n ? tile_n : ((N > 64) ? 128 : (N > 32 ? 64 : 32));
            rc = make_tmap_im2col<E>(&ta, act, Ca, Wa, Ha, B, ld_act, -PW, -PH, PW - (KW - 1), PH - (KH - 1), SW, SH, kBM, false);
            if (rc) return rc;
            rc = make_tmap_2d<E>(&tb, other, (uint64_t)Kd, (uint64_t)N, (uint64_t)ld_other, BK, (uint32_t)BN, false);
            if (rc) return rc;
            if (fin != nullptr && (rc = apply_finish(ep, *fin, (int)Mpix, N, BN)) != 0) return rc;
            if (BN == 32) return launch_gemm<E, 32, 2, 0>(ta, tb, ep, cv, (int)Mpix, N, Kd, split_k, stream);
            if (BN == 64) return launch_gemm<E, 64, 2, 0>(ta, tb, ep, cv, (int)Mpix, N, Kd, split_k, stream);
            if (BN == 128) return launch_gemm<E, 128, 2, 0>(ta, tb, ep, cv, (int)Mpix, N, Kd, split_k, stream);
            if constexpr (!E::kTf32) {
                if (BN == 256) return launch_gemm<E, 256, 2, 0>(ta, tb, ep, cv, (int)Mpix, N, Kd, split_k, stream);
            }
        } else {
            const int Cout = n_out;
            int BN = tile_n ? tile_n : 128;
            DMLB_REQUIRE(fin == nullptr, "weight gradients accumulate into the arena; there is nothing to finish");
            rc = make_tmap_2d<E>(&ta, other, (uint64_t)Cout, (uint64_t)Mpix, (uint64_t)ld_other, BK, BK, true);   // dY as MN-major A
            if (rc) return rc;
            rc = make_tmap_im2col<E>(&tb, act, Ca, Wa, Ha, B, ld_act, -PW, -PH, PW - (KW - 1), PH - (KH - 1), SW, SH, BK, true);
            if (rc) return rc;
            if (BN == 64) return launch_gemm<E, 64, 1, 2>(ta, tb, ep, cv, Cout, Kd, (int)Mpix, split_k, stream);
            if (BN == 128) return launch_gemm<E, 128, 1, 2>(ta, tb, ep, cv, Cout, Kd, (int)Mpix, split_k, stream);
        }
    } else {
        DMLB_REQUIRE(SH == 1 && SW == 1, "implicit dgrad supports stride 1 (strided convolutions use the explicit path)");
        const int Cin = n_out, Cout = Ca;
        DMLB_REQUIRE((Cin * E::kBytes) % 16 == 0, "implicit dgrad needs 16-byte filter taps (Cin % 8 == 0 for bf16, % 4 for fp32)");
        const int Hin = Ha + KH - 1 - 2 * PH, Win = Wa + KW - 1 - 2 * PW;     // conv input extents (stride 1)
        DMLB_REQUIRE(Hin > 0 && Win > 0, "bad dgrad geometry");
        cv.P = Hin; cv.Q = Win; cv.low_h = -(KH - 1 - PH); cv.low_w = -(KW - 1 - PW); cv.flip = 1;
        const int Kd = KH * KW * Cout;
        const long long Mpix = 1ll * B * Hin * Win;
        DMLB_REQUIRE(Mpix < (1ll << 31), "too many pixels");
        int BN = tile_n ? tile_n : (Cin > 64 ? 128 : 64);
        rc = make_tmap_im2col<E>(&ta, act, Cout, Wa, Ha, B, ld_act, cv.low_w, cv.low_h, -PW, -PH, 1, 1, kBM, false);
        if (rc) return rc;
        rc = make_tmap_w3d<E>(&tb, other, Cin, KH * KW, Cout, ld_other);
        if (rc) return rc;
        if (fin != nullptr && (rc = apply_finish(ep, *fin, (int)Mpix, Cin, BN)) != 0) return rc;
        if (BN == 64) return launch_gemm<E, 64, 2, 3>(ta, tb, ep, cv, (int)Mpix, Cin, Kd, split_k, stream);
        if (BN == 128) return launch_gemm<E, 128, 2, 3>(ta, tb, ep, cv, (int)Mpix, Cin, Kd, split_k, stream);
    }
    dmlb_set_error("dmlb_conv_implicit: unsupported tile_n for this kind");
    return -4;
}

// R consecutive output pixels (in (n, ho, wo) order, starting at a multiple of R) as a box {qb, pb, nb} of the (wo, ho, n) index space.
static bool pixel_box(int R, int Q, int P, int* qb, int* pb, int* nb) {
    if (Q >= R) {
        if (Q % R) return false;
        *qb = R; *pb = 1; *nb = 1;
        return true;
    }
    if (R % Q) return false;
    const int a = R / Q;
    if (a <= P) {
        if (P % a) return false;
        *qb = Q; *pb = a; *nb = 1;
        return true;
    }
    if (a % P) return false;
    *qb = Q; *pb = P; *nb = a / P;
    return *nb <= 256;
}

// Overlapping-window tensor map over a zero-padded granule buffer [B, Hp, Wp, granule]: dim 0 walks the bytes of one image (so a
// coordinate (kh*Wp + kw0)*granule selects the filter tap), dims 1-3 are wo / ho / n with strides of `sw` pixels / `sh` rows / one image.
template <typename E>
static int make_tmap_rowwin(CUtensorMap* map, const void* base, int B, int Hp, int Wp, int P, int Q, int sh, int sw, int R, bool mn_major) {
    EncodeTiledFn fn = get_encode_fn();
    DMLB_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled entry point not available");
    DMLB_REQUIRE((reinterpret_cast<uintptr_t>(base) & 127) == 0, "the padded input buffer must be 128-byte aligned");
    constexpr int G = 16 / E::kBytes;
    int qb, pb, nb;
    DMLB_REQUIRE(pixel_box(R, Q, P, &qb, &pb, &nb), "output extents do not tile into boxes of consecutive pixels");
    cuuint64_t dims[4] = {(cuuint64_t)Hp * Wp * G, (cuuint64_t)Q, (cuuint64_t)P, (cuuint64_t)B};
    cuuint64_t strides[3] = {(cuuint64_t)sw * 16, (cuuint64_t)sh * Wp * 16, (cuuint64_t)Hp * Wp * 16};
    cuuint32_t box[4] = {(cuuint32_t)(8 * G), (cuuint32_t)qb, (cuuint32_t)pb, (cuuint32_t)nb};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = fn(map, tma_dtype<E>(), 4, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    mn_major ? E::kTmaMnSwizzle : CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        char buf[256];
        snprintf(buf, sizeof(buf), "cuTensorMapEncodeTiled (row-window map) failed: CUresult %d (B=%d Hp=%d Wp=%d P=%d Q=%d stride=%d,%d box=%d,%d,%d)",
                 (int)r, B, Hp, Wp, P, Q, sh, sw, qb, pb, nb);
        dmlb_set_error(buf);
        return -3;
    }
    return 0;
}

// Implicit-GEMM convolution for SMALL channel counts (the first layer of every model here: Cin = 3).  The activation is channels-last with
// exactly one 16-byte granule per pixel (4 fp32 / 8 bf16 channels, the unused ones zero); a k-block gathers eight filter taps, so the
// reduction is kh*kw*granule long instead of kh*kw*64 — see the A = 3 / B = 4 operand forms of the kernel.
//   kind 0  fprop: out[B*Ho*Wo, Cout] = conv(x, W);   other = W as [Cout, kh*kw*granule] (pad channels zero), K-major
//                  pool_w > 0 additionally writes maxpool2x2(out) to pooled (fused in the epilogue; needs Wo | 16, Ho even)
//   kind 1  wgrad: out[Cout, kh*kw*cin] (+)= dY^T * im2col(x);  other = dY [B*Ho*Wo, ld_other]; pad-channel columns are dropped;
//                  pt_layout = 1 writes PyTorch's [Cout, Cin, kh, kw] order instead of [Cout, kh, kw, Cin]
// Reference: the first nn.Conv2d of example/models.py:8,29 (3 -> 6 5x5 and 3 -> 64 11x11 stride 4).
// Row-window variant (hp > 0): act is the zero-PADDED buffer [B, hp, wp, granule] (image at rows PH.., columns PW..), the filter width is
// padded to kwpad = 8 * ceil(KW / 8) in the weight layout [n_out, KH * kwpad * granule], and one tiled TMA load fetches a whole k-block.
template <typename E>
static int conv_smallc_impl(int kind, const void* act, int B, int Ha, int Wa, int cin, const void* other, int ld_other, int KH, int KW, int SH,
                            int SW, int PH, int PW, int n_out, void* out, int out_is_f32, int ldo, const float* bias, int relu, int split_k,
                            int pt_layout, void* pooled, int ldp, int pool_w, int tile_n, int hp, int wp, void* stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    constexpr int BK = E::kBK;
    constexpr int G = 16 / E::kBytes;            // channels per granule = channel pitch of the activation
    DMLB_REQUIRE(kind == 0 || kind == 1, "kind must be 0 (fprop) or 1 (wgrad)");
    DMLB_REQUIRE(cin >= 1 && cin <= G, "small-channel convolution: Cin must fit one 16-byte granule (4 fp32 / 8 bf16 channels)");
    DMLB_REQUIRE(!E::kTf32 || out_is_f32, "the tf32 convolution writes fp32");
    const int Ho = (Ha + 2 * PH - KH) / SH + 1, Wo = (Wa + 2 * PW - KW) / SW + 1;
    DMLB_REQUIRE(Ho > 0 && Wo > 0, "empty conv output");
    const bool rowwin = hp > 0;
    const int kwb = (KW + 7) / 8;                // k-blocks per filter row of the row-window form
    const int ntaps = rowwin ? KH * kwb * 8 : KH * KW;
    const int Kg = ntaps * G;                    // GEMM reduction length of fprop = column count of wgrad
    if (rowwin) {
        DMLB_REQUIRE(hp >= (Ho - 1) * SH + KH && wp >= (Wo - 1) * SW + kwb * 8 && hp >= Ha + 2 * PH && wp >= Wa + 2 * PW,
                     "padded input buffer too small for the row windows");
    }
    const long long Mpix = 1ll * B * Ho * Wo;
    DMLB_REQUIRE(Mpix < (1ll << 31), "too many output pixels");
    if (split_k < 1) split_k = 1;
    GemmEpilogue ep = {};
    ConvK cv = {};
    cv.C = G; cv.KH = KH; cv.KW = KW; cv.sh = SH; cv.sw = SW; cv.P = Ho; cv.Q = Wo; cv.low_h = -PH; cv.low_w = -PW; cv.flip = 0;
    cv.wp = wp; cv.kwb = kwb;
    CUtensorMap ta, tb;
    int rc;
    if (kind == 0) {
        const int N = n_out;
        int BN = tile_n ? tile_n : ((N > 64) ? 128 : (N > 32 ? 64 : 32));
        ep.out = out; ep.bias = bias; ep.ldo = ldo; ep.out_is_f32 = out_is_f32; ep.relu = relu;
        if (split_k > 1) {
            // few row tiles and a long reduction (AlexNet conv1: 32 tiles x 22 k-blocks): spread K over the SMs into an fp32 workspace
            DMLB_REQUIRE(out_is_f32 && bias == nullptr && !relu && pool_w == 0, "split-K small-channel fprop accumulates raw sums into an fp32 workspace");
            ep.atomic = 1;
        }
        if (pool_w > 0) {
            DMLB_REQUIRE(pool_w == Wo && (pool_w & (pool_w - 1)) == 0 && pool_w >= 2 && pool_w <= 16 && Ho % 2 == 0 && pooled != nullptr,
                         "fused 2x2 pool needs a power-of-two output width <= 16 and an even height");
            ep.pooled = pooled; ep.ldp = ldp; ep.pool_w = pool_w;
        }
        rc = make_tmap_2d<E>(&tb, other, (uint64_t)Kg, (uint64_t)N, (uint64_t)ld_other, BK, (uint32_t)BN, false);
        if (rc) return rc;
        if (rowwin) {
            rc = make_tmap_rowwin<E>(&ta, act, B, hp, wp, Ho, Wo, SH, SW, kBM, false);
            if (rc) return rc;
            if (BN == 32) return launch_gemm<E, 32, 4, 0>(ta, tb, ep, cv, (int)Mpix, N, Kg, split_k, stream);
            if (BN == 64) return launch_gemm<E, 64, 4, 0>(ta, tb, ep, cv, (int)Mpix, N, Kg, split_k, stream);
            if (BN == 128) return launch_gemm<E, 128, 4, 0>(ta, tb, ep, cv, (int)Mpix, N, Kg, split_k, stream);
            dmlb_set_error("dmlb_conv_smallc: unsupported tile_n");
            return -4;
        }
        rc = make_tmap_im2col<E>(&ta, act, G, Wa, Ha, B, G, -PW, -PH, PW - (KW - 1), PH - (KH - 1), SW, SH, kBM, false, true);
        if (rc) return rc;
        if (BN == 32) return launch_gemm<E, 32, 3, 0>(ta, tb, ep, cv, (int)Mpix, N, Kg, split_k, stream);
        if (BN == 64) return launch_gemm<E, 64, 3, 0>(ta, tb, ep, cv, (int)Mpix, N, Kg, split_k, stream);
        if (BN == 128) return launch_gemm<E, 128, 3, 0>(ta, tb, ep, cv, (int)Mpix, N, Kg, split_k, stream);
    } else {
        const int Cout = n_out;
        int BN = tile_n ? tile_n : 128;
        DMLB_REQUIRE(out_is_f32, "weight gradients accumulate in fp32");
        ep.out = out; ep.ldo = ldo; ep.out_is_f32 = 1; ep.atomic = 1; ep.cpad = G; ep.cin = cin; ep.perm_khw = pt_layout ? KH * KW : 0;
        rc = make_tmap_2d<E>(&ta, other, (uint64_t)Cout, (uint64_t)Mpix, (uint64_t)ld_other, BK, BK, true);   // dY as MN-major A
        if (rc) return rc;
        if (rowwin) {
            ep.kw = KW; ep.kwpad = kwb * 8;
            rc = make_tmap_rowwin<E>(&tb, act, B, hp, wp, Ho, Wo, SH, SW, BK, true);
            if (rc) return rc;
            if (BN == 64) return launch_gemm<E, 64, 1, 5>(ta, tb, ep, cv, Cout, Kg, (int)Mpix, split_k, stream);
            if (BN == 128) return launch_gemm<E, 128, 1, 5>(ta, tb, ep, cv, Cout, Kg, (int)Mpix, split_k, stream);
            dmlb_set_error("dmlb_conv_smallc: unsupported tile_n");
            return -4;
        }
        DMLB_REQUIRE(!E::kTf32, "the granule form has no tf32 weight-gradient kernel (32-bit MN-major operands exist only in the 32-byte-granule "
                                "swizzle): use the row-window form or the explicit column matrix");
        rc = make_tmap_im2col<E>(&tb, act, G, Wa, Ha, B, G, -PW, -PH, PW - (KW - 1), PH - (KH - 1), SW, SH, BK, true, true);
        if (rc) return rc;
        if (BN == 64) return launch_gemm<E, 64, 1, 4>(ta, tb, ep, cv, Cout, Kg, (int)Mpix, split_k, stream);
        if (BN == 128) return launch_gemm<E, 128, 1, 4>(ta, tb, ep, cv, Cout, Kg, (int)Mpix, split_k, stream);
    }
    dmlb_set_error("dmlb_conv_smallc: unsupported tile_n for this kind");
    return -4;
}

}  // namespace dmlb

// BatchNorm2d (training + inference), residual add + ReLU fused into the normalisation pass, and global average
// pooling — the extra ops of the NORTH-STAR ResNet-18 / ResNet-50 configurations (BASELINE.json configs[2,3]; the
// reference's own models, example/models.py, have no normalisation layers — SURVEY.md §8 row K13).
//
// Activations are bf16 channels-last rows [M = B*H*W, ld] (see conv_ops.cu); statistics and affine parameters are
// fp32.  Fast paths need C % 8 == 0 and C <= 2048 (every ResNet layer); other shapes take the scalar kernels.
//
//   forward :  bn_stats   sums[c] = sum_m x, sums[C+c] = sum_m x^2            (block partials -> red.add)
//              bn_apply   y = act((x - mean) * rstd * gamma + beta [+ residual]); saves mean/rstd, updates running stats
//   backward:  bn_bwd_reduce   g = dy * (y > 0 if relu);  dsum[c] = sum g, dsum[C+c] = sum g * xhat;  writes g (bf16)
//              bn_bwd_apply    dx = gamma * rstd * (g - dsum[c]/M - xhat * dsum[C+c]/M);  dgamma/dbeta += dsum
//   The masked gradient g is also the gradient of the residual input, so no separate add/ReLU backward kernels exist.
#include "common.cuh"

namespace dmlb {

__device__ __forceinline__ void unpack8f(const uint4& v, float (&f)[8]) {
    const float2 a = unpack_bf16x2(v.x), b = unpack_bf16x2(v.y), c = unpack_bf16x2(v.z), d = unpack_bf16x2(v.w);
    f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
__device__ __forceinline__ uint4 pack8f(const float (&f)[8]) {
    uint4 o;
    o.x = pack_bf16x2(f[0], f[1]); o.y = pack_bf16x2(f[2], f[3]); o.z = pack_bf16x2(f[4], f[5]); o.w = pack_bf16x2(f[6], f[7]);
    return o;
}

// Block layout shared by the reductions: tx = channel group (8 channels), ty = row lane.
// blockIdx.y selects a slab of up to 32 channel groups (256 channels), so wide layers keep several row lanes per block and
// each block ends with at most 512 global reductions.
struct RowTiling {
    int c8n;    // channel groups handled by this block (<= 32)
    int lanes;  // row lanes per block (c8n * lanes <= blockDim.x)
    int c8_0;   // first channel group of this block's slab
};
__device__ __forceinline__ RowTiling row_tiling(int C) {
    RowTiling t;
    const int total = C >> 3;
    const int per = total < 32 ? total : 32;
    t.c8_0 = blockIdx.y * per;
    t.c8n = min(per, total - t.c8_0);
    t.lanes = blockDim.x / per;
    return t;
}

// sums[0..C) += sum over rows of a, sums[C..2C) += sum over rows of a*b'   (generic two-moment reduction)
//   mode 0 (forward stats):  a = x,  second = x * x
//   mode 1 (backward):       a = g = dy * mask,  second = g * xhat;  optionally stores g
// The block's partial sums go through shared memory WITHOUT atomics: every thread parks its 16 register sums in its own slot, then each
// of the slab's 2 x 8 x c8n columns is summed over the row lanes by one thread and leaves the block as one red.add.  (The first version
// used shared-memory atomics: 4096 per block with up to 32 threads on one address, which cost more than streaming the block's rows.)
constexpr int kBnPartFloats = 256 * 16;
template <int mode, int kU>
__global__ void __launch_bounds__(256) bn_reduce_kernel(const __nv_bfloat16* __restrict__ x, int ldx, const __nv_bfloat16* __restrict__ dy, int lddy,
                                                        const __nv_bfloat16* __restrict__ yact, int ldya, __nv_bfloat16* __restrict__ gout, int ldg,
                                                        const float* __restrict__ mean_rstd, float* __restrict__ sums, long long M, int C,
                                                        int relu) {
    pdl_prologue();
    extern __shared__ __align__(16) float part[];  // [256 threads][16]: s1[8], s2[8] of each thread
    const RowTiling t = row_tiling(C);
    const int per = (C >> 3) < 32 ? (C >> 3) : 32;
    const int txl = threadIdx.x % per, ty = threadIdx.x / per;
    const int tx = t.c8_0 + txl;                       // global channel group
    const bool live = ty < t.lanes && txl < t.c8n;
    float s1[8] = {0, 0, 0, 0, 0, 0, 0, 0}, s2[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    float mu[8], rs[8];
    if (mode == 1 && live) {
#pragma unroll
        for (int j = 0; j < 8; ++j) { mu[j] = mean_rstd[tx * 8 + j]; rs[j] = mean_rstd[C + tx * 8 + j]; }
    }
    if (live) {
        // kU rows per trip, every load issued before the first use: the kernel is a pure stream and lives on loads in flight
        const long long rstride = 1ll * gridDim.x * t.lanes;
        for (long long r0 = blockIdx.x * 1ll * t.lanes + ty; r0 < M; r0 += rstride * kU) {
            uint4 rx[kU], rd[kU], ry[kU];
#pragma unroll
            for (int u = 0; u < kU; ++u) {
                const long long r = r0 + u * rstride;
                if (r < M) {
                    rx[u] = __ldg(reinterpret_cast<const uint4*>(x + r * ldx + tx * 8));
                    if (mode == 1) {
                        rd[u] = __ldg(reinterpret_cast<const uint4*>(dy + r * lddy + tx * 8));
                        if (relu) ry[u] = __ldg(reinterpret_cast<const uint4*>(yact + r * ldya + tx * 8));
                    }
                }
            }
#pragma unroll
            for (int u = 0; u < kU; ++u) {
                const long long r = r0 + u * rstride;
                if (r >= M) continue;
                float xv[8];
                unpack8f(rx[u], xv);
                if (mode == 0) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) { s1[j] += xv[j]; s2[j] += xv[j] * xv[j]; }
                } else {
                    float g[8];
                    unpack8f(rd[u], g);
                    if (relu) {
                        float ya[8];
                        unpack8f(ry[u], ya);
#pragma unroll
                        for (int j = 0; j < 8; ++j)
                            if (!(ya[j] > 0.f)) g[j] = 0.f;
                    }
                    if (gout != nullptr) *reinterpret_cast<uint4*>(gout + r * ldg + tx * 8) = pack8f(g);
#pragma unroll
                    for (int j = 0; j < 8; ++j) { s1[j] += g[j]; s2[j] += g[j] * (xv[j] - mu[j]) * rs[j]; }
                }
            }
        }
        float4* slot = reinterpret_cast<float4*>(part + (ty * per + txl) * 16);
        slot[0] = make_float4(s1[0], s1[1], s1[2], s1[3]);
        slot[1] = make_float4(s1[4], s1[5], s1[6], s1[7]);
        slot[2] = make_float4(s2[0], s2[1], s2[2], s2[3]);
        slot[3] = make_float4(s2[4], s2[5], s2[6], s2[7]);
    }
    __syncthreads();
    for (int col = threadIdx.x; col < t.c8n * 16; col += blockDim.x) {
        float acc = 0.f;
        for (int l = 0; l < t.lanes; ++l) acc += part[l * per * 16 + col];
        const int g8 = col >> 4, k = col & 15;
        red_add_f32(sums + (k < 8 ? 0 : C) + (t.c8_0 + g8) * 8 + (k & 7), acc);
    }
}

// Forward normalisation.  stats: if `sums` != null the batch statistics are derived from it (training) and written to
// save_mean_rstd / folded into the running statistics by block 0; otherwise mean_rstd_in (inference: running mean and
// 1/sqrt(running_var + eps), prepared by the caller) is used as is.
__global__ void __launch_bounds__(256) bn_apply_kernel(const __nv_bfloat16* __restrict__ x, int ldx, __nv_bfloat16* __restrict__ y, int ldy,
                                                       const __nv_bfloat16* __restrict__ res, int ldr, const float* __restrict__ sums,
                                                       const float* __restrict__ mean_rstd_in, const float* __restrict__ gamma,
                                                       const float* __restrict__ beta, float* __restrict__ save_mean_rstd,
                                                       float* __restrict__ running_mean, float* __restrict__ running_var, long long M, int C,
                                                       float eps, float momentum, int relu, float* __restrict__ zero_2c) {
    pdl_prologue();
    const RowTiling t = row_tiling(C);
    const int per = (C >> 3) < 32 ? (C >> 3) : 32;
    const int txl = threadIdx.x % per, ty = threadIdx.x / per;
    if (ty >= t.lanes || txl >= t.c8n) return;
    const int tx = t.c8_0 + txl;
    // The first trip's rows are requested BEFORE the per-channel constants are loaded and combined (~30 dependent scalar loads per thread):
    // with 16 rows per thread that prologue would otherwise be a quarter of the kernel.
    constexpr int kU = 8;
    const long long rstride = 1ll * gridDim.x * t.lanes;
    long long r0 = blockIdx.x * 1ll * t.lanes + ty;
    uint4 rx[kU], rr[kU];
    auto request = [&](long long base) {
#pragma unroll
        for (int u = 0; u < kU; ++u) {
            const long long r = base + u * rstride;
            if (r < M) {
                rx[u] = __ldg(reinterpret_cast<const uint4*>(x + r * ldx + tx * 8));
                if (res != nullptr) rr[u] = __ldg(reinterpret_cast<const uint4*>(res + r * ldr + tx * 8));
            }
        }
    };
    if (r0 < M) request(r0);
    float sc[8], sh[8];
    const float inv_m = 1.f / static_cast<float>(M);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int c = tx * 8 + j;
        float mean, rstd;
        if (sums != nullptr) {
            mean = sums[c] * inv_m;
            const float var = fmaxf(sums[C + c] * inv_m - mean * mean, 0.f);
            rstd = rsqrtf(var + eps);
            if (blockIdx.x == 0 && ty == 0) {
                if (zero_2c != nullptr) { zero_2c[c] = 0.f; zero_2c[C + c] = 0.f; }   // leaves the backward scratch clean (no memset launch)
                save_mean_rstd[c] = mean;
                save_mean_rstd[C + c] = rstd;
                if (running_mean != nullptr) {
                    const float unbiased = M > 1 ? var * static_cast<float>(M) / static_cast<float>(M - 1) : var;
                    running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
                    running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
                }
            }
        } else {
            mean = mean_rstd_in[c];
            rstd = mean_rstd_in[C + c];
        }
        const float g = gamma != nullptr ? gamma[c] : 1.f;
        sc[j] = g * rstd;
        sh[j] = (beta != nullptr ? beta[c] : 0.f) - mean * g * rstd;
    }
    while (r0 < M) {
#pragma unroll
        for (int u = 0; u < kU; ++u) {
            const long long r = r0 + u * rstride;
            if (r >= M) continue;
            float v[8];
            unpack8f(rx[u], v);
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = v[j] * sc[j] + sh[j];
            if (res != nullptr) {
                float q[8];
                unpack8f(rr[u], q);
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] += q[j];
            }
            if (relu) {
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] = fmaxf(v[j], 0.f);
            }
            *reinterpret_cast<uint4*>(y + r * ldy + tx * 8) = pack8f(v);
        }
        r0 += rstride * kU;
        if (r0 < M) request(r0);
    }
}

__global__ void __launch_bounds__(256) bn_bwd_apply_kernel(const __nv_bfloat16* __restrict__ x, int ldx, const __nv_bfloat16* __restrict__ g,
                                                           int ldg, __nv_bfloat16* __restrict__ dx, int lddx, const float* __restrict__ mean_rstd,
                                                           const float* __restrict__ gamma, const float* __restrict__ dsum,
                                                           float* __restrict__ dgamma, float* __restrict__ dbeta, long long M, int C,
                                                           float* __restrict__ zero_2c) {
    pdl_prologue();
    const RowTiling t = row_tiling(C);
    const int per = (C >> 3) < 32 ? (C >> 3) : 32;
    const int txl = threadIdx.x % per, ty = threadIdx.x / per;
    if (ty >= t.lanes || txl >= t.c8n) return;
    const int tx = t.c8_0 + txl;
    constexpr int kU = 4;
    const long long rstride = 1ll * gridDim.x * t.lanes;
    long long r0 = blockIdx.x * 1ll * t.lanes + ty;
    uint4 rx[kU], rg[kU];
    auto request = [&](long long base) {       // see bn_apply_kernel: rows first, constants second
#pragma unroll
        for (int u = 0; u < kU; ++u) {
            const long long r = base + u * rstride;
            if (r < M) {
                rx[u] = __ldg(reinterpret_cast<const uint4*>(x + r * ldx + tx * 8));
                rg[u] = __ldg(reinterpret_cast<const uint4*>(g + r * ldg + tx * 8));
            }
        }
    };
    if (r0 < M) request(r0);
    float mu[8], rs[8], k0[8], k1[8], k2[8];
    const float inv_m = 1.f / static_cast<float>(M);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int c = tx * 8 + j;
        mu[j] = mean_rstd[c];
        rs[j] = mean_rstd[C + c];
        const float gm = gamma != nullptr ? gamma[c] : 1.f;
        k0[j] = gm * rs[j];
        k1[j] = dsum[c] * inv_m;
        k2[j] = dsum[C + c] * inv_m;
        if (blockIdx.x == 0 && ty == 0) {
            if (zero_2c != nullptr) { zero_2c[c] = 0.f; zero_2c[C + c] = 0.f; }       // leaves the forward statistics scratch clean
            if (dgamma != nullptr) dgamma[c] += dsum[C + c];
            if (dbeta != nullptr) dbeta[c] += dsum[c];
        }
    }
    while (r0 < M) {
#pragma unroll
        for (int u = 0; u < kU; ++u) {
            const long long r = r0 + u * rstride;
            if (r >= M) continue;
            float xv[8], gv[8];
            unpack8f(rx[u], xv);
            unpack8f(rg[u], gv);
#pragma unroll
            for (int j = 0; j < 8; ++j) gv[j] = k0[j] * (gv[j] - k1[j] - (xv[j] - mu[j]) * rs[j] * k2[j]);
            *reinterpret_cast<uint4*>(dx + r * lddx + tx * 8) = pack8f(gv);
        }
        r0 += rstride * kU;
        if (r0 < M) request(r0);
    }
}

// Global average pooling over H*W pixels per sample: y[b, c] = mean_p x[b, p, c]; backward spreads dy / P.
__global__ void __launch_bounds__(256) avgpool_fwd_kernel(const __nv_bfloat16* __restrict__ x, int ldx, __nv_bfloat16* __restrict__ y, int ldy, int B,
                                                          int P, int C) {
    pdl_prologue();
    const int c8n = C >> 3;
    const int total = B * c8n;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int c8 = i % c8n, b = i / c8n;
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int p = 0; p < P; ++p) {
            float v[8];
            unpack8f(__ldg(reinterpret_cast<const uint4*>(x + (1ll * b * P + p) * ldx + c8 * 8)), v);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] += v[j];
        }
        const float inv = 1.f / static_cast<float>(P);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] *= inv;
        *reinterpret_cast<uint4*>(y + 1ll * b * ldy + c8 * 8) = pack8f(acc);
    }
}
__global__ void __launch_bounds__(256) avgpool_bwd_kernel(const __nv_bfloat16* __restrict__ dy, int lddy, __nv_bfloat16* __restrict__ dx, int lddx,
                                                          int B, int P, int C) {
    pdl_prologue();
    const int c8n = C >> 3;
    const long long total = 1ll * B * P * c8n;
    const float inv = 1.f / static_cast<float>(P);
    for (long long i = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i < total; i += 1ll * gridDim.x * blockDim.x) {
        const int c8 = static_cast<int>(i % c8n);
        const long long bp = i / c8n;
        const int b = static_cast<int>(bp / P);
        float v[8];
        unpack8f(__ldg(reinterpret_cast<const uint4*>(dy + 1ll * b * lddy + c8 * 8)), v);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] *= inv;
        *reinterpret_cast<uint4*>(dx + bp * lddx + c8 * 8) = pack8f(v);
    }
}

static inline dim3 rows_grid(long long M, int C) {
    const int total = C / 8;
    const int per = total < 32 ? total : 32;
    const int slabs = (total + per - 1) / per;
    const int lanes = 256 / per;
    long long gx = (M + 1ll * lanes * 16 - 1) / (1ll * lanes * 16);     // >= 16 rows per thread before the block reduces
    const long long cap = (148ll * 8 + slabs - 1) / slabs;
    if (gx > cap) gx = cap;
    if (gx < 1) gx = 1;
    return dim3(static_cast<unsigned>(gx), static_cast<unsigned>(slabs), 1);
}

}  // namespace dmlb

using namespace dmlb;

#define BN_SHAPE_OK(C) ((C) % 8 == 0 && (C) >= 8 && (C) <= 2048)

// sums: fp32 [2C], zero on entry.
extern "C" int dmlb_bn_stats(const void* x, int ldx, float* sums, long long M, int C, void* stream) {
    DMLB_REQUIRE(BN_SHAPE_OK(C) && ldx % 8 == 0, "batch-norm kernels need C % 8 == 0, C <= 2048 and 16-byte rows");
    DMLB_CHECK_CUDA(launch_k(bn_reduce_kernel<0, 8>, rows_grid(M, C), 256, kBnPartFloats * sizeof(float), static_cast<cudaStream_t>(stream),
                             static_cast<const __nv_bfloat16*>(x), ldx, static_cast<const __nv_bfloat16*>(nullptr), 0,
                             static_cast<const __nv_bfloat16*>(nullptr), 0, static_cast<__nv_bfloat16*>(nullptr), 0,
                             static_cast<const float*>(nullptr), sums, M, C, 0));
    DMLB_LAUNCH_CHECK();
    return 0;
}

extern "C" int dmlb_bn_apply(const void* x, int ldx, void* y, int ldy, const void* res, int ldr, const float* sums, const float* mean_rstd_in,
                             const float* gamma, const float* beta, float* save_mean_rstd, float* running_mean, float* running_var, long long M,
                             int C, float eps, float momentum, int relu, float* zero_2c, void* stream) {
    DMLB_REQUIRE(BN_SHAPE_OK(C) && ldx % 8 == 0 && ldy % 8 == 0, "batch-norm kernels need C % 8 == 0, C <= 2048 and 16-byte rows");
    DMLB_REQUIRE(sums != nullptr || mean_rstd_in != nullptr, "either batch sums or prepared statistics are required");
    DMLB_CHECK_CUDA(launch_k(bn_apply_kernel, rows_grid(M, C), 256, 0, static_cast<cudaStream_t>(stream), static_cast<const __nv_bfloat16*>(x), ldx,
                             static_cast<__nv_bfloat16*>(y), ldy, static_cast<const __nv_bfloat16*>(res), ldr, sums, mean_rstd_in, gamma, beta,
                             save_mean_rstd, running_mean, running_var, M, C, eps, momentum, relu, zero_2c));
    DMLB_LAUNCH_CHECK();
    return 0;
}

// dsum: fp32 [2C], zero on entry.  g_out (bf16 rows, pitch ldg) receives dy * relu-mask; it doubles as the residual gradient.
extern "C" int dmlb_bn_bwd_reduce(const void* x, int ldx, const void* dy, int lddy, const void* yact, int ldya, void* g_out, int ldg,
                                  const float* mean_rstd, float* dsum, long long M, int C, int relu, void* stream) {
    DMLB_REQUIRE(BN_SHAPE_OK(C) && ldx % 8 == 0 && lddy % 8 == 0, "batch-norm kernels need C % 8 == 0, C <= 2048 and 16-byte rows");
    DMLB_CHECK_CUDA(launch_k(bn_reduce_kernel<1, 4>, rows_grid(M, C), 256, kBnPartFloats * sizeof(float), static_cast<cudaStream_t>(stream),
                             static_cast<const __nv_bfloat16*>(x), ldx, static_cast<const __nv_bfloat16*>(dy), lddy,
                             static_cast<const __nv_bfloat16*>(yact), ldya, static_cast<__nv_bfloat16*>(g_out), ldg, mean_rstd, dsum, M, C, relu));
    DMLB_LAUNCH_CHECK();
    return 0;
}

extern "C" int dmlb_bn_bwd_apply(const void* x, int ldx, const void* g, int ldg, void* dx, int lddx, const float* mean_rstd, const float* gamma,
                                 const float* dsum, float* dgamma, float* dbeta, long long M, int C, float* zero_2c, void* stream) {
    DMLB_REQUIRE(BN_SHAPE_OK(C) && ldx % 8 == 0 && ldg % 8 == 0 && lddx % 8 == 0, "batch-norm kernels need C % 8 == 0, C <= 2048 and 16-byte rows");
    DMLB_CHECK_CUDA(launch_k(bn_bwd_apply_kernel, rows_grid(M, C), 256, 0, static_cast<cudaStream_t>(stream), static_cast<const __nv_bfloat16*>(x),
                             ldx, static_cast<const __nv_bfloat16*>(g), ldg, static_cast<__nv_bfloat16*>(dx), lddx, mean_rstd, gamma, dsum, dgamma,
                             dbeta, M, C, zero_2c));
    DMLB_LAUNCH_CHECK();
    return 0;
}

extern "C" int dmlb_avgpool_fwd_bf16(const void* x, int ldx, void* y, int ldy, int B, int P, int C, void* stream) {
    DMLB_REQUIRE(C % 8 == 0 && ldx % 8 == 0 && ldy % 8 == 0, "average pooling needs C % 8 == 0");
    const int total = B * (C / 8);
    DMLB_CHECK_CUDA(launch_k(avgpool_fwd_kernel, (total + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream), static_cast<const __nv_bfloat16*>(x),
                             ldx, static_cast<__nv_bfloat16*>(y), ldy, B, P, C));
    DMLB_LAUNCH_CHECK();
    return 0;
}

extern "C" int dmlb_avgpool_bwd_bf16(const void* dy, int lddy, void* dx, int lddx, int B, int P, int C, void* stream) {
    DMLB_REQUIRE(C % 8 == 0 && lddy % 8 == 0 && lddx % 8 == 0, "average pooling needs C % 8 == 0");
    const long long total = 1ll * B * P * (C / 8);
    long long blocks = (total + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    DMLB_CHECK_CUDA(launch_k(avgpool_bwd_kernel, static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream),
                             static_cast<const __nv_bfloat16*>(dy), lddy, static_cast<__nv_bfloat16*>(dx), lddx, B, P, C));
    DMLB_LAUNCH_CHECK();
    return 0;
}

// Data-movement and activation kernels around the tcgen05 GEMM for Conv2d / MaxPool2d / ReLU.
//
// Activations are bf16, channels-last: a logical [B, C, H, W] tensor is stored as [B, H, W, ldc] with
// ldc >= C (ldc % 8 == 0 so every row is a legal TMA / 16-byte-vector operand).  Column matrices are
// [M = B*Ho*Wo, Kpad] with K ordered (kh, kw, cin).
//
// Reference call sites replaced: nn.Conv2d / nn.MaxPool2d / nn.ReLU in example/models.py:16-17,28-41
// (cuDNN + ATen there).
#include "common.cuh"

namespace dmlb {

struct ConvGeom {
    int B, H, W, C;      // input
    int KH, KW, SH, SW, PH, PW;
    int Ho, Wo;
    int ldx;             // input channel pitch (NHWC) — unused for NCHW fp32 input
    int Kpad;            // column-matrix pitch
};

// ---------------------------------------------------------------------------------------------
// im2col, bf16 NHWC input, 8 channels (16 B) per thread.  Requires C % 8 == 0.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) im2col_nhwc_vec8_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ col, ConvGeom g) {
    pdl_prologue();
    const int c8n = g.C >> 3;
    const long long total = 1ll * g.B * g.Ho * g.Wo * g.KH * g.KW * c8n;
    for (long long i = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i < total; i += 1ll * gridDim.x * blockDim.x) {
        const int c8 = static_cast<int>(i % c8n);
        long long t = i / c8n;
        const int kwi = static_cast<int>(t % g.KW); t /= g.KW;
        const int khi = static_cast<int>(t % g.KH); t /= g.KH;
        const int wo = static_cast<int>(t % g.Wo); t /= g.Wo;
        const int ho = static_cast<int>(t % g.Ho);
        const int b = static_cast<int>(t / g.Ho);
        const int hi = ho * g.SH - g.PH + khi;
        const int wi = wo * g.SW - g.PW + kwi;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (hi >= 0 && hi < g.H && wi >= 0 && wi < g.W)
            v = __ldg(reinterpret_cast<const uint4*>(x + ((1ll * b * g.H + hi) * g.W + wi) * g.ldx + c8 * 8));
        const long long m = (1ll * b * g.Ho + ho) * g.Wo + wo;
        *reinterpret_cast<uint4*>(col + m * g.Kpad + (khi * g.KW + kwi) * g.C + c8 * 8) = v;
    }
}

// im2col, bf16 NHWC input, any C (one element per thread).
__global__ void __launch_bounds__(256) im2col_nhwc_scalar_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ col, ConvGeom g) {
    pdl_prologue();
    const int K = g.KH * g.KW * g.C;
    const long long total = 1ll * g.B * g.Ho * g.Wo * K;
    for (long long i = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i < total; i += 1ll * gridDim.x * blockDim.x) {
        const int k = static_cast<int>(i % K);
        const long long m = i / K;
        const int c = k % g.C;
        const int kwi = (k / g.C) % g.KW;
        const int khi = k / (g.C * g.KW);
        const int wo = static_cast<int>(m % g.Wo);
        const int ho = static_cast<int>((m / g.Wo) % g.Ho);
        const int b = static_cast<int>(m / (1ll * g.Wo * g.Ho));
        const int hi = ho * g.SH - g.PH + khi;
        const int wi = wo * g.SW - g.PW + kwi;
        __nv_bfloat16 v = __float2bfloat16(0.f);
        if (hi >= 0 && hi < g.H && wi >= 0 && wi < g.W) v = x[((1ll * b * g.H + hi) * g.W + wi) * g.ldx + c];
        col[m * g.Kpad + k] = v;
    }
}

// im2col straight from the user's fp32 NCHW batch (first layer): fuses the fp32->bf16 cast and the
// NCHW->NHWC re-layout into the gather (reference main.py:64-70 feeds such a batch to conv1).
__global__ void __launch_bounds__(256) im2col_nchw_f32_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ col, ConvGeom g) {
    pdl_prologue();
    // one thread per (output pixel, filter tap): the index arithmetic is paid once for all C channels, neighbouring
    // threads read neighbouring input columns (coalesced) and write neighbouring C-element groups of the same row
    const int taps = g.KH * g.KW;
    const long long total = 1ll * g.B * g.Ho * g.Wo * taps;
    const long long plane = 1ll * g.H * g.W;
    for (long long i = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i < total; i += 1ll * gridDim.x * blockDim.x) {
        const int tap = static_cast<int>(i % taps);
        const long long m = i / taps;
        const int kwi = tap % g.KW;
        const int khi = tap / g.KW;
        const int wo = static_cast<int>(m % g.Wo);
        const int ho = static_cast<int>((m / g.Wo) % g.Ho);
        const int b = static_cast<int>(m / (1ll * g.Wo * g.Ho));
        const int hi = ho * g.SH - g.PH + khi;
        const int wi = wo * g.SW - g.PW + kwi;
        const bool ok = hi >= 0 && hi < g.H && wi >= 0 && wi < g.W;
        const float* src = x + (1ll * b * g.C * g.H + hi) * g.W + wi;
        __nv_bfloat16* dst = col + m * g.Kpad + tap * g.C;
        for (int c = 0; c < g.C; ++c) dst[c] = __float2bfloat16_rn(ok ? __ldg(src + c * plane) : 0.f);
    }
}

// ---------------------------------------------------------------------------------------------
// col2im (gather form, no atomics): dx[b,h,w,c] = sum over taps of dcol[m(b,ho,wo), (kh,kw,c)].
// ---------------------------------------------------------------------------------------------
template <int VEC>
__global__ void __launch_bounds__(256) col2im_nhwc_kernel(const __nv_bfloat16* __restrict__ dcol, __nv_bfloat16* __restrict__ dx, ConvGeom g) {
    pdl_prologue();
    const int cvn = (g.C + VEC - 1) / VEC;
    const long long total = 1ll * g.B * g.H * g.W * cvn;
    for (long long i = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i < total; i += 1ll * gridDim.x * blockDim.x) {
        const int cv = static_cast<int>(i % cvn);
        long long t = i / cvn;
        const int w = static_cast<int>(t % g.W); t /= g.W;
        const int h = static_cast<int>(t % g.H);
        const int b = static_cast<int>(t / g.H);
        float acc[VEC];
#pragma unroll
        for (int j = 0; j < VEC; ++j) acc[j] = 0.f;
        for (int khi = 0; khi < g.KH; ++khi) {
            const int hn = h + g.PH - khi;
            if (hn < 0 || hn % g.SH) continue;
            const int ho = hn / g.SH;
            if (ho >= g.Ho) continue;
            for (int kwi = 0; kwi < g.KW; ++kwi) {
                const int wn = w + g.PW - kwi;
                if (wn < 0 || wn % g.SW) continue;
                const int wo = wn / g.SW;
                if (wo >= g.Wo) continue;
                const long long m = (1ll * b * g.Ho + ho) * g.Wo + wo;
                const __nv_bfloat16* src = dcol + m * g.Kpad + (khi * g.KW + kwi) * g.C + cv * VEC;
                if constexpr (VEC == 8) {
                    const uint4 v = __ldg(reinterpret_cast<const uint4*>(src));
                    const float2 a = unpack_bf16x2(v.x), bq = unpack_bf16x2(v.y), c = unpack_bf16x2(v.z), d = unpack_bf16x2(v.w);
                    acc[0] += a.x; acc[1] += a.y; acc[2] += bq.x; acc[3] += bq.y;
                    acc[4] += c.x; acc[5] += c.y; acc[6] += d.x; acc[7] += d.y;
                } else {
                    acc[0] += __bfloat162float(src[0]);
                }
            }
        }
        __nv_bfloat16* dst = dx + ((1ll * b * g.H + h) * g.W + w) * g.ldx + cv * VEC;
        if constexpr (VEC == 8) {
            uint4 o;
            o.x = pack_bf16x2(acc[0], acc[1]); o.y = pack_bf16x2(acc[2], acc[3]);
            o.z = pack_bf16x2(acc[4], acc[5]); o.w = pack_bf16x2(acc[6], acc[7]);
            *reinterpret_cast<uint4*>(dst) = o;
        } else {
            dst[0] = __float2bfloat16_rn(acc[0]);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// MaxPool2d forward (NHWC bf16), window k x k, stride s, padding p (implicit -inf).
// ---------------------------------------------------------------------------------------------
struct PoolGeom {
    int B, H, W, C, ldx;   // input
    int Ho, Wo, ldo;       // output
    int k, s, p;
};

template <int VEC>
__global__ void __launch_bounds__(256) maxpool_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, PoolGeom g) {
    pdl_prologue();
    const int cvn = (g.C + VEC - 1) / VEC;
    const long long total = 1ll * g.B * g.Ho * g.Wo * cvn;
    for (long long i = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i < total; i += 1ll * gridDim.x * blockDim.x) {
        const int cv = static_cast<int>(i % cvn);
        long long t = i / cvn;
        const int wo = static_cast<int>(t % g.Wo); t /= g.Wo;
        const int ho = static_cast<int>(t % g.Ho);
        const int b = static_cast<int>(t / g.Ho);
        float mx[VEC];
#pragma unroll
        for (int j = 0; j < VEC; ++j) mx[j] = -INFINITY;
        for (int dh = 0; dh < g.k; ++dh) {
            const int h = ho * g.s - g.p + dh;
            if (h < 0 || h >= g.H) continue;
            for (int dw = 0; dw < g.k; ++dw) {
                const int w = wo * g.s - g.p + dw;
                if (w < 0 || w >= g.W) continue;
                const __nv_bfloat16* src = x + ((1ll * b * g.H + h) * g.W + w) * g.ldx + cv * VEC;
                if constexpr (VEC == 8) {
                    const uint4 v = __ldg(reinterpret_cast<const uint4*>(src));
                    const float2 a = unpack_bf16x2(v.x), bq = unpack_bf16x2(v.y), c = unpack_bf16x2(v.z), d = unpack_bf16x2(v.w);
                    mx[0] = fmaxf(mx[0], a.x); mx[1] = fmaxf(mx[1], a.y); mx[2] = fmaxf(mx[2], bq.x); mx[3] = fmaxf(mx[3], bq.y);
                    mx[4] = fmaxf(mx[4], c.x); mx[5] = fmaxf(mx[5], c.y); mx[6] = fmaxf(mx[6], d.x); mx[7] = fmaxf(mx[7], d.y);
                } else {
                    mx[0] = fmaxf(mx[0], __bfloat162float(src[0]));
                }
            }
        }
        __nv_bfloat16* dst = y + ((1ll * b * g.Ho + ho) * g.Wo + wo) * g.ldo + cv * VEC;
        if constexpr (VEC == 8) {
            uint4 o;
            o.x = pack_bf16x2(mx[0], mx[1]); o.y = pack_bf16x2(mx[2], mx[3]);
            o.z = pack_bf16x2(mx[4], mx[5]); o.w = pack_bf16x2(mx[6], mx[7]);
            *reinterpret_cast<uint4*>(dst) = o;
        } else {
            dst[0] = __float2bfloat16_rn(mx[0]);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Fused activation backward: (optional max-pool backward) o (optional ReLU mask) + bias gradient.
//   y    : forward post-activation conv output [B,H,W,ldy]   (the pool's input when pooling)
//   dout : gradient w.r.t. the block output: pooled [B,Ho,Wo,ldd] when has_pool else [B,H,W,ldd]
//   dy   : gradient w.r.t. the pre-activation conv output [B,H,W,ldy]  (GEMM operand for wgrad/dgrad)
//   dbias: fp32 [C], accumulated with red.add (may be null)
// Pool backward recomputes the window arg-max from y with PyTorch's "first maximum wins" rule, so no
// index tensor is stored in the forward pass.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) act_bwd_kernel(const __nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ dout,
                                                      __nv_bfloat16* __restrict__ dy, float* __restrict__ dbias, PoolGeom g, int ldd,
                                                      int has_pool, int relu) {
    pdl_prologue();
    extern __shared__ float sbias[];  // [C]
    for (int c = threadIdx.x; c < g.C; c += blockDim.x) sbias[c] = 0.f;
    __syncthreads();
    const long long total = 1ll * g.B * g.H * g.W * g.C;
    for (long long i = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i < total; i += 1ll * gridDim.x * blockDim.x) {
        const int c = static_cast<int>(i % g.C);
        long long t = i / g.C;
        const int w = static_cast<int>(t % g.W); t /= g.W;
        const int h = static_cast<int>(t % g.H);
        const int b = static_cast<int>(t / g.H);
        const long long xi = ((1ll * b * g.H + h) * g.W + w) * g.ldx + c;
        const float yv = __bfloat162float(y[xi]);
        float gsum = 0.f;
        if (!has_pool) {
            gsum = __bfloat162float(dout[((1ll * b * g.H + h) * g.W + w) * ldd + c]);
        } else {
            // windows (ho, wo) containing (h, w): ho*s - p <= h < ho*s - p + k
            const int ho_hi = min(g.Ho - 1, (h + g.p) / g.s);
            const int wo_hi = min(g.Wo - 1, (w + g.p) / g.s);
            for (int ho = ho_hi; ho >= 0 && ho * g.s - g.p + g.k > h; --ho)
                for (int wo = wo_hi; wo >= 0 && wo * g.s - g.p + g.k > w; --wo) {
                    // is (h, w) the first maximum of this window?
                    bool winner = true;
                    for (int dh = 0; dh < g.k && winner; ++dh) {
                        const int hh = ho * g.s - g.p + dh;
                        if (hh < 0 || hh >= g.H) continue;
                        for (int dw = 0; dw < g.k; ++dw) {
                            const int ww = wo * g.s - g.p + dw;
                            if (ww < 0 || ww >= g.W) continue;
                            if (hh == h && ww == w) continue;
                            const float o = __bfloat162float(y[((1ll * b * g.H + hh) * g.W + ww) * g.ldx + c]);
                            const bool before = (hh < h) || (hh == h && ww < w);
                            if (o > yv || (before && o == yv)) { winner = false; break; }
                        }
                    }
                    if (winner) gsum += __bfloat162float(dout[((1ll * b * g.Ho + ho) * g.Wo + wo) * ldd + c]);
                }
        }
        if (relu && !(yv > 0.f)) gsum = 0.f;
        dy[xi] = __float2bfloat16_rn(gsum);
        if (dbias != nullptr && gsum != 0.f) atomicAdd(&sbias[c], gsum);
    }
    if (dbias != nullptr) {
        __syncthreads();
        for (int c = threadIdx.x; c < g.C; c += blockDim.x)
            if (sbias[c] != 0.f) red_add_f32(dbias + c, sbias[c]);
    }
}

// Vectorised variant (8 channels = 16 bytes per thread) for the common cases: no pooling, or the 2x2 / stride-2 pool of
// AlexNet and LeNet.  Bias-gradient partial sums stay in registers while a thread keeps visiting the same channel group.
__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
    const float2 a = unpack_bf16x2(v.x), b = unpack_bf16x2(v.y), c = unpack_bf16x2(v.z), d = unpack_bf16x2(v.w);
    f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}

__global__ void __launch_bounds__(256) act_bwd_vec8_kernel(const __nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ dout,
                                                           __nv_bfloat16* __restrict__ dy, float* __restrict__ dbias, PoolGeom g, int ldd,
                                                           int has_pool, int relu) {
    pdl_prologue();
    extern __shared__ float sbias[];  // [C]
    for (int c = threadIdx.x; c < g.C; c += blockDim.x) sbias[c] = 0.f;
    __syncthreads();
    const int c8n = g.C >> 3;
    const long long total = 1ll * g.B * g.H * g.W * c8n;
    float bacc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    int cur_c8 = -1;
    for (long long i = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i < total; i += 1ll * gridDim.x * blockDim.x) {
        const int c8 = static_cast<int>(i % c8n);
        long long t = i / c8n;
        const int w = static_cast<int>(t % g.W); t /= g.W;
        const int h = static_cast<int>(t % g.H);
        const int b = static_cast<int>(t / g.H);
        const long long xi = ((1ll * b * g.H + h) * g.W + w) * g.ldx + c8 * 8;
        float yv[8], gv[8];
        unpack8(__ldg(reinterpret_cast<const uint4*>(y + xi)), yv);
        if (!has_pool) {
            unpack8(__ldg(reinterpret_cast<const uint4*>(dout + ((1ll * b * g.H + h) * g.W + w) * ldd + c8 * 8)), gv);
        } else {
            const int hp = h >> 1, wp = w >> 1;
            if (hp >= g.Ho || wp >= g.Wo) {
#pragma unroll
                for (int j = 0; j < 8; ++j) gv[j] = 0.f;
            } else {
                unpack8(__ldg(reinterpret_cast<const uint4*>(dout + ((1ll * b * g.Ho + hp) * g.Wo + wp) * ldd + c8 * 8)), gv);
                const int self = ((h & 1) << 1) | (w & 1);   // scan position inside the window
#pragma unroll
                for (int o = 0; o < 4; ++o) {
                    if (o == self) continue;
                    float ov[8];
                    const long long oi = ((1ll * b * g.H + (2 * hp + (o >> 1))) * g.W + (2 * wp + (o & 1))) * g.ldx + c8 * 8;
                    unpack8(__ldg(reinterpret_cast<const uint4*>(y + oi)), ov);
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        if (ov[j] > yv[j] || (o < self && ov[j] == yv[j])) gv[j] = 0.f;   // PyTorch: first maximum wins
                }
            }
        }
        if (relu) {
#pragma unroll
            for (int j = 0; j < 8; ++j)
                if (!(yv[j] > 0.f)) gv[j] = 0.f;
        }
        uint4 o;
        o.x = pack_bf16x2(gv[0], gv[1]); o.y = pack_bf16x2(gv[2], gv[3]);
        o.z = pack_bf16x2(gv[4], gv[5]); o.w = pack_bf16x2(gv[6], gv[7]);
        *reinterpret_cast<uint4*>(dy + xi) = o;
        if (dbias != nullptr) {
            if (c8 != cur_c8) {
                if (cur_c8 >= 0) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        if (bacc[j] != 0.f) atomicAdd(&sbias[cur_c8 * 8 + j], bacc[j]);
                        bacc[j] = 0.f;
                    }
                }
                cur_c8 = c8;
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) bacc[j] += gv[j];
        }
    }
    if (dbias != nullptr) {
        if (cur_c8 >= 0) {
#pragma unroll
            for (int j = 0; j < 8; ++j)
                if (bacc[j] != 0.f) atomicAdd(&sbias[cur_c8 * 8 + j], bacc[j]);
        }
        __syncthreads();
        for (int c = threadIdx.x; c < g.C; c += blockDim.x)
            if (sbias[c] != 0.f) red_add_f32(dbias + c, sbias[c]);
    }
}

// Standalone ReLU (bf16, padded rows): y = max(x, 0); backward dx = dy * (y > 0).
__global__ void __launch_bounds__(256) relu_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, long long rows,
                                                       int C, int ldx, int ldy) {
    pdl_prologue();
    const long long total = rows * C;
    for (long long i = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i < total; i += 1ll * gridDim.x * blockDim.x) {
        const long long r = i / C;
        const int c = static_cast<int>(i % C);
        y[r * ldy + c] = __float2bfloat16_rn(fmaxf(__bfloat162float(x[r * ldx + c]), 0.f));
    }
}
__global__ void __launch_bounds__(256) relu_bwd_kernel(const __nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ dy,
                                                       __nv_bfloat16* __restrict__ dx, long long rows, int C, int ldy, int lddy, int lddx) {
    pdl_prologue();
    const long long total = rows * C;
    for (long long i = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i < total; i += 1ll * gridDim.x * blockDim.x) {
        const long long r = i / C;
        const int c = static_cast<int>(i % C);
        const float yv = __bfloat162float(y[r * ldy + c]);
        dx[r * lddx + c] = yv > 0.f ? dy[r * lddy + c] : __float2bfloat16(0.f);
    }
}

static inline int grid_for(long long total, int threads = 256) {
    long long blocks = (total + threads - 1) / threads;
    const long long cap = 148ll * 16;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return static_cast<int>(blocks);
}

}  // namespace dmlb

using namespace dmlb;

static ConvGeom make_geom(int B, int H, int W, int C, int KH, int KW, int SH, int SW, int PH, int PW, int ldx, int Kpad) {
    ConvGeom g;
    g.B = B; g.H = H; g.W = W; g.C = C; g.KH = KH; g.KW = KW; g.SH = SH; g.SW = SW; g.PH = PH; g.PW = PW;
    g.Ho = (H + 2 * PH - KH) / SH + 1;
    g.Wo = (W + 2 * PW - KW) / SW + 1;
    g.ldx = ldx; g.Kpad = Kpad;
    return g;
}

extern "C" int dmlb_im2col_nhwc_bf16(const void* x, void* col, int B, int H, int W, int C, int ldx, int KH, int KW, int SH, int SW, int PH,
                                     int PW, int Kpad, void* stream) {
    ConvGeom g = make_geom(B, H, W, C, KH, KW, SH, SW, PH, PW, ldx, Kpad);
    DMLB_REQUIRE(g.Ho > 0 && g.Wo > 0, "empty conv output");
    DMLB_REQUIRE(Kpad >= KH * KW * C && Kpad % 8 == 0, "Kpad must cover K and be a multiple of 8");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (C % 8 == 0 && ldx % 8 == 0) {
        const long long total = 1ll * B * g.Ho * g.Wo * KH * KW * (C / 8);
        DMLB_CHECK_CUDA(launch_k(im2col_nhwc_vec8_kernel, grid_for(total), 256, 0, s, static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(col), g));
    } else {
        const long long total = 1ll * B * g.Ho * g.Wo * KH * KW * C;
        DMLB_CHECK_CUDA(launch_k(im2col_nhwc_scalar_kernel, grid_for(total), 256, 0, s, static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(col), g));
    }
    DMLB_LAUNCH_CHECK();
    return 0;
}

extern "C" int dmlb_im2col_nchw_f32(const float* x, void* col, int B, int H, int W, int C, int KH, int KW, int SH, int SW, int PH, int PW,
                                    int Kpad, void* stream) {
    ConvGeom g = make_geom(B, H, W, C, KH, KW, SH, SW, PH, PW, 0, Kpad);
    DMLB_REQUIRE(g.Ho > 0 && g.Wo > 0, "empty conv output");
    DMLB_REQUIRE(Kpad >= KH * KW * C && Kpad % 8 == 0, "Kpad must cover K and be a multiple of 8");
    const long long total = 1ll * B * g.Ho * g.Wo * KH * KW;
    DMLB_CHECK_CUDA(launch_k(im2col_nchw_f32_kernel, grid_for(total), 256, 0, static_cast<cudaStream_t>(stream), x, static_cast<__nv_bfloat16*>(col), g));
    DMLB_LAUNCH_CHECK();
    return 0;
}

extern "C" int dmlb_col2im_nhwc_bf16(const void* dcol, void* dx, int B, int H, int W, int C, int ldx, int KH, int KW, int SH, int SW, int PH,
                                     int PW, int Kpad, void* stream) {
    ConvGeom g = make_geom(B, H, W, C, KH, KW, SH, SW, PH, PW, ldx, Kpad);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (C % 8 == 0 && ldx % 8 == 0) {
        const long long total = 1ll * B * H * W * (C / 8);
        DMLB_CHECK_CUDA(launch_k(col2im_nhwc_kernel<8>, grid_for(total), 256, 0, s, static_cast<const __nv_bfloat16*>(dcol), static_cast<__nv_bfloat16*>(dx), g));
    } else {
        const long long total = 1ll * B * H * W * C;
        DMLB_CHECK_CUDA(launch_k(col2im_nhwc_kernel<1>, grid_for(total), 256, 0, s, static_cast<const __nv_bfloat16*>(dcol), static_cast<__nv_bfloat16*>(dx), g));
    }
    DMLB_LAUNCH_CHECK();
    return 0;
}

static PoolGeom make_pool(int B, int H, int W, int C, int ldx, int ldo, int k, int s, int p) {
    PoolGeom g;
    g.B = B; g.H = H; g.W = W; g.C = C; g.ldx = ldx; g.ldo = ldo; g.k = k; g.s = s; g.p = p;
    g.Ho = (H + 2 * p - k) / s + 1;
    g.Wo = (W + 2 * p - k) / s + 1;
    return g;
}

extern "C" int dmlb_maxpool_fwd_bf16(const void* x, void* y, int B, int H, int W, int C, int ldx, int ldo, int k, int s, int p, void* stream) {
    PoolGeom g = make_pool(B, H, W, C, ldx, ldo, k, s, p);
    DMLB_REQUIRE(g.Ho > 0 && g.Wo > 0, "empty pool output");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (C % 8 == 0 && ldx % 8 == 0 && ldo % 8 == 0) {
        const long long total = 1ll * B * g.Ho * g.Wo * (C / 8);
        DMLB_CHECK_CUDA(launch_k(maxpool_fwd_kernel<8>, grid_for(total), 256, 0, st, static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), g));
    } else {
        const long long total = 1ll * B * g.Ho * g.Wo * C;
        DMLB_CHECK_CUDA(launch_k(maxpool_fwd_kernel<1>, grid_for(total), 256, 0, st, static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), g));
    }
    DMLB_LAUNCH_CHECK();
    return 0;
}

// has_pool=0: dout has the geometry of y (k/s/p ignored).  dbias may be NULL.
extern "C" int dmlb_act_bwd_bf16(const void* y, const void* dout, void* dy, float* dbias, int B, int H, int W, int C, int ldy, int ldd,
                                 int has_pool, int k, int s, int p, int relu, void* stream) {
    PoolGeom g = has_pool ? make_pool(B, H, W, C, ldy, ldd, k, s, p) : make_pool(B, H, W, C, ldy, ldd, 1, 1, 0);
    DMLB_REQUIRE(C * 4 <= 48 * 1024, "channel count too large for the bias-gradient staging buffer");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const bool pool_ok = !has_pool || (k == 2 && s == 2 && p == 0);
    if (C % 8 == 0 && ldy % 8 == 0 && ldd % 8 == 0 && pool_ok) {
        const int c8n = C / 8;
        const long long total = 1ll * B * H * W * c8n;
        int blocks = grid_for(total);
        if (blocks > 148 * 2) blocks = 148 * 2;
        // keep (threads in the grid) a multiple of the channel-group count so a thread's channel group never changes
        int a = c8n, bq = 256;
        while (bq) { const int tmp = a % bq; a = bq; bq = tmp; }
        const int mult = c8n / a;
        if (blocks > mult) blocks = blocks / mult * mult;
        DMLB_CHECK_CUDA(launch_k(act_bwd_vec8_kernel, blocks, 256, C * sizeof(float), st, static_cast<const __nv_bfloat16*>(y),
                                 static_cast<const __nv_bfloat16*>(dout), static_cast<__nv_bfloat16*>(dy), dbias, g, ldd, has_pool, relu));
        DMLB_LAUNCH_CHECK();
        return 0;
    }
    const long long total = 1ll * B * H * W * C;
    int blocks = grid_for(total);
    if (blocks > 148 * 4) blocks = 148 * 4;  // fewer, fatter CTAs: fewer global bias atomics
    DMLB_CHECK_CUDA(launch_k(act_bwd_kernel, blocks, 256, C * sizeof(float), st, static_cast<const __nv_bfloat16*>(y),
                             static_cast<const __nv_bfloat16*>(dout), static_cast<__nv_bfloat16*>(dy), dbias, g, ldd, has_pool, relu));
    DMLB_LAUNCH_CHECK();
    return 0;
}

extern "C" int dmlb_relu_fwd_bf16(const void* x, void* y, long long rows, int C, int ldx, int ldy, void* stream) {
    DMLB_CHECK_CUDA(launch_k(relu_fwd_kernel, grid_for(rows * C), 256, 0, static_cast<cudaStream_t>(stream), static_cast<const __nv_bfloat16*>(x),
                                                                                       static_cast<__nv_bfloat16*>(y), rows, C, ldx, ldy));
    DMLB_LAUNCH_CHECK();
    return 0;
}

extern "C" int dmlb_relu_bwd_bf16(const void* y, const void* dy, void* dx, long long rows, int C, int ldy, int lddy, int lddx, void* stream) {
    DMLB_CHECK_CUDA(launch_k(relu_bwd_kernel, grid_for(rows * C), 256, 0, static_cast<cudaStream_t>(stream), 
        static_cast<const __nv_bfloat16*>(y), static_cast<const __nv_bfloat16*>(dy), static_cast<__nv_bfloat16*>(dx), rows, C, ldy, lddy, lddx));
    DMLB_LAUNCH_CHECK();
    return 0;
}

// Data-movement and activation kernels around the tcgen05 GEMM for Conv2d / MaxPool2d / ReLU.
//
// Activations are channels-last: a logical [B, C, H, W] tensor is stored as [B, H, W, ldc] with ldc >= C and a 16-byte row granule
// (ldc % 8 == 0 for bf16, ldc % 4 == 0 for fp32) so every row is a legal TMA / 16-byte-vector operand.  Every kernel is templated on
// the element type T: __nv_bfloat16 (bf16 path) or float (tf32 path — fp32 activations, the reference's storage precision).
// Column matrices are [M = B*Ho*Wo, Kpad] with K ordered (kh, kw, cin).
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
// im2col, NHWC input, one 16-byte channel granule per thread.  Requires C % Vec16<T>::N == 0.
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) im2col_nhwc_vec_kernel(const T* __restrict__ x, T* __restrict__ col, ConvGeom g) {
    pdl_prologue();
    constexpr int V = Vec16<T>::N;
    const int cvn = g.C / V;
    const long long total = 1ll * g.B * g.Ho * g.Wo * g.KH * g.KW * cvn;
    for (long long i = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i < total; i += 1ll * gridDim.x * blockDim.x) {
        const int cv = static_cast<int>(i % cvn);
        long long t = i / cvn;
        const int kwi = static_cast<int>(t % g.KW); t /= g.KW;
        const int khi = static_cast<int>(t % g.KH); t /= g.KH;
        const int wo = static_cast<int>(t % g.Wo); t /= g.Wo;
        const int ho = static_cast<int>(t % g.Ho);
        const int b = static_cast<int>(t / g.Ho);
        const int hi = ho * g.SH - g.PH + khi;
        const int wi = wo * g.SW - g.PW + kwi;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (hi >= 0 && hi < g.H && wi >= 0 && wi < g.W)
            v = __ldg(reinterpret_cast<const uint4*>(x + ((1ll * b * g.H + hi) * g.W + wi) * g.ldx + cv * V));
        const long long m = (1ll * b * g.Ho + ho) * g.Wo + wo;
        *reinterpret_cast<uint4*>(col + m * g.Kpad + (khi * g.KW + kwi) * g.C + cv * V) = v;
    }
}

// im2col, NHWC input, any C (one element per thread).
template <typename T>
__global__ void __launch_bounds__(256) im2col_nhwc_scalar_kernel(const T* __restrict__ x, T* __restrict__ col, ConvGeom g) {
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
        T v = from_f32<T>(0.f);
        if (hi >= 0 && hi < g.H && wi >= 0 && wi < g.W) v = x[((1ll * b * g.H + hi) * g.W + wi) * g.ldx + c];
        col[m * g.Kpad + k] = v;
    }
}

// im2col straight from the user's fp32 NCHW batch (first layer): fuses the cast (bf16 path) and the
// NCHW->NHWC re-layout into the gather (reference main.py:64-70 feeds such a batch to conv1).
template <typename T>
__global__ void __launch_bounds__(256) im2col_nchw_f32_kernel(const float* __restrict__ x, T* __restrict__ col, ConvGeom g) {
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
        T* dst = col + m * g.Kpad + tap * g.C;
        for (int c = 0; c < g.C; ++c) dst[c] = from_f32<T>(ok ? __ldg(src + c * plane) : 0.f);
    }
}

// ---------------------------------------------------------------------------------------------
// col2im (gather form, no atomics): dx[b,h,w,c] = sum over taps of dcol[m(b,ho,wo), (kh,kw,c)].
// VEC = true: one 16-byte channel granule per thread.
// ---------------------------------------------------------------------------------------------
template <typename T, bool VEC>
__global__ void __launch_bounds__(256) col2im_nhwc_kernel(const T* __restrict__ dcol, T* __restrict__ dx, ConvGeom g) {
    pdl_prologue();
    constexpr int V = VEC ? Vec16<T>::N : 1;
    const int cvn = (g.C + V - 1) / V;
    const long long total = 1ll * g.B * g.H * g.W * cvn;
    for (long long i = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i < total; i += 1ll * gridDim.x * blockDim.x) {
        const int cv = static_cast<int>(i % cvn);
        long long t = i / cvn;
        const int w = static_cast<int>(t % g.W); t /= g.W;
        const int h = static_cast<int>(t % g.H);
        const int b = static_cast<int>(t / g.H);
        float acc[Vec16<T>::N];
#pragma unroll
        for (int j = 0; j < V; ++j) acc[j] = 0.f;
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
                const T* src = dcol + m * g.Kpad + (khi * g.KW + kwi) * g.C + cv * V;
                if constexpr (VEC) {
                    float f[Vec16<T>::N];
                    Vec16<T>::load(src, f);
#pragma unroll
                    for (int j = 0; j < V; ++j) acc[j] += f[j];
                } else {
                    acc[0] += to_f32<T>(src[0]);
                }
            }
        }
        T* dst = dx + ((1ll * b * g.H + h) * g.W + w) * g.ldx + cv * V;
        if constexpr (VEC) Vec16<T>::store(dst, acc);
        else dst[0] = from_f32<T>(acc[0]);
    }
}

// ---------------------------------------------------------------------------------------------
// MaxPool2d forward (NHWC), window k x k, stride s, padding p (implicit -inf).
// ---------------------------------------------------------------------------------------------
struct PoolGeom {
    int B, H, W, C, ldx;   // input
    int Ho, Wo, ldo;       // output
    int k, s, p;
};

template <typename T, bool VEC>
__global__ void __launch_bounds__(256) maxpool_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, PoolGeom g) {
    pdl_prologue();
    constexpr int V = VEC ? Vec16<T>::N : 1;
    const int cvn = (g.C + V - 1) / V;
    const long long total = 1ll * g.B * g.Ho * g.Wo * cvn;
    for (long long i = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i < total; i += 1ll * gridDim.x * blockDim.x) {
        const int cv = static_cast<int>(i % cvn);
        long long t = i / cvn;
        const int wo = static_cast<int>(t % g.Wo); t /= g.Wo;
        const int ho = static_cast<int>(t % g.Ho);
        const int b = static_cast<int>(t / g.Ho);
        float mx[Vec16<T>::N];
#pragma unroll
        for (int j = 0; j < V; ++j) mx[j] = -INFINITY;
        for (int dh = 0; dh < g.k; ++dh) {
            const int h = ho * g.s - g.p + dh;
            if (h < 0 || h >= g.H) continue;
            for (int dw = 0; dw < g.k; ++dw) {
                const int w = wo * g.s - g.p + dw;
                if (w < 0 || w >= g.W) continue;
                const T* src = x + ((1ll * b * g.H + h) * g.W + w) * g.ldx + cv * V;
                if constexpr (VEC) {
                    float f[Vec16<T>::N];
                    Vec16<T>::load(src, f);
#pragma unroll
                    for (int j = 0; j < V; ++j) mx[j] = fmaxf(mx[j], f[j]);
                } else {
                    mx[0] = fmaxf(mx[0], to_f32<T>(src[0]));
                }
            }
        }
        T* dst = y + ((1ll * b * g.Ho + ho) * g.Wo + wo) * g.ldo + cv * V;
        if constexpr (VEC) Vec16<T>::store(dst, mx);
        else dst[0] = from_f32<T>(mx[0]);
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
template <typename T>
__global__ void __launch_bounds__(256) act_bwd_kernel(const T* __restrict__ y, const T* __restrict__ dout, T* __restrict__ dy,
                                                      float* __restrict__ dbias, PoolGeom g, int ldd, int has_pool, int relu) {
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
        const float yv = to_f32<T>(y[xi]);
        float gsum = 0.f;
        if (!has_pool) {
            gsum = to_f32<T>(dout[((1ll * b * g.H + h) * g.W + w) * ldd + c]);
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
                            const float o = to_f32<T>(y[((1ll * b * g.H + hh) * g.W + ww) * g.ldx + c]);
                            const bool before = (hh < h) || (hh == h && ww < w);
                            if (o > yv || (before && o == yv)) { winner = false; break; }
                        }
                    }
                    if (winner) gsum += to_f32<T>(dout[((1ll * b * g.Ho + ho) * g.Wo + wo) * ldd + c]);
                }
        }
        if (relu && !(yv > 0.f)) gsum = 0.f;
        dy[xi] = from_f32<T>(gsum);
        if (dbias != nullptr && gsum != 0.f) atomicAdd(&sbias[c], gsum);
    }
    if (dbias != nullptr) {
        __syncthreads();
        for (int c = threadIdx.x; c < g.C; c += blockDim.x)
            if (sbias[c] != 0.f) red_add_f32(dbias + c, sbias[c]);
    }
}

// Vectorised variant (one 16-byte channel granule per thread) for the common cases: no pooling, or the 2x2 / stride-2 pool of
// AlexNet and LeNet.  Bias-gradient partial sums stay in registers while a thread keeps visiting the same channel group.
template <typename T>
__global__ void __launch_bounds__(256) act_bwd_vec_kernel(const T* __restrict__ y, const T* __restrict__ dout, T* __restrict__ dy,
                                                          float* __restrict__ dbias, PoolGeom g, int ldd, int has_pool, int relu) {
    pdl_prologue();
    constexpr int V = Vec16<T>::N;
    extern __shared__ float sbias[];  // [C]
    for (int c = threadIdx.x; c < g.C; c += blockDim.x) sbias[c] = 0.f;
    __syncthreads();
    const int cvn = g.C / V;
    const long long total = 1ll * g.B * g.H * g.W * cvn;
    float bacc[V];
#pragma unroll
    for (int j = 0; j < V; ++j) bacc[j] = 0.f;
    int cur_cv = -1;
    for (long long i = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i < total; i += 1ll * gridDim.x * blockDim.x) {
        const int cv = static_cast<int>(i % cvn);
        long long t = i / cvn;
        const int w = static_cast<int>(t % g.W); t /= g.W;
        const int h = static_cast<int>(t % g.H);
        const int b = static_cast<int>(t / g.H);
        const long long xi = ((1ll * b * g.H + h) * g.W + w) * g.ldx + cv * V;
        float yv[V], gv[V];
        Vec16<T>::load(y + xi, yv);
        if (!has_pool) {
            Vec16<T>::load(dout + ((1ll * b * g.H + h) * g.W + w) * ldd + cv * V, gv);
        } else {
            const int hp = h >> 1, wp = w >> 1;
            if (hp >= g.Ho || wp >= g.Wo) {
#pragma unroll
                for (int j = 0; j < V; ++j) gv[j] = 0.f;
            } else {
                Vec16<T>::load(dout + ((1ll * b * g.Ho + hp) * g.Wo + wp) * ldd + cv * V, gv);
                const int self = ((h & 1) << 1) | (w & 1);   // scan position inside the window
#pragma unroll
                for (int o = 0; o < 4; ++o) {
                    if (o == self) continue;
                    float ov[V];
                    const long long oi = ((1ll * b * g.H + (2 * hp + (o >> 1))) * g.W + (2 * wp + (o & 1))) * g.ldx + cv * V;
                    Vec16<T>::load(y + oi, ov);
#pragma unroll
                    for (int j = 0; j < V; ++j)
                        if (ov[j] > yv[j] || (o < self && ov[j] == yv[j])) gv[j] = 0.f;   // PyTorch: first maximum wins
                }
            }
        }
        if (relu) {
#pragma unroll
            for (int j = 0; j < V; ++j)
                if (!(yv[j] > 0.f)) gv[j] = 0.f;
        }
        Vec16<T>::store(dy + xi, gv);
        if (dbias != nullptr) {
            if (cv != cur_cv) {
                if (cur_cv >= 0) {
#pragma unroll
                    for (int j = 0; j < V; ++j) {
                        if (bacc[j] != 0.f) atomicAdd(&sbias[cur_cv * V + j], bacc[j]);
                        bacc[j] = 0.f;
                    }
                }
                cur_cv = cv;
            }
#pragma unroll
            for (int j = 0; j < V; ++j) bacc[j] += gv[j];
        }
    }
    if (dbias != nullptr) {
        if (cur_cv >= 0) {
#pragma unroll
            for (int j = 0; j < V; ++j)
                if (bacc[j] != 0.f) atomicAdd(&sbias[cur_cv * V + j], bacc[j]);
        }
        __syncthreads();
        for (int c = threadIdx.x; c < g.C; c += blockDim.x)
            if (sbias[c] != 0.f) red_add_f32(dbias + c, sbias[c]);
    }
}

// Standalone ReLU (padded rows): y = max(x, 0); backward dx = dy * (y > 0).
template <typename T>
__global__ void __launch_bounds__(256) relu_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, long long rows, int C, int ldx, int ldy) {
    pdl_prologue();
    const long long total = rows * C;
    for (long long i = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i < total; i += 1ll * gridDim.x * blockDim.x) {
        const long long r = i / C;
        const int c = static_cast<int>(i % C);
        y[r * ldy + c] = from_f32<T>(fmaxf(to_f32<T>(x[r * ldx + c]), 0.f));
    }
}
template <typename T>
__global__ void __launch_bounds__(256) relu_bwd_kernel(const T* __restrict__ y, const T* __restrict__ dy, T* __restrict__ dx, long long rows,
                                                       int C, int ldy, int lddy, int lddx) {
    pdl_prologue();
    const long long total = rows * C;
    for (long long i = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i < total; i += 1ll * gridDim.x * blockDim.x) {
        const long long r = i / C;
        const int c = static_cast<int>(i % C);
        const float yv = to_f32<T>(y[r * ldy + c]);
        dx[r * lddx + c] = yv > 0.f ? dy[r * lddy + c] : from_f32<T>(0.f);
    }
}

static inline int grid_for(long long total, int threads = 256) {
    long long blocks = (total + threads - 1) / threads;
    const long long cap = 148ll * 16;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return static_cast<int>(blocks);
}

static ConvGeom make_geom(int B, int H, int W, int C, int KH, int KW, int SH, int SW, int PH, int PW, int ldx, int Kpad) {
    ConvGeom g;
    g.B = B; g.H = H; g.W = W; g.C = C; g.KH = KH; g.KW = KW; g.SH = SH; g.SW = SW; g.PH = PH; g.PW = PW;
    g.Ho = (H + 2 * PH - KH) / SH + 1;
    g.Wo = (W + 2 * PW - KW) / SW + 1;
    g.ldx = ldx; g.Kpad = Kpad;
    return g;
}

static PoolGeom make_pool(int B, int H, int W, int C, int ldx, int ldo, int k, int s, int p) {
    PoolGeom g;
    g.B = B; g.H = H; g.W = W; g.C = C; g.ldx = ldx; g.ldo = ldo; g.k = k; g.s = s; g.p = p;
    g.Ho = (H + 2 * p - k) / s + 1;
    g.Wo = (W + 2 * p - k) / s + 1;
    return g;
}

// ---- host launchers, one per op, templated on the element type ----
template <typename T>
static int im2col_nhwc_t(const void* x_, void* col_, int B, int H, int W, int C, int ldx, int KH, int KW, int SH, int SW, int PH, int PW, int Kpad,
                         void* stream) {
    constexpr int V = 16 / sizeof(T);
    const T* x = static_cast<const T*>(x_);
    T* col = static_cast<T*>(col_);
    ConvGeom g = make_geom(B, H, W, C, KH, KW, SH, SW, PH, PW, ldx, Kpad);
    DMLB_REQUIRE(g.Ho > 0 && g.Wo > 0, "empty conv output");
    DMLB_REQUIRE(Kpad >= KH * KW * C && Kpad % V == 0, "Kpad must cover K and keep rows 16-byte aligned");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (C % V == 0 && ldx % V == 0) {
        const long long total = 1ll * B * g.Ho * g.Wo * KH * KW * (C / V);
        DMLB_CHECK_CUDA(launch_k(im2col_nhwc_vec_kernel<T>, grid_for(total), 256, 0, s, x, col, g));
    } else {
        const long long total = 1ll * B * g.Ho * g.Wo * KH * KW * C;
        DMLB_CHECK_CUDA(launch_k(im2col_nhwc_scalar_kernel<T>, grid_for(total), 256, 0, s, x, col, g));
    }
    DMLB_LAUNCH_CHECK();
    return 0;
}

template <typename T>
static int im2col_nchw_t(const float* x, void* col, int B, int H, int W, int C, int KH, int KW, int SH, int SW, int PH, int PW, int Kpad,
                         void* stream) {
    constexpr int V = 16 / sizeof(T);
    ConvGeom g = make_geom(B, H, W, C, KH, KW, SH, SW, PH, PW, 0, Kpad);
    DMLB_REQUIRE(g.Ho > 0 && g.Wo > 0, "empty conv output");
    DMLB_REQUIRE(Kpad >= KH * KW * C && Kpad % V == 0, "Kpad must cover K and keep rows 16-byte aligned");
    const long long total = 1ll * B * g.Ho * g.Wo * KH * KW;
    DMLB_CHECK_CUDA(launch_k(im2col_nchw_f32_kernel<T>, grid_for(total), 256, 0, static_cast<cudaStream_t>(stream), x, static_cast<T*>(col), g));
    DMLB_LAUNCH_CHECK();
    return 0;
}

template <typename T>
static int col2im_nhwc_t(const void* dcol_, void* dx_, int B, int H, int W, int C, int ldx, int KH, int KW, int SH, int SW, int PH, int PW,
                         int Kpad, void* stream) {
    constexpr int V = 16 / sizeof(T);
    const T* dcol = static_cast<const T*>(dcol_);
    T* dx = static_cast<T*>(dx_);
    ConvGeom g = make_geom(B, H, W, C, KH, KW, SH, SW, PH, PW, ldx, Kpad);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (C % V == 0 && ldx % V == 0 && Kpad % V == 0) {
        const long long total = 1ll * B * H * W * (C / V);
        DMLB_CHECK_CUDA(launch_k(col2im_nhwc_kernel<T, true>, grid_for(total), 256, 0, s, dcol, dx, g));
    } else {
        const long long total = 1ll * B * H * W * C;
        DMLB_CHECK_CUDA(launch_k(col2im_nhwc_kernel<T, false>, grid_for(total), 256, 0, s, dcol, dx, g));
    }
    DMLB_LAUNCH_CHECK();
    return 0;
}

template <typename T>
static int maxpool_fwd_t(const void* x_, void* y_, int B, int H, int W, int C, int ldx, int ldo, int k, int s, int p, void* stream) {
    constexpr int V = 16 / sizeof(T);
    const T* x = static_cast<const T*>(x_);
    T* y = static_cast<T*>(y_);
    PoolGeom g = make_pool(B, H, W, C, ldx, ldo, k, s, p);
    DMLB_REQUIRE(g.Ho > 0 && g.Wo > 0, "empty pool output");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (C % V == 0 && ldx % V == 0 && ldo % V == 0) {
        const long long total = 1ll * B * g.Ho * g.Wo * (C / V);
        DMLB_CHECK_CUDA(launch_k(maxpool_fwd_kernel<T, true>, grid_for(total), 256, 0, st, x, y, g));
    } else {
        const long long total = 1ll * B * g.Ho * g.Wo * C;
        DMLB_CHECK_CUDA(launch_k(maxpool_fwd_kernel<T, false>, grid_for(total), 256, 0, st, x, y, g));
    }
    DMLB_LAUNCH_CHECK();
    return 0;
}

template <typename T>
static int act_bwd_t(const void* y_, const void* dout_, void* dy_, float* dbias, int B, int H, int W, int C, int ldy, int ldd, int has_pool, int k,
                     int s, int p, int relu, void* stream) {
    constexpr int V = 16 / sizeof(T);
    const T* y = static_cast<const T*>(y_);
    const T* dout = static_cast<const T*>(dout_);
    T* dy = static_cast<T*>(dy_);
    PoolGeom g = has_pool ? make_pool(B, H, W, C, ldy, ldd, k, s, p) : make_pool(B, H, W, C, ldy, ldd, 1, 1, 0);
    DMLB_REQUIRE(C * 4 <= 48 * 1024, "channel count too large for the bias-gradient staging buffer");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const bool pool_ok = !has_pool || (k == 2 && s == 2 && p == 0);
    if (C % V == 0 && ldy % V == 0 && ldd % V == 0 && pool_ok) {
        const int cvn = C / V;
        const long long total = 1ll * B * H * W * cvn;
        int blocks = grid_for(total);
        if (blocks > 148 * 2) blocks = 148 * 2;
        // keep (threads in the grid) a multiple of the channel-group count so a thread's channel group never changes
        int a = cvn, bq = 256;
        while (bq) { const int tmp = a % bq; a = bq; bq = tmp; }
        const int mult = cvn / a;
        if (blocks > mult) blocks = blocks / mult * mult;
        DMLB_CHECK_CUDA(launch_k(act_bwd_vec_kernel<T>, blocks, 256, C * sizeof(float), st, y, dout, dy, dbias, g, ldd, has_pool, relu));
        DMLB_LAUNCH_CHECK();
        return 0;
    }
    const long long total = 1ll * B * H * W * C;
    int blocks = grid_for(total);
    if (blocks > 148 * 4) blocks = 148 * 4;  // fewer, fatter CTAs: fewer global bias atomics
    DMLB_CHECK_CUDA(launch_k(act_bwd_kernel<T>, blocks, 256, C * sizeof(float), st, y, dout, dy, dbias, g, ldd, has_pool, relu));
    DMLB_LAUNCH_CHECK();
    return 0;
}

template <typename T>
static int relu_fwd_t(const void* x, void* y, long long rows, int C, int ldx, int ldy, void* stream) {
    DMLB_CHECK_CUDA(launch_k(relu_fwd_kernel<T>, grid_for(rows * C), 256, 0, static_cast<cudaStream_t>(stream), static_cast<const T*>(x),
                             static_cast<T*>(y), rows, C, ldx, ldy));
    DMLB_LAUNCH_CHECK();
    return 0;
}

template <typename T>
static int relu_bwd_t(const void* y, const void* dy, void* dx, long long rows, int C, int ldy, int lddy, int lddx, void* stream) {
    DMLB_CHECK_CUDA(launch_k(relu_bwd_kernel<T>, grid_for(rows * C), 256, 0, static_cast<cudaStream_t>(stream), static_cast<const T*>(y),
                             static_cast<const T*>(dy), static_cast<T*>(dx), rows, C, ldy, lddy, lddx));
    DMLB_LAUNCH_CHECK();
    return 0;
}

}  // namespace dmlb

using namespace dmlb;
typedef __nv_bfloat16 bf16_t;

extern "C" int dmlb_im2col_nhwc_bf16(const void* x, void* col, int B, int H, int W, int C, int ldx, int KH, int KW, int SH, int SW, int PH,
                                     int PW, int Kpad, void* stream) {
    return im2col_nhwc_t<bf16_t>(x, col, B, H, W, C, ldx, KH, KW, SH, SW, PH, PW, Kpad, stream);
}
extern "C" int dmlb_im2col_nhwc_f32(const void* x, void* col, int B, int H, int W, int C, int ldx, int KH, int KW, int SH, int SW, int PH,
                                    int PW, int Kpad, void* stream) {
    return im2col_nhwc_t<float>(x, col, B, H, W, C, ldx, KH, KW, SH, SW, PH, PW, Kpad, stream);
}

extern "C" int dmlb_im2col_nchw_f32(const float* x, void* col, int B, int H, int W, int C, int KH, int KW, int SH, int SW, int PH, int PW,
                                    int Kpad, void* stream) {
    return im2col_nchw_t<bf16_t>(x, col, B, H, W, C, KH, KW, SH, SW, PH, PW, Kpad, stream);
}
extern "C" int dmlb_im2col_nchw_f32_f32(const float* x, void* col, int B, int H, int W, int C, int KH, int KW, int SH, int SW, int PH, int PW,
                                        int Kpad, void* stream) {
    return im2col_nchw_t<float>(x, col, B, H, W, C, KH, KW, SH, SW, PH, PW, Kpad, stream);
}

extern "C" int dmlb_col2im_nhwc_bf16(const void* dcol, void* dx, int B, int H, int W, int C, int ldx, int KH, int KW, int SH, int SW, int PH,
                                     int PW, int Kpad, void* stream) {
    return col2im_nhwc_t<bf16_t>(dcol, dx, B, H, W, C, ldx, KH, KW, SH, SW, PH, PW, Kpad, stream);
}
extern "C" int dmlb_col2im_nhwc_f32(const void* dcol, void* dx, int B, int H, int W, int C, int ldx, int KH, int KW, int SH, int SW, int PH,
                                    int PW, int Kpad, void* stream) {
    return col2im_nhwc_t<float>(dcol, dx, B, H, W, C, ldx, KH, KW, SH, SW, PH, PW, Kpad, stream);
}

extern "C" int dmlb_maxpool_fwd_bf16(const void* x, void* y, int B, int H, int W, int C, int ldx, int ldo, int k, int s, int p, void* stream) {
    return maxpool_fwd_t<bf16_t>(x, y, B, H, W, C, ldx, ldo, k, s, p, stream);
}
extern "C" int dmlb_maxpool_fwd_f32(const void* x, void* y, int B, int H, int W, int C, int ldx, int ldo, int k, int s, int p, void* stream) {
    return maxpool_fwd_t<float>(x, y, B, H, W, C, ldx, ldo, k, s, p, stream);
}

// has_pool=0: dout has the geometry of y (k/s/p ignored).  dbias may be NULL.
extern "C" int dmlb_act_bwd_bf16(const void* y, const void* dout, void* dy, float* dbias, int B, int H, int W, int C, int ldy, int ldd,
                                 int has_pool, int k, int s, int p, int relu, void* stream) {
    return act_bwd_t<bf16_t>(y, dout, dy, dbias, B, H, W, C, ldy, ldd, has_pool, k, s, p, relu, stream);
}
extern "C" int dmlb_act_bwd_f32(const void* y, const void* dout, void* dy, float* dbias, int B, int H, int W, int C, int ldy, int ldd,
                                int has_pool, int k, int s, int p, int relu, void* stream) {
    return act_bwd_t<float>(y, dout, dy, dbias, B, H, W, C, ldy, ldd, has_pool, k, s, p, relu, stream);
}

extern "C" int dmlb_relu_fwd_bf16(const void* x, void* y, long long rows, int C, int ldx, int ldy, void* stream) {
    return relu_fwd_t<bf16_t>(x, y, rows, C, ldx, ldy, stream);
}
extern "C" int dmlb_relu_fwd_f32(const void* x, void* y, long long rows, int C, int ldx, int ldy, void* stream) {
    return relu_fwd_t<float>(x, y, rows, C, ldx, ldy, stream);
}
extern "C" int dmlb_relu_bwd_bf16(const void* y, const void* dy, void* dx, long long rows, int C, int ldy, int lddy, int lddx, void* stream) {
    return relu_bwd_t<bf16_t>(y, dy, dx, rows, C, ldy, lddy, lddx, stream);
}
extern "C" int dmlb_relu_bwd_f32(const void* y, const void* dy, void* dx, long long rows, int C, int ldy, int lddy, int lddx, void* stream) {
    return relu_bwd_t<float>(y, dy, dx, rows, C, ldy, lddy, lddx, stream);
}

// Split-K forward GEMM with a fused finishing pass.
//
// The reference's conv / linear layers at its default batch of 64 (example/main.py:142) are tiny GEMMs with a
// long reduction: AlexNet conv3-5 are 256 x {384,256} x {1728,3456,2304}.  A 128 x 128 tile
// grid has 4-6 CTAs, each limited by what one SM can pull from L2, so the tensor cores idle.  Splitting K over
// gridDim.z spreads the same work over ~100 SMs; partial tiles are summed with vector fp32 red.add into an fp32
// workspace (the scratch memory of this path), and one finishing kernel applies bias + ReLU, rounds to bf16,
// optionally performs the 2x2/2 max-pool that follows the conv in the model, and writes zeros back so the
// workspace is clean for the next layer (no memset launches in the step).
#include "common.cuh"

extern "C" int dmlb_gemm_bf16(const void* A, int lda, int a_mn_major, const void* B, int ldb, int b_mn_major, int M, int N, int K,
                              void* out, int out_is_f32, int ldo, const float* bias, int relu, int split_k, int atomic, int perm_cin,
                              int perm_khw, int tile_n, void* stream);

namespace dmlb {

// ws [M, ldw] fp32 -> y [M, ldy] (bf16 or fp32) = act(ws + bias); ws <- 0.   N % 4 == 0 fast path, scalar otherwise.
template <typename T>
__global__ void __launch_bounds__(256) splitk_finish_kernel(float* __restrict__ ws, int ldw, const float* __restrict__ bias, int relu,
                                                            T* __restrict__ y, int ldy, long long M, int N) {
    pdl_prologue();
    const int n4 = (N + 3) / 4;
    const long long total = M * n4;
    for (long long i = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i < total; i += 1ll * gridDim.x * blockDim.x) {
        const long long m = i / n4;
        const int c = static_cast<int>(i % n4) * 4;
        float* w = ws + m * ldw + c;
        float4 v = *reinterpret_cast<float4*>(w);            // ldw % 4 == 0: always in bounds of the padded row
        *reinterpret_cast<float4*>(w) = make_float4(0.f, 0.f, 0.f, 0.f);
        float f[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (c + j < N) {
                float x = f[j] + (bias != nullptr ? __ldg(bias + c + j) : 0.f);
                if (relu) x = fmaxf(x, 0.f);
                y[m * ldy + c + j] = from_f32<T>(x);
            }
        }
    }
}

// Same, plus the 2x2 stride-2 max-pool of the activation: one thread per pooled pixel and 4 channels.
//   ws rows are NHWC pixels of a [B, H, W] map (H, W even); pooled is [B, H/2, W/2, ldp].
template <typename T>
__global__ void __launch_bounds__(256) splitk_finish_pool2_kernel(float* __restrict__ ws, int ldw, const float* __restrict__ bias, int relu,
                                                                  T* __restrict__ y, int ldy, T* __restrict__ pooled, int ldp, int B, int H,
                                                                  int W, int N) {
    pdl_prologue();
    const int n4 = (N + 3) / 4;
    const int Hp = H / 2, Wp = W / 2;
    const long long total = 1ll * B * Hp * Wp * n4;
    for (long long i = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i < total; i += 1ll * gridDim.x * blockDim.x) {
        const int c = static_cast<int>(i % n4) * 4;
        long long t = i / n4;
        const int wp = static_cast<int>(t % Wp); t /= Wp;
        const int hp = static_cast<int>(t % Hp);
        const int b = static_cast<int>(t / Hp);
        float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
        float bb[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) bb[j] = (bias != nullptr && c + j < N) ? __ldg(bias + c + j) : 0.f;
#pragma unroll
        for (int dh = 0; dh < 2; ++dh)
#pragma unroll
            for (int dw = 0; dw < 2; ++dw) {
                const long long m = (1ll * b * H + (2 * hp + dh)) * W + (2 * wp + dw);
                float* w = ws + m * ldw + c;
                const float4 v = *reinterpret_cast<float4*>(w);
                *reinterpret_cast<float4*>(w) = make_float4(0.f, 0.f, 0.f, 0.f);
                const float f[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    if (c + j < N) {
                        float x = f[j] + bb[j];
                        if (relu) x = fmaxf(x, 0.f);
                        const T h = from_f32<T>(x);
                        y[m * ldy + c + j] = h;
                        mx[j] = fmaxf(mx[j], to_f32<T>(h));
                    }
                }
            }
        const long long mp = (1ll * b * Hp + hp) * Wp + wp;
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (c + j < N) pooled[mp * ldp + c + j] = from_f32<T>(mx[j]);
    }
}

// ---------------------------------------------------------------------------------------------
// Finish of a split-K *data-gradient* GEMM fused with the activation backward of the layer that produced the GEMM's input
// ("conv_dgrad fused with the ReLU-mask of the previous layer").
//   ws   : fp32 [Mx, ldw], the summed dgrad = gradient w.r.t. the producer block's OUTPUT (pooled when the producer pools); zeroed here
//   y    : the producer's post-activation conv output [B, H, W, ldy]       (H = Hx, or 2*Hx with the 2x2/2 pool)
//   dy   : gradient w.r.t. the producer's pre-activation output [B, H, W, ldy] — the operand of its wgrad / dgrad GEMMs
//   dbias: fp32 [C] += column sums of dy (may be null)
// One thread handles one ws pixel x 4 channels: with pooling it routes the value to the first maximum of the 2x2 window of y
// (PyTorch's rule) and writes all four dy pixels.  Replaces splitk_finish + act_bwd (two launches, one extra round trip of dx).
// ---------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ void load4(const T* p, float (&f)[4]);
template <>
__device__ __forceinline__ void load4<float>(const float* p, float (&f)[4]) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(p));
    f[0] = v.x; f[1] = v.y; f[2] = v.z; f[3] = v.w;
}
template <>
__device__ __forceinline__ void load4<__nv_bfloat16>(const __nv_bfloat16* p, float (&f)[4]) {
    const uint2 v = __ldg(reinterpret_cast<const uint2*>(p));
    const float2 a = unpack_bf16x2(v.x), b = unpack_bf16x2(v.y);
    f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y;
}
template <typename T>
__device__ __forceinline__ void store4(T* p, const float (&f)[4]);
template <>
__device__ __forceinline__ void store4<float>(float* p, const float (&f)[4]) {
    *reinterpret_cast<float4*>(p) = make_float4(f[0], f[1], f[2], f[3]);
}
template <>
__device__ __forceinline__ void store4<__nv_bfloat16>(__nv_bfloat16* p, const float (&f)[4]) {
    uint2 o;
    o.x = pack_bf16x2(f[0], f[1]);
    o.y = pack_bf16x2(f[2], f[3]);
    *reinterpret_cast<uint2*>(p) = o;
}

template <typename T, bool POOL>
__global__ void __launch_bounds__(256) splitk_finish_actbwd_kernel(float* __restrict__ ws, int ldw, const T* __restrict__ y, T* __restrict__ dy,
                                                                   int ldy, float* __restrict__ dbias, int B, int Hx, int Wx, int C, int relu) {
    pdl_prologue();
    extern __shared__ float sbias[];  // [C]
    if (dbias != nullptr) {
        for (int c = threadIdx.x; c < C; c += blockDim.x) sbias[c] = 0.f;
        __syncthreads();
    }
    const int c4n = C >> 2;                       // C % 4 == 0 (host checks)
    const long long total = 1ll * B * Hx * Wx * c4n;
    const int H = POOL ? 2 * Hx : Hx, W = POOL ? 2 * Wx : Wx;
    float bacc[4] = {0.f, 0.f, 0.f, 0.f};
    int cur_c4 = -1;
    for (long long i = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i < total; i += 1ll * gridDim.x * blockDim.x) {
        const int c4 = static_cast<int>(i % c4n);
        long long t = i / c4n;
        const int wx = static_cast<int>(t % Wx); t /= Wx;
        const int hx = static_cast<int>(t % Hx);
        const int b = static_cast<int>(t / Hx);
        float* wp = ws + ((1ll * b * Hx + hx) * Wx + wx) * ldw + c4 * 4;
        const float4 gq = *reinterpret_cast<const float4*>(wp);
        *reinterpret_cast<float4*>(wp) = make_float4(0.f, 0.f, 0.f, 0.f);
        const float g[4] = {gq.x, gq.y, gq.z, gq.w};
        float sum[4] = {0.f, 0.f, 0.f, 0.f};
        if constexpr (!POOL) {
            const long long yi = ((1ll * b * H + hx) * W + wx) * ldy + c4 * 4;
            float yv[4], o[4];
            load4<T>(y + yi, yv);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                o[j] = (relu && !(yv[j] > 0.f)) ? 0.f : g[j];
                sum[j] = o[j];
            }
            store4<T>(dy + yi, o);
        } else {
            float yv[4][4];
            long long yi[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                yi[q] = ((1ll * b * H + (2 * hx + (q >> 1))) * W + (2 * wx + (q & 1))) * ldy + c4 * 4;
                load4<T>(y + yi[q], yv[q]);
            }
            int win[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                int best = 0;
#pragma unroll
                for (int q = 1; q < 4; ++q)
                    if (yv[q][j] > yv[best][j]) best = q;          // strict: the first maximum wins
                win[j] = best;
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                float o[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    o[j] = (win[j] == q && !(relu && !(yv[q][j] > 0.f))) ? g[j] : 0.f;
                    sum[j] += o[j];
                }
                store4<T>(dy + yi[q], o);
            }
        }
        if (dbias != nullptr) {
            if (c4 != cur_c4) {
                if (cur_c4 >= 0) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        if (bacc[j] != 0.f) atomicAdd(&sbias[cur_c4 * 4 + j], bacc[j]);
                        bacc[j] = 0.f;
                    }
                }
                cur_c4 = c4;
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) bacc[j] += sum[j];
        }
    }
    if (dbias != nullptr) {
        if (cur_c4 >= 0) {
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if (bacc[j] != 0.f) atomicAdd(&sbias[cur_c4 * 4 + j], bacc[j]);
        }
        __syncthreads();
        for (int c = threadIdx.x; c < C; c += blockDim.x)
            if (sbias[c] != 0.f) red_add_f32(dbias + c, sbias[c]);
    }
}

static inline int finish_grid(long long total) {
    long long blocks = (total + 255) / 256;
    const long long cap = 148ll * 8;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return static_cast<int>(blocks);
}

}  // namespace dmlb

using namespace dmlb;

// ws: fp32 [M, ldw], all zeros on entry and again on exit.  pool2 = 1 additionally writes the 2x2/2 max-pool of y
// (rows of y are the NHWC pixels of a [B, H, W] map with B*H*W == M) into pooled [B, H/2, W/2, ldp].
template <typename T>
static int splitk_finish_t(float* ws, int ldw, const float* bias, int relu, void* y, int ldy, long long M, int N, int pool2, int B, int H, int W,
                           void* pooled, int ldp, void* stream) {
    DMLB_REQUIRE(ldw % 4 == 0 && ldw >= N, "workspace pitch must be a multiple of 4 and cover N");
    DMLB_REQUIRE((reinterpret_cast<uintptr_t>(ws) & 15) == 0, "workspace must be 16-byte aligned");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (pool2) {
        DMLB_REQUIRE(1ll * B * H * W == M && H % 2 == 0 && W % 2 == 0, "pooled finish needs an even [B,H,W] map with B*H*W == M");
        const long long total = 1ll * B * (H / 2) * (W / 2) * ((N + 3) / 4);
        DMLB_CHECK_CUDA(launch_k(splitk_finish_pool2_kernel<T>, finish_grid(total), 256, 0, st, ws, ldw, bias, relu, static_cast<T*>(y), ldy,
                                 static_cast<T*>(pooled), ldp, B, H, W, N));
    } else {
        const long long total = M * ((N + 3) / 4);
        DMLB_CHECK_CUDA(launch_k(splitk_finish_kernel<T>, finish_grid(total), 256, 0, st, ws, ldw, bias, relu, static_cast<T*>(y), ldy, M, N));
    }
    DMLB_LAUNCH_CHECK();
    return 0;
}

extern "C" int dmlb_splitk_finish(float* ws, int ldw, const float* bias, int relu, void* y, int ldy, long long M, int N, int pool2, int B,
                                  int H, int W, void* pooled, int ldp, void* stream) {
    return splitk_finish_t<__nv_bfloat16>(ws, ldw, bias, relu, y, ldy, M, N, pool2, B, H, W, pooled, ldp, stream);
}
// fp32 outputs (tf32 path)
extern "C" int dmlb_splitk_finish_f32(float* ws, int ldw, const float* bias, int relu, void* y, int ldy, long long M, int N, int pool2, int B,
                                      int H, int W, void* pooled, int ldp, void* stream) {
    return splitk_finish_t<float>(ws, ldw, bias, relu, y, ldy, M, N, pool2, B, H, W, pooled, ldp, stream);
}

template <typename T>
static int splitk_finish_actbwd_t(float* ws, int ldw, const void* y, void* dy, int ldy, float* dbias, int B, int Hx, int Wx, int C, int pool2,
                                  int relu, void* stream) {
    DMLB_REQUIRE(ldw % 4 == 0 && ldw >= C && C % 4 == 0 && ldy % 4 == 0, "channels and pitches must be multiples of 4");
    DMLB_REQUIRE((reinterpret_cast<uintptr_t>(ws) & 15) == 0 && (reinterpret_cast<uintptr_t>(y) & 7) == 0 && (reinterpret_cast<uintptr_t>(dy) & 7) == 0,
                 "workspace / activation alignment");
    DMLB_REQUIRE(C * 4 <= 48 * 1024, "channel count too large for the bias-gradient staging buffer");
    const int c4n = C / 4;
    const long long total = 1ll * B * Hx * Wx * c4n;
    int blocks = finish_grid(total);
    if (blocks > 148 * 2) blocks = 148 * 2;
    // keep (threads in the grid) a multiple of the channel-group count so a thread's channel group never changes
    int a = c4n, bq = 256;
    while (bq) { const int tmp = a % bq; a = bq; bq = tmp; }
    const int mult = c4n / a;
    if (blocks > mult) blocks = blocks / mult * mult;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t smem = dbias != nullptr ? C * sizeof(float) : 0;
    if (pool2)
        DMLB_CHECK_CUDA(launch_k(splitk_finish_actbwd_kernel<T, true>, blocks, 256, smem, st, ws, ldw, static_cast<const T*>(y), static_cast<T*>(dy), ldy,
                                 dbias, B, Hx, Wx, C, relu));
    else
        DMLB_CHECK_CUDA(launch_k(splitk_finish_actbwd_kernel<T, false>, blocks, 256, smem, st, ws, ldw, static_cast<const T*>(y), static_cast<T*>(dy), ldy,
                                 dbias, B, Hx, Wx, C, relu));
    DMLB_LAUNCH_CHECK();
    return 0;
}

// ws: fp32 [B*Hx*Wx, ldw] summed split-K dgrad (zeroed again here); y / dy: [B, H, W, ldy] with (H, W) = (Hx, Wx), or (2Hx, 2Wx) when
// pool2 = 1 (the producer block ends in a 2x2/2 max-pool).  is_f32 selects fp32 (tf32 path) or bf16 activations.
extern "C" int dmlb_splitk_finish_actbwd(float* ws, int ldw, const void* y, void* dy, int ldy, float* dbias, int B, int Hx, int Wx, int C,
                                         int pool2, int relu, int is_f32, void* stream) {
    if (is_f32) return splitk_finish_actbwd_t<float>(ws, ldw, y, dy, ldy, dbias, B, Hx, Wx, C, pool2, relu, stream);
    return splitk_finish_actbwd_t<__nv_bfloat16>(ws, ldw, y, dy, ldy, dbias, B, Hx, Wx, C, pool2, relu, stream);
}

// y[M,N] (bf16, pitch ldy) = act(A[M,K] * W[N,K]^T + bias) with K split `split_k` ways.
// Scratch: ws, fp32 [M, round_up(N,4)] — must be all zeros on entry; it is all zeros again when the call's work retires.
extern "C" int dmlb_gemm_bias_act_splitk(const void* A, int lda, const void* Wt, int ldb, int M, int N, int K, float* ws, const float* bias,
                                         int relu, void* y, int ldy, int split_k, void* stream) {
    const int ldw = (N + 3) / 4 * 4;
    int rc = dmlb_gemm_bf16(A, lda, 0, Wt, ldb, 0, M, N, K, ws, 1, ldw, nullptr, 0, split_k, 1, 0, 0, 0, stream);
    if (rc) return rc;
    return dmlb_splitk_finish(ws, ldw, bias, relu, y, ldy, M, N, 0, 0, 0, 0, nullptr, 0, stream);
}

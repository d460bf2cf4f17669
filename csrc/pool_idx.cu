// Max-pool with saved arg-max indices, for general windows (ResNet stem: 3x3, stride 2, padding 1).
//
// The recompute-from-activations backward in conv_ops.cu (act_bwd) is ideal for the non-overlapping 2x2/2 pools of the
// reference's AlexNet / LeNet (example/models.py:16-17,31,34,41), but for overlapping windows it re-reads up to
// 4 windows x 9 taps per input element.  Here the forward pass records, per output element, which tap of its window held
// the first maximum (one byte); the backward pass then needs at most ceil(k/s)^2 index + gradient loads per input
// element, 8 channels (16 bytes) at a time, with no atomics.
#include "common.cuh"

namespace dmlb {

struct PoolIdxGeom {
    int B, H, W, C, ldx;
    int Ho, Wo, ldo;
    int k, s, p;
};

__device__ __forceinline__ void pi_unpack8(const uint4& v, float (&f)[8]) {
    const float2 a = unpack_bf16x2(v.x), b = unpack_bf16x2(v.y), c = unpack_bf16x2(v.z), d = unpack_bf16x2(v.w);
    f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}

__global__ void __launch_bounds__(256) maxpool_fwd_idx_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y,
                                                              uint8_t* __restrict__ idx, PoolIdxGeom g) {
    pdl_prologue();
    const int c8n = g.C >> 3;
    const long long total = 1ll * g.B * g.Ho * g.Wo * c8n;
    for (long long i = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i < total; i += 1ll * gridDim.x * blockDim.x) {
        const int c8 = static_cast<int>(i % c8n);
        long long t = i / c8n;
        const int wo = static_cast<int>(t % g.Wo); t /= g.Wo;
        const int ho = static_cast<int>(t % g.Ho);
        const int b = static_cast<int>(t / g.Ho);
        float mx[8];
        int arg[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) { mx[j] = -INFINITY; arg[j] = 0; }
        for (int dh = 0; dh < g.k; ++dh) {
            const int h = ho * g.s - g.p + dh;
            if (h < 0 || h >= g.H) continue;
            for (int dw = 0; dw < g.k; ++dw) {
                const int w = wo * g.s - g.p + dw;
                if (w < 0 || w >= g.W) continue;
                float v[8];
                pi_unpack8(__ldg(reinterpret_cast<const uint4*>(x + ((1ll * b * g.H + h) * g.W + w) * g.ldx + c8 * 8)), v);
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    if (v[j] > mx[j]) { mx[j] = v[j]; arg[j] = dh * g.k + dw; }   // strict >: the first maximum wins (PyTorch)
            }
        }
        const long long o = (1ll * b * g.Ho + ho) * g.Wo + wo;
        uint4 out;
        out.x = pack_bf16x2(mx[0], mx[1]); out.y = pack_bf16x2(mx[2], mx[3]);
        out.z = pack_bf16x2(mx[4], mx[5]); out.w = pack_bf16x2(mx[6], mx[7]);
        *reinterpret_cast<uint4*>(y + o * g.ldo + c8 * 8) = out;
        uint2 pk;
        pk.x = arg[0] | (arg[1] << 8) | (arg[2] << 16) | (arg[3] << 24);
        pk.y = arg[4] | (arg[5] << 8) | (arg[6] << 16) | (arg[7] << 24);
        *reinterpret_cast<uint2*>(idx + o * g.C + c8 * 8) = pk;
    }
}

__global__ void __launch_bounds__(256) maxpool_bwd_idx_kernel(const __nv_bfloat16* __restrict__ dout, const uint8_t* __restrict__ idx,
                                                              __nv_bfloat16* __restrict__ dx, PoolIdxGeom g, int ldd) {
    pdl_prologue();
    const int c8n = g.C >> 3;
    const long long total = 1ll * g.B * g.H * g.W * c8n;
    for (long long i = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i < total; i += 1ll * gridDim.x * blockDim.x) {
        const int c8 = static_cast<int>(i % c8n);
        long long t = i / c8n;
        const int w = static_cast<int>(t % g.W); t /= g.W;
        const int h = static_cast<int>(t % g.H);
        const int b = static_cast<int>(t / g.H);
        float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        const int ho_hi = min(g.Ho - 1, (h + g.p) / g.s);
        const int wo_hi = min(g.Wo - 1, (w + g.p) / g.s);
        for (int ho = ho_hi; ho >= 0 && ho * g.s - g.p + g.k > h; --ho)
            for (int wo = wo_hi; wo >= 0 && wo * g.s - g.p + g.k > w; --wo) {
                const int me = (h - (ho * g.s - g.p)) * g.k + (w - (wo * g.s - g.p));
                const long long o = (1ll * b * g.Ho + ho) * g.Wo + wo;
                const uint2 pk = __ldg(reinterpret_cast<const uint2*>(idx + o * g.C + c8 * 8));
                float gv[8];
                pi_unpack8(__ldg(reinterpret_cast<const uint4*>(dout + o * ldd + c8 * 8)), gv);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int a = ((j < 4 ? pk.x : pk.y) >> (8 * (j & 3))) & 0xFF;
                    if (a == me) acc[j] += gv[j];
                }
            }
        uint4 out;
        out.x = pack_bf16x2(acc[0], acc[1]); out.y = pack_bf16x2(acc[2], acc[3]);
        out.z = pack_bf16x2(acc[4], acc[5]); out.w = pack_bf16x2(acc[6], acc[7]);
        *reinterpret_cast<uint4*>(dx + ((1ll * b * g.H + h) * g.W + w) * g.ldx + c8 * 8) = out;
    }
}

static inline int pool_grid(long long total) {
    long long blocks = (total + 255) / 256;
    const long long cap = 148ll * 16;
    if (blocks > cap) blocks = cap;
    return static_cast<int>(blocks < 1 ? 1 : blocks);
}

static PoolIdxGeom pool_geom(int B, int H, int W, int C, int ldx, int ldo, int k, int s, int p) {
    PoolIdxGeom g;
    g.B = B; g.H = H; g.W = W; g.C = C; g.ldx = ldx; g.ldo = ldo; g.k = k; g.s = s; g.p = p;
    g.Ho = (H + 2 * p - k) / s + 1;
    g.Wo = (W + 2 * p - k) / s + 1;
    return g;
}

}  // namespace dmlb

using namespace dmlb;

// idx: uint8 [B, Ho, Wo, C] (dense) <- tap (dh * k + dw) of the first maximum of each window.
extern "C" int dmlb_maxpool_fwd_idx_bf16(const void* x, void* y, void* idx, int B, int H, int W, int C, int ldx, int ldo, int k, int s, int p,
                                         void* stream) {
    DMLB_REQUIRE(C % 8 == 0 && ldx % 8 == 0 && ldo % 8 == 0 && k * k <= 255, "indexed max-pool needs C % 8 == 0 and k*k <= 255");
    PoolIdxGeom g = pool_geom(B, H, W, C, ldx, ldo, k, s, p);
    DMLB_REQUIRE(g.Ho > 0 && g.Wo > 0, "empty pool output");
    const long long total = 1ll * B * g.Ho * g.Wo * (C / 8);
    DMLB_CHECK_CUDA(launch_k(maxpool_fwd_idx_kernel, pool_grid(total), 256, 0, static_cast<cudaStream_t>(stream),
                             static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), static_cast<uint8_t*>(idx), g));
    DMLB_LAUNCH_CHECK();
    return 0;
}

extern "C" int dmlb_maxpool_bwd_idx_bf16(const void* dout, const void* idx, void* dx, int B, int H, int W, int C, int lddx, int ldd, int k,
                                         int s, int p, void* stream) {
    DMLB_REQUIRE(C % 8 == 0 && lddx % 8 == 0 && ldd % 8 == 0, "indexed max-pool needs C % 8 == 0");
    PoolIdxGeom g = pool_geom(B, H, W, C, lddx, ldd, k, s, p);
    const long long total = 1ll * B * H * W * (C / 8);
    DMLB_CHECK_CUDA(launch_k(maxpool_bwd_idx_kernel, pool_grid(total), 256, 0, static_cast<cudaStream_t>(stream),
                             static_cast<const __nv_bfloat16*>(dout), static_cast<const uint8_t*>(idx), static_cast<__nv_bfloat16*>(dx), g, ldd));
    DMLB_LAUNCH_CHECK();
    return 0;
}

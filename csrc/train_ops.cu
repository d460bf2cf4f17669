// Loss, optimizer, dropout and cast kernels (single-GPU part of the training step).
//
// Reference call sites replaced:
//   F.cross_entropy + torch.max(outputs, 1)      example/main.py:71,75,124-125   -> softmax_xent (K11)
//   per-tensor p.data.add_(-lr, grad) loop       asgd/optim/Asynchronous.py:63-68 -> sgd_flat (K5)
//   accumulated_gradients.add_(-lr, gradients)   asgd/optim/Asynchronous.py:55    -> folded into sgd_flat
//   Dropout2d / F.dropout                        example/models.py:10,17,20       -> dropout (K12)
#include "common.cuh"

namespace dmlb {

// ---------------------------------------------------------------------------------------------
// softmax cross-entropy, forward + backward in one launch (one CTA, one warp per row at a time).
//   logits : fp32 or bf16 [B, ldl]
//   labels : int64 [B]
//   metrics: fp32 [2] <- { mean loss, number of rows whose arg-max equals the label }
//   dlogits: same dtype as logits, [B, ldd] <- (softmax - onehot) / B   (may be null: evaluation)
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(1024) softmax_xent_kernel(const T* __restrict__ logits, const long long* __restrict__ labels,
                                                            float* __restrict__ metrics, T* __restrict__ dlogits, int B, int C,
                                                            int ldl, int ldd, volatile float* host_ring, unsigned long long* ring_ctr,
                                                            int ring_slots) {
    pdl_prologue();
    __shared__ float s_loss[32];
    __shared__ float s_hit[32];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
    float loss_acc = 0.f, hit_acc = 0.f;
    const float inv_b = 1.f / static_cast<float>(B);
    for (int r = warp; r < B; r += nwarps) {
        const T* row = logits + static_cast<size_t>(r) * ldl;
        const int label = static_cast<int>(labels[r]);
        float mx = -INFINITY;
        int arg = 0x7fffffff;
        for (int c = lane; c < C; c += 32) {
            const float v = to_f32<T>(row[c]);
            if (v > mx) { mx = v; arg = c; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float om = __shfl_xor_sync(0xffffffffu, mx, o);
            const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
            if (om > mx || (om == mx && oa < arg)) { mx = om; arg = oa; }
        }
        float se = 0.f;
        for (int c = lane; c < C; c += 32) se += __expf(to_f32<T>(row[c]) - mx);
        se = warp_sum(se);
        const float lse = mx + __logf(se);
        if (dlogits != nullptr) {
            T* drow = dlogits + static_cast<size_t>(r) * ldd;
            for (int c = lane; c < C; c += 32) {
                const float p = __expf(to_f32<T>(row[c]) - lse);
                drow[c] = from_f32<T>((p - (c == label ? 1.f : 0.f)) * inv_b);
            }
        }
        if (lane == 0) {
            loss_acc += (label >= 0 && label < C) ? (lse - to_f32<T>(row[label])) : 0.f;
            hit_acc += (arg == label) ? 1.f : 0.f;
        }
    }
    if (lane == 0) { s_loss[warp] = loss_acc; s_hit[warp] = hit_acc; }
    __syncthreads();
    if (warp == 0) {
        float l = lane < nwarps ? s_loss[lane] : 0.f;
        float h = lane < nwarps ? s_hit[lane] : 0.f;
        l = warp_sum(l);
        h = warp_sum(h);
        if (lane == 0) {
            metrics[0] = l * inv_b;
            metrics[1] = h;
            if (host_ring != nullptr) {
                // result mailbox in pinned host memory: slot = launch number mod ring_slots, {loss, hits, launch number + 1}; the sequence
                // word is written last, so a host that sees it sees the values (no D2H copy, no event on the step stream)
                const unsigned long long n = *ring_ctr;
                volatile float* slot = host_ring + 4 * (n % static_cast<unsigned long long>(ring_slots));
                slot[0] = l * inv_b;
                slot[1] = h;
                __threadfence_system();
                slot[2] = static_cast<float>((n + 1ull) & 0xFFFFFFull);
                *ring_ctr = n + 1ull;
            }
        }
    }
}

// x[r, c] *= *scale  (bf16, padded rows) — applies an upstream scalar gradient to the saved dlogits.
template <typename T>
__global__ void __launch_bounds__(256) scale_rows_kernel(T* __restrict__ x, const float* __restrict__ scale, long long rows, int C, int ld) {
    pdl_prologue();
    const float s = *scale;
    const long long total = rows * C;
    for (long long i = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i < total; i += 1ll * gridDim.x * blockDim.x) {
        const long long r = i / C;
        const int c = static_cast<int>(i % C);
        x[r * ld + c] = from_f32<T>(to_f32<T>(x[r * ld + c]) * s);
    }
}

// ---------------------------------------------------------------------------------------------
// Flat SGD over the parameter arena, one launch for the whole model:
//   g  = grad + wd * theta;   v = mu * v + g (if momentum);   theta -= lr * g;   acc -= lr * g (if acc);
//   shadow = bf16(theta) (if shadow: the flat bf16 mirror the tensor-core GEMMs read)
// lr is read from device memory so a captured CUDA graph follows LR-scheduler changes.
// n must be a multiple of 4 (the arena pads to that).
// ---------------------------------------------------------------------------------------------
struct ShadowEntry {
    long long src_off;   // elements into the fp32 arena
    long long dst_off;   // elements into the shadow arena
    int rows, cin, khw, ld;
    int cpad;            // 0: columns (khw, cin) dense.  > 0: small-channel layout, cpad channels per tap (the extra ones stay zero)
    int src_gemm_order;  // source is stored [rows, khw, cin] (channels-last arena) instead of PyTorch's [rows, cin, khw]
    int kw, kwpad;       // kwpad > 0 (with cpad): taps are laid out (kh, kw padded to kwpad) — the row-window convolution's weight order
};

// Destination column of element k = (kk, ci) of a GEMM-layout shadow row (see ShadowEntry).
__device__ __forceinline__ int shadow_col(const ShadowEntry& e, int kk, int ci, int k) {
    if (e.cpad <= 0) return k;
    int tap = kk;
    if (e.kwpad > 0) tap = (kk / e.kw) * e.kwpad + kk % e.kw;
    return tap * e.cpad + ci;
}

// One updated master value -> the table-backed shadow it belongs to, if any (weights whose GEMM-layout copy is not a plain cast of the
// arena: padded rows, PyTorch (cin,kh,kw) order, first-layer tap padding).  Folding this scatter into the optimizer kernel removes the
// separate refresh_shadows launch from the end of every training step.
template <typename T>
__device__ __forceinline__ void scatter_to_shadows(const ShadowEntry* __restrict__ table, int n_entries, long long elem, float v) {
    for (int t = 0; t < n_entries; ++t) {
        const ShadowEntry e = table[t];
        const int K = e.cin * e.khw;
        const long long local = elem - e.src_off;
        if (local < 0 || local >= 1ll * e.rows * K) continue;
        const int r = static_cast<int>(local / K), rem = static_cast<int>(local - 1ll * r * K);
        int kk, ci;
        if (e.src_gemm_order) { kk = rem / e.cin; ci = rem - kk * e.cin; }
        else                  { ci = rem / e.khw; kk = rem - ci * e.khw; }
        // dst_off is an absolute address in units of T (not an offset from a null base: that would be undefined behaviour)
        T* dst = reinterpret_cast<T*>(static_cast<uintptr_t>(e.dst_off) * sizeof(T));
        dst[1ll * r * e.ld + shadow_col(e, kk, ci, kk * e.cin + ci)] = from_f32<T>(v);
        return;
    }
}

__global__ void __launch_bounds__(256) sgd_flat_kernel(float* __restrict__ params, float* __restrict__ grads, float* __restrict__ acc,
                                                       float* __restrict__ mom, __nv_bfloat16* __restrict__ shadow, long long n4,
                                                       const float* __restrict__ lr_ptr, float mu, float wd, int zero_grads,
                                                       const ShadowEntry* __restrict__ table, int n_entries, int table_f32, long long table_lo,
                                                       long long table_hi) {
    pdl_prologue();
    const float lr = *lr_ptr;
    for (long long i = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i < n4; i += 1ll * gridDim.x * blockDim.x) {
        float4 p = reinterpret_cast<float4*>(params)[i];
        float4 g = ld_stream_f32x4(grads + 4 * i);
        if (zero_grads) reinterpret_cast<float4*>(grads)[i] = make_float4(0.f, 0.f, 0.f, 0.f);  // next step starts clean: no memset node
        if (wd != 0.f) { g.x += wd * p.x; g.y += wd * p.y; g.z += wd * p.z; g.w += wd * p.w; }
        if (mom != nullptr) {
            float4 v = reinterpret_cast<float4*>(mom)[i];
            v.x = mu * v.x + g.x; v.y = mu * v.y + g.y; v.z = mu * v.z + g.z; v.w = mu * v.w + g.w;
            reinterpret_cast<float4*>(mom)[i] = v;
            g = v;
        }
        if (acc != nullptr) {
            float4 a = reinterpret_cast<float4*>(acc)[i];
            a.x -= lr * g.x; a.y -= lr * g.y; a.z -= lr * g.z; a.w -= lr * g.w;
            reinterpret_cast<float4*>(acc)[i] = a;
        }
        p.x -= lr * g.x; p.y -= lr * g.y; p.z -= lr * g.z; p.w -= lr * g.w;
        reinterpret_cast<float4*>(params)[i] = p;
        if (shadow != nullptr) {  // bf16 mirror for the tensor cores, same flat offsets
            uint2 h;
            h.x = pack_bf16x2(p.x, p.y);
            h.y = pack_bf16x2(p.z, p.w);
            reinterpret_cast<uint2*>(shadow)[i] = h;
        }
        if (n_entries > 0 && 4 * i + 3 >= table_lo && 4 * i < table_hi) {      // [table_lo, table_hi): hull of the table-backed weights
            const float pv[4] = {p.x, p.y, p.z, p.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                // table offsets are absolute (address / element size): rebase the arena index
                const long long abs_elem = (reinterpret_cast<long long>(params) >> 2) + 4 * i + j;
                if (table_f32) scatter_to_shadows<float>(table, n_entries, abs_elem, pv[j]);
                else           scatter_to_shadows<__nv_bfloat16>(table, n_entries, abs_elem, pv[j]);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// bf16 "shadow" weights for the tensor cores, refreshed from the fp32 master arena in one launch.
// Each table entry describes one weight tensor [rows, cin, khw] (PyTorch layout, khw = kh*kw, 1 for
// Linear) and where its GEMM-layout copy [rows, ld] with columns ordered (khw, cin) lives.
// ---------------------------------------------------------------------------------------------


template <typename T>
__global__ void __launch_bounds__(256) refresh_shadows_kernel(const float* __restrict__ params, T* __restrict__ shadow,
                                                              const ShadowEntry* __restrict__ table) {
    pdl_prologue();
    const ShadowEntry e = table[blockIdx.y];
    const int K = e.cin * e.khw;
    const long long total = 1ll * e.rows * K;
    const float* src = params + e.src_off;
    T* dst = shadow + e.dst_off;
    for (long long i = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i < total; i += 1ll * gridDim.x * blockDim.x) {
        // iterate in destination order so the bf16 stores coalesce; sources are gathered (L2-resident)
        const int r = static_cast<int>(i / K);
        const int k = static_cast<int>(i % K);
        const int ci = k % e.cin;
        const int kk = k / e.cin;
        const float v = e.src_gemm_order ? __ldg(src + 1ll * r * K + k) : __ldg(src + (1ll * r * e.cin + ci) * e.khw + kk);
        dst[1ll * r * e.ld + shadow_col(e, kk, ci, k)] = from_f32<T>(v);
    }
}

// Generic 2-D cast fp32 -> bf16 with row pitches (inputs of Linear layers fed with fp32 tensors).
__global__ void __launch_bounds__(256) cast_f32_bf16_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ y, long long rows, int C,
                                                            long long ldx, int ldy) {
    pdl_prologue();
    const long long total = rows * C;
    for (long long i = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i < total; i += 1ll * gridDim.x * blockDim.x) {
        const long long r = i / C;
        const int c = static_cast<int>(i % C);
        y[r * ldy + c] = __float2bfloat16_rn(x[r * ldx + c]);
    }
}
__global__ void __launch_bounds__(256) cast_bf16_f32_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ y, long long rows, int C,
                                                            int ldx, long long ldy) {
    pdl_prologue();
    const long long total = rows * C;
    for (long long i = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i < total; i += 1ll * gridDim.x * blockDim.x) {
        const long long r = i / C;
        const int c = static_cast<int>(i % C);
        y[r * ldy + c] = __bfloat162float(x[r * ldx + c]);
    }
}

// ---------------------------------------------------------------------------------------------
// Dropout (element) / Dropout2d (whole channel of one sample), Philox-free counter hash so the mask is
// a pure function of (seed, *step, unit index): CUDA-graph replays draw fresh masks because *step
// lives in device memory and is bumped by dmlb_counter_bump after each use.
//   rows x C elements, pitch ld.  unit = element index (mode 0) or (sample, channel) (mode 1; rows_per_sample = H*W).
//   mask: uint8 [rows * C] (1 = kept), saved for the backward pass.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t mix32(uint64_t z) {  // splitmix64 finaliser
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= z >> 31;
    return static_cast<uint32_t>(z >> 32);
}

template <typename T>
__global__ void __launch_bounds__(256) dropout_fwd_kernel(const T* __restrict__ x, T* __restrict__ y,
                                                          uint8_t* __restrict__ mask, long long rows, int C, int ldx, int ldy,
                                                          int rows_per_sample, int channel_mode, float p, uint64_t seed,
                                                          const unsigned long long* __restrict__ step) {
    pdl_prologue();
    const uint64_t st = *step;
    const float scale = 1.f / (1.f - p);
    const uint32_t thresh = static_cast<uint32_t>(static_cast<double>(p) * 4294967296.0);
    const long long total = rows * C;
    for (long long i = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i < total; i += 1ll * gridDim.x * blockDim.x) {
        const long long r = i / C;
        const int c = static_cast<int>(i % C);
        const uint64_t unit = channel_mode ? static_cast<uint64_t>((r / rows_per_sample) * C + c) : static_cast<uint64_t>(i);
        const uint32_t rnd = mix32(seed ^ (st * 0xD1B54A32D192ED03ull) ^ (unit * 0x9E3779B97F4A7C15ull + 0x632BE59BD9B4E019ull));
        const bool keep = rnd >= thresh;
        mask[i] = keep ? 1 : 0;
        y[r * ldy + c] = keep ? from_f32<T>(to_f32<T>(x[r * ldx + c]) * scale) : from_f32<T>(0.f);
    }
}
template <typename T>
__global__ void __launch_bounds__(256) dropout_bwd_kernel(const T* __restrict__ dy, const uint8_t* __restrict__ mask,
                                                          T* __restrict__ dx, long long rows, int C, int lddy, int lddx, float p) {
    pdl_prologue();
    const float scale = 1.f / (1.f - p);
    const long long total = rows * C;
    for (long long i = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i < total; i += 1ll * gridDim.x * blockDim.x) {
        const long long r = i / C;
        const int c = static_cast<int>(i % C);
        dx[r * lddx + c] = mask[i] ? from_f32<T>(to_f32<T>(dy[r * lddy + c]) * scale) : from_f32<T>(0.f);
    }
}
__global__ void counter_bump_kernel(unsigned long long* c) { *c += 1ull; }

static inline int flat_grid(long long total, int threads = 256) {
    long long blocks = (total + threads - 1) / threads;
    const long long cap = 148ll * 8;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return static_cast<int>(blocks);
}

}  // namespace dmlb

using namespace dmlb;

static int softmax_xent_impl(const void* logits, int logits_is_bf16, const long long* labels, float* metrics, void* dlogits, int B, int C, int ldl,
                             int ldd, float* host_ring, void* ring_ctr, int ring_slots, void* stream);

extern "C" int dmlb_softmax_xent(const void* logits, int logits_is_bf16, const long long* labels, float* metrics, void* dlogits, int B, int C,
                                 int ldl, int ldd, void* stream) {
    return softmax_xent_impl(logits, logits_is_bf16, labels, metrics, dlogits, B, C, ldl, ldd, nullptr, nullptr, 0, stream);
}

// The same kernel additionally posts {loss, hits, sequence number} into a ring of `ring_slots` 16-byte slots in pinned (device-visible)
// host memory; ring_ctr is a device counter of launches.  Replaces the reference's blocking loss.item() (example/main.py:79) without
// putting a D2H copy or an event between two replays of the training step.
extern "C" int dmlb_softmax_xent_ring(const void* logits, int logits_is_bf16, const long long* labels, float* metrics, void* dlogits, int B, int C,
                                      int ldl, int ldd, float* host_ring, void* ring_ctr, int ring_slots, void* stream) {
    DMLB_REQUIRE(host_ring != nullptr && ring_ctr != nullptr && ring_slots > 0, "ring buffer arguments");
    return softmax_xent_impl(logits, logits_is_bf16, labels, metrics, dlogits, B, C, ldl, ldd, host_ring, ring_ctr, ring_slots, stream);
}

static int softmax_xent_impl(const void* logits, int logits_is_bf16, const long long* labels, float* metrics, void* dlogits, int B, int C, int ldl,
                             int ldd, float* host_ring, void* ring_ctr, int ring_slots, void* stream) {
    DMLB_REQUIRE(B > 0 && C > 0, "empty logits");
    int threads = 32 * ((B < 32) ? B : 32);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (logits_is_bf16)
        DMLB_CHECK_CUDA(launch_k(softmax_xent_kernel<__nv_bfloat16>, 1, threads, 0, st, static_cast<const __nv_bfloat16*>(logits), labels, metrics,
                                                                  static_cast<__nv_bfloat16*>(dlogits), B, C, ldl, ldd, static_cast<volatile float*>(host_ring),
                                                                  static_cast<unsigned long long*>(ring_ctr), ring_slots));
    else
        DMLB_CHECK_CUDA(launch_k(softmax_xent_kernel<float>, 1, threads, 0, st, static_cast<const float*>(logits), labels, metrics, static_cast<float*>(dlogits), B, C,
                                                          ldl, ldd, static_cast<volatile float*>(host_ring), static_cast<unsigned long long*>(ring_ctr), ring_slots));
    DMLB_LAUNCH_CHECK();
    return 0;
}

extern "C" int dmlb_scale_rows_bf16(void* x, const float* scale, long long rows, int C, int ld, void* stream) {
    DMLB_CHECK_CUDA(launch_k(scale_rows_kernel<__nv_bfloat16>, flat_grid(rows * C), 256, 0, static_cast<cudaStream_t>(stream), static_cast<__nv_bfloat16*>(x), scale, rows, C, ld));
    DMLB_LAUNCH_CHECK();
    return 0;
}
extern "C" int dmlb_scale_rows_f32(void* x, const float* scale, long long rows, int C, int ld, void* stream) {
    DMLB_CHECK_CUDA(launch_k(scale_rows_kernel<float>, flat_grid(rows * C), 256, 0, static_cast<cudaStream_t>(stream), static_cast<float*>(x), scale, rows, C, ld));
    DMLB_LAUNCH_CHECK();
    return 0;
}

// dmlb_sgd_flat plus the table-backed weight shadows (dmlb_refresh_shadows' table with ABSOLUTE addresses: src_off / dst_off = address /
// element size) refreshed by the same launch.  [table_lo, table_hi) is the element range of the arena (relative to params) that holds
// those weights.
extern "C" int dmlb_sgd_flat_ex(float* params, float* grads, float* acc, float* mom, void* shadow, long long n, const float* lr_ptr, float mu,
                                float wd, int zero_grads, const void* table, int n_entries, int table_is_f32, long long table_lo,
                                long long table_hi, void* stream) {
    DMLB_REQUIRE(n % 4 == 0, "flat SGD length must be a multiple of 4 (arena padding)");
    DMLB_REQUIRE(((reinterpret_cast<uintptr_t>(params) | reinterpret_cast<uintptr_t>(grads)) & 15) == 0, "arena must be 16-byte aligned");
    DMLB_CHECK_CUDA(launch_k(sgd_flat_kernel, flat_grid(n / 4), 256, 0, static_cast<cudaStream_t>(stream), params, grads, acc, mom,
                             static_cast<__nv_bfloat16*>(shadow), n / 4, lr_ptr, mu, wd, zero_grads, static_cast<const ShadowEntry*>(table),
                             table != nullptr ? n_entries : 0, table_is_f32, table_lo, table_hi));
    DMLB_LAUNCH_CHECK();
    return 0;
}

extern "C" int dmlb_sgd_flat(float* params, float* grads, float* acc, float* mom, void* shadow, long long n, const float* lr_ptr, float mu,
                             float wd, int zero_grads, void* stream) {
    return dmlb_sgd_flat_ex(params, grads, acc, mom, shadow, n, lr_ptr, mu, wd, zero_grads, nullptr, 0, 0, 0, 0, stream);
}

namespace dmlb {
// fp32 NCHW batch -> channels-last activation with ONE 16-byte granule per pixel (C <= 4 fp32 / 8 bf16 channels, the rest zero): the
// input layout of the small-channel first-layer convolution.  One thread per pixel: C strided reads, one 16-byte store.
// The destination may be a zero-padded buffer [B, Hp, Wp, G] (image at rows ph.., columns pw..; the border is zeroed once, by its owner).
template <typename T>
__global__ void __launch_bounds__(256) nchw_to_act_kernel(const float* __restrict__ x, T* __restrict__ y, int B, int C, int H, int W, int Hp, int Wp,
                                                          int ph, int pw) {
    pdl_prologue();
    constexpr int G = 16 / sizeof(T);
    const long long plane = 1ll * H * W;
    const long long total = 1ll * B * plane;
    for (long long i = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i < total; i += 1ll * gridDim.x * blockDim.x) {
        const long long b = i / plane, p = i - b * plane;
        const int h = static_cast<int>(p / W), w = static_cast<int>(p - 1ll * h * W);
        const float* src = x + b * C * plane + p;
        float f[G];
#pragma unroll
        for (int c = 0; c < G; ++c) f[c] = c < C ? __ldg(src + c * plane) : 0.f;
        Vec16<T>::store(y + ((b * Hp + h + ph) * Wp + w + pw) * G, f);
    }
}

}  // namespace dmlb

template <typename T>
static int nchw_to_act_t(const float* x, void* y, int B, int C, int H, int W, int hp, int wp, int ph, int pw, void* stream) {
    DMLB_REQUIRE(C >= 1 && C <= static_cast<int>(16 / sizeof(T)), "one 16-byte granule holds at most 8 bf16 / 4 fp32 channels");
    if (hp <= 0) { hp = H; wp = W; ph = 0; pw = 0; }
    DMLB_REQUIRE(hp >= H + ph && wp >= W + pw && ph >= 0 && pw >= 0, "the image does not fit the padded buffer");
    DMLB_CHECK_CUDA(launch_k(nchw_to_act_kernel<T>, flat_grid(1ll * B * H * W), 256, 0, static_cast<cudaStream_t>(stream), x, static_cast<T*>(y), B, C,
                             H, W, hp, wp, ph, pw));
    DMLB_LAUNCH_CHECK();
    return 0;
}
extern "C" int dmlb_nchw_to_act_bf16(const float* x, void* y, int B, int C, int H, int W, int hp, int wp, int ph, int pw, void* stream) {
    return nchw_to_act_t<__nv_bfloat16>(x, y, B, C, H, W, hp, wp, ph, pw, stream);
}
extern "C" int dmlb_nchw_to_act_f32(const float* x, void* y, int B, int C, int H, int W, int hp, int wp, int ph, int pw, void* stream) {
    return nchw_to_act_t<float>(x, y, B, C, H, W, hp, wp, ph, pw, stream);
}

namespace dmlb {
// params = pulled + (params - snapshot);  snapshot = pulled;  shadow = bf16(params)
__global__ void __launch_bounds__(256) land_pull_kernel(float* __restrict__ params, const float* __restrict__ pulled, float* __restrict__ snapshot,
                                                        __nv_bfloat16* __restrict__ shadow, long long n4) {
    pdl_prologue();
    for (long long i = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i < n4; i += 1ll * gridDim.x * blockDim.x) {
        const float4 s = ld_stream_f32x4(pulled + 4 * i);
        const float4 q = reinterpret_cast<const float4*>(snapshot)[i];
        float4 p = reinterpret_cast<const float4*>(params)[i];
        p.x = s.x + (p.x - q.x); p.y = s.y + (p.y - q.y); p.z = s.z + (p.z - q.z); p.w = s.w + (p.w - q.w);
        reinterpret_cast<float4*>(params)[i] = p;
        reinterpret_cast<float4*>(snapshot)[i] = s;
        if (shadow != nullptr) {
            uint2 h;
            h.x = pack_bf16x2(p.x, p.y);
            h.y = pack_bf16x2(p.z, p.w);
            reinterpret_cast<uint2*>(shadow)[i] = h;
        }
    }
}
}  // namespace dmlb

// Landing of a fetched pull in the delta scheme (see dmlb_push_delta): the replica is re-based on the server's parameters while keeping
// exactly the local updates the server has not seen, params = pulled + (params - snapshot), and the snapshot becomes the server value.
// Reference: the listener overwrites the model with the server's reply (Asynchronous.py:17-18) and thereby drops those updates from the replica.
extern "C" int dmlb_land_pull(float* params, const float* pulled, float* snapshot, void* shadow_bf16, long long n, void* stream) {
    DMLB_REQUIRE(n % 4 == 0 && params != nullptr && pulled != nullptr && snapshot != nullptr, "bad arguments");
    DMLB_CHECK_CUDA(launch_k(land_pull_kernel, flat_grid(n / 4), 256, 0, static_cast<cudaStream_t>(stream), params, pulled, snapshot,
                             static_cast<__nv_bfloat16*>(shadow_bf16), n / 4));
    DMLB_LAUNCH_CHECK();
    return 0;
}


// table: device array of n_entries ShadowEntry {int64 src_off, int64 dst_off, int32 rows, cin, khw, ld, cpad, src_gemm_order}.
extern "C" int dmlb_refresh_shadows(const float* params, void* shadow, const void* table, int n_entries, void* stream) {
    if (n_entries <= 0) return 0;
    dim3 grid(96, n_entries);
    DMLB_CHECK_CUDA(launch_k(refresh_shadows_kernel<__nv_bfloat16>, grid, 256, 0, static_cast<cudaStream_t>(stream), params, static_cast<__nv_bfloat16*>(shadow),
                                                                               static_cast<const ShadowEntry*>(table)));
    DMLB_LAUNCH_CHECK();
    return 0;
}
// fp32 GEMM-layout copies (tf32 path): only weights whose arena storage is not already a legal TMA operand need one
// (row length not a multiple of 4, PyTorch (cin,kh,kw) order, in_map linears); dst_off counts fp32 elements.
extern "C" int dmlb_refresh_shadows_f32(const float* params, void* shadow, const void* table, int n_entries, void* stream) {
    if (n_entries <= 0) return 0;
    dim3 grid(96, n_entries);
    DMLB_CHECK_CUDA(launch_k(refresh_shadows_kernel<float>, grid, 256, 0, static_cast<cudaStream_t>(stream), params, static_cast<float*>(shadow),
                             static_cast<const ShadowEntry*>(table)));
    DMLB_LAUNCH_CHECK();
    return 0;
}

extern "C" int dmlb_cast_f32_bf16(const float* x, void* y, long long rows, int C, long long ldx, int ldy, void* stream) {
    DMLB_CHECK_CUDA(launch_k(cast_f32_bf16_kernel, flat_grid(rows * C), 256, 0, static_cast<cudaStream_t>(stream), x, static_cast<__nv_bfloat16*>(y), rows, C, ldx, ldy));
    DMLB_LAUNCH_CHECK();
    return 0;
}
extern "C" int dmlb_cast_bf16_f32(const void* x, float* y, long long rows, int C, int ldx, long long ldy, void* stream) {
    DMLB_CHECK_CUDA(launch_k(cast_bf16_f32_kernel, flat_grid(rows * C), 256, 0, static_cast<cudaStream_t>(stream), static_cast<const __nv_bfloat16*>(x), y, rows, C, ldx, ldy));
    DMLB_LAUNCH_CHECK();
    return 0;
}

template <typename T>
static int dropout_fwd_t(const void* x, void* y, void* mask, long long rows, int C, int ldx, int ldy, int rows_per_sample, int channel_mode, float p,
                         unsigned long long seed, const void* step, void* stream) {
    DMLB_REQUIRE(p >= 0.f && p < 1.f, "dropout probability must be in [0, 1)");
    DMLB_CHECK_CUDA(launch_k(dropout_fwd_kernel<T>, flat_grid(rows * C), 256, 0, static_cast<cudaStream_t>(stream), static_cast<const T*>(x),
                             static_cast<T*>(y), static_cast<uint8_t*>(mask), rows, C, ldx, ldy, rows_per_sample, channel_mode, p, seed,
                             static_cast<const unsigned long long*>(step)));
    DMLB_LAUNCH_CHECK();
    return 0;
}
template <typename T>
static int dropout_bwd_t(const void* dy, const void* mask, void* dx, long long rows, int C, int lddy, int lddx, float p, void* stream) {
    DMLB_CHECK_CUDA(launch_k(dropout_bwd_kernel<T>, flat_grid(rows * C), 256, 0, static_cast<cudaStream_t>(stream), static_cast<const T*>(dy),
                             static_cast<const uint8_t*>(mask), static_cast<T*>(dx), rows, C, lddy, lddx, p));
    DMLB_LAUNCH_CHECK();
    return 0;
}
extern "C" int dmlb_dropout_fwd_bf16(const void* x, void* y, void* mask, long long rows, int C, int ldx, int ldy, int rows_per_sample,
                                     int channel_mode, float p, unsigned long long seed, const void* step, void* stream) {
    return dropout_fwd_t<__nv_bfloat16>(x, y, mask, rows, C, ldx, ldy, rows_per_sample, channel_mode, p, seed, step, stream);
}
extern "C" int dmlb_dropout_fwd_f32(const void* x, void* y, void* mask, long long rows, int C, int ldx, int ldy, int rows_per_sample,
                                    int channel_mode, float p, unsigned long long seed, const void* step, void* stream) {
    return dropout_fwd_t<float>(x, y, mask, rows, C, ldx, ldy, rows_per_sample, channel_mode, p, seed, step, stream);
}
extern "C" int dmlb_dropout_bwd_bf16(const void* dy, const void* mask, void* dx, long long rows, int C, int lddy, int lddx, float p, void* stream) {
    return dropout_bwd_t<__nv_bfloat16>(dy, mask, dx, rows, C, lddy, lddx, p, stream);
}
extern "C" int dmlb_dropout_bwd_f32(const void* dy, const void* mask, void* dx, long long rows, int C, int lddy, int lddx, float p, void* stream) {
    return dropout_bwd_t<float>(dy, mask, dx, rows, C, lddy, lddx, p, stream);
}
extern "C" int dmlb_counter_bump(void* counter, void* stream) {
    DMLB_CHECK_CUDA(launch_k(counter_bump_kernel, 1, 1, 0, static_cast<cudaStream_t>(stream), static_cast<unsigned long long*>(counter)));
    DMLB_LAUNCH_CHECK();
    return 0;
}

// dmlb_gemm_bias_act_splitk_noalloc — the split-K forward GEMM of splitk.cu for callers that cannot allocate.
//
// dmlb_gemm_bias_act_splitk needs an fp32 workspace that somebody has to cudaMalloc.  Some users of this library
// have nobody to ask: a C or C++ host that links libdmlb without torch and owns no allocator; a service that
// hands the whole GPU to its own memory pool at start-up, after which a late cudaMalloc can simply fail; a
// latency-critical loop that cannot afford the device-wide synchronisation a cudaMalloc / cudaFree implies.
// For them the workspace lives here, in a fixed array inside the library image.  Inputs and outputs are still
// the caller's.
//
// What that means for you as a caller:
//   * Capacity is fixed: M * round_up(N, 4) <= DMLB_NOALLOC_SCRATCH_FLOATS (1 Mi floats = 4 MiB).  Larger problems
//     are rejected with an error, never truncated.  Every GPU that loads the library carries these 4 MiB until
//     the process exits, whether or not this entry point is ever called.
//   * The first call on each device looks up the array's device address (cudaGetSymbolAddress with that device
//     current).  Make that first call during start-up.
//   * This function calls no cudaMalloc, cudaFree or any allocating library routine (torch's included).
//   * All calls on a device share the one array, and it keeps whatever the previous call left in it.  Each call
//     therefore clears the part it is about to use (on `stream`) before accumulating into it.  Two calls whose
//     GPU work overlaps — different streams without an ordering between them, or different host threads — race
//     on the array and produce wrong sums.  Keep the calls of one device on one stream, or order them yourself.
//
// The reference has nothing like this (it has no native code at all).
#include <mutex>

#include "common.cuh"

#define DMLB_NOALLOC_SCRATCH_FLOATS (1 << 20)

extern "C" int dmlb_gemm_bf16(const void* A, int lda, int a_mn_major, const void* B, int ldb, int b_mn_major, int M, int N, int K,
                              void* out, int out_is_f32, int ldo, const float* bias, int relu, int split_k, int atomic, int perm_cin,
                              int perm_khw, int tile_n, void* stream);
extern "C" int dmlb_splitk_finish(float* ws, int ldw, const float* bias, int relu, void* y, int ldy, long long M, int N, int pool2, int B,
                                  int H, int W, void* pooled, int ldp, void* stream);

// File-scope device array.  In host code this name is only a stand-in for the symbol — NOT a device address;
// host code must go through cudaGetSymbolAddress (below) and hand the returned pointer to kernels.
static __device__ __align__(256) float g_noalloc_scratch[DMLB_NOALLOC_SCRATCH_FLOATS];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_setup_mu;
float* g_scratch_addr[kMaxDevices] = {nullptr};

// Once per device, with that device current: resolve the device address of g_noalloc_scratch.
int scratch_for_current_device(float** out) {
    int dev = -1;
    DMLB_CHECK_CUDA(cudaGetDevice(&dev));
    DMLB_REQUIRE(dev >= 0 && dev < kMaxDevices, "device ordinal out of range for the no-alloc scratch table");
    std::lock_guard<std::mutex> lk(g_setup_mu);
    if (g_scratch_addr[dev] == nullptr) {
        void* p = nullptr;
        DMLB_CHECK_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        DMLB_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned a null address for the no-alloc scratch");
        g_scratch_addr[dev] = static_cast<float*>(p);
    }
    *out = g_scratch_addr[dev];
    return 0;
}
}  // namespace

extern "C" long long dmlb_noalloc_scratch_capacity() { return DMLB_NOALLOC_SCRATCH_FLOATS; }

extern "C" int dmlb_gemm_bias_act_splitk_noalloc(const void* A, int lda, const void* Wt, int ldb, int M, int N, int K, const float* bias,
                                                 int relu, void* y, int ldy, int split_k, void* stream) {
    DMLB_REQUIRE(M > 0 && N > 0 && K > 0, "empty GEMM");
    const long long ldw = (static_cast<long long>(N) + 3) / 4 * 4;
    const long long need = static_cast<long long>(M) * ldw;
    if (need > DMLB_NOALLOC_SCRATCH_FLOATS) {
        char buf[256];
        snprintf(buf, sizeof(buf),
                 "dmlb_gemm_bias_act_splitk_noalloc: M*round_up(N,4) = %lld floats exceeds the built-in scratch of %d floats; "
                 "use dmlb_gemm_bias_act_splitk with your own workspace",
                 need, DMLB_NOALLOC_SCRATCH_FLOATS);
        dmlb_set_error(buf);
        return -5;
    }
    float* ws = nullptr;
    int rc = scratch_for_current_device(&ws);
    if (rc) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    // The array still holds whatever the last call (or a failed one) left behind: start this call's region from zero.
    DMLB_CHECK_CUDA(cudaMemsetAsync(ws, 0, static_cast<size_t>(need) * sizeof(float), st));
    rc = dmlb_gemm_bf16(A, lda, 0, Wt, ldb, 0, M, N, K, ws, 1, static_cast<int>(ldw), nullptr, 0, split_k, 1, 0, 0, 0, stream);
    if (rc) return rc;
    return dmlb_splitk_finish(ws, static_cast<int>(ldw), bias, relu, y, ldy, M, N, 0, 0, 0, 0, nullptr, 0, stream);
}

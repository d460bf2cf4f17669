// Host-side runtime glue shared by every kernel launcher: last-error string, kernel-launch
// accounting (bench.py reports "gpu_launches" from it) and device queries.
// No counterpart in the reference (it has no native code).
#include <stdlib.h>

#include <atomic>
#include <mutex>
#include <string>

#include "common.cuh"

namespace {
std::mutex g_err_mu;
std::string g_last_error;
std::atomic<long long> g_launches{0};
}  // namespace

extern "C" void dmlb_set_error(const char* msg) {
    std::lock_guard<std::mutex> lk(g_err_mu);
    g_last_error = msg ? msg : "";
}

extern "C" const char* dmlb_last_error() {
    std::lock_guard<std::mutex> lk(g_err_mu);
    static thread_local std::string copy;
    copy = g_last_error;
    return copy.c_str();
}

extern "C" void dmlb_count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
extern "C" long long dmlb_launch_count() { return g_launches.load(std::memory_order_relaxed); }
extern "C" void dmlb_reset_launch_count() { g_launches.store(0, std::memory_order_relaxed); }

extern "C" int dmlb_abi_version() { return 1; }

// Programmatic dependent launch switch (default on; DMLB_PDL=0 in the environment or dmlb_set_pdl(0) turns it off).
static int g_pdl = -1;
extern "C" int dmlb_pdl_enabled() {
    if (g_pdl < 0) {
        const char* e = getenv("DMLB_PDL");
        g_pdl = (e != nullptr && e[0] == '0') ? 0 : 1;
    }
    return g_pdl;
}
extern "C" void dmlb_set_pdl(int on) { g_pdl = on ? 1 : 0; }

// Persistent GEMM for multi-wave, short-K tile grids.  Opt-in (DMLB_PERSISTENT=1 or dmlb_set_gemm_persistent(1)): measured in
// isolation it wins (M=200704,N=64,K=576: 61 -> 52 us, 0.76 of HBM peak) but inside a ResNet-50 step it loses (9.6 -> 10.4 ms)
// because a grid that pins all 148 SMs serialises against the side-stream wgrad GEMMs that otherwise run concurrently.
static int g_persist = -1;
extern "C" int dmlb_gemm_persistent_enabled() {
    if (g_persist < 0) {
        const char* e = getenv("DMLB_PERSISTENT");
        g_persist = (e != nullptr && e[0] == '1') ? 1 : 0;
    }
    return g_persist;
}
extern "C" void dmlb_set_gemm_persistent(int on) { g_persist = on ? 1 : 0; }

// CTA-pair (cta_group::2) GEMM for 256-wide tiles: 1 = on (default), 0 = off, 2 = on but never the persistent form
// (env DMLB_PAIR or dmlb_set_gemm_pair).
static int g_pair = -1;
extern "C" int dmlb_gemm_pair_enabled() {
    if (g_pair < 0) {
        const char* e = getenv("DMLB_PAIR");
        g_pair = (e != nullptr && e[0] == '0') ? 0 : ((e != nullptr && e[0] == '2') ? 2 : 1);
    }
    return g_pair;
}
extern "C" void dmlb_set_gemm_pair(int on) { g_pair = on < 0 ? 0 : (on > 2 ? 1 : on); }

// Number of SMs of the current device (148 on B200); used to size persistent / flat grids.
extern "C" int dmlb_sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

// Fused communication + update kernels over NVLink peer memory (one process per GPU, memory mapped with
// CUDA IPC by peer.cu).  These replace the reference's message-based parameter-server traffic:
//
//   GradientUpdate  send of N floats + server `theta += payload`    asgd/optim/Asynchronous.py:55,58-60  -> push_add (K1)
//   ParameterRequest (N dummy floats up) + ParameterUpdate reply    asgd/optim/Asynchronous.py:17-18,48-49 -> pull (K3)
//   NORTH-STAR momentum / staleness-scaled server update                                               -> push_mailbox + owner_apply (K2)
//   NORTH-STAR synchronous mode (all-reduce + SGD)                                                      -> sync_rs_update_ag (K4)
//
// The flat parameter vector [0, n) is cut into `world` contiguous shards of `shard_size` elements; GPU g owns
// shard g (the "server" for that slice).  No kernel ever spins on, or reads flags from, remote memory in the
// asynchronous paths: every cross-GPU flag is *written* remotely and *polled* locally.
#include "common.cuh"

namespace dmlb {

constexpr int kMaxPeers = 16;
constexpr int kChunk = 4096;  // elements per mailbox chunk (16 KiB)

struct PeerTable {
    void* ptr[kMaxPeers];
};

// ---------------------------------------------------------------------------------------------
// K1  push_add: master[owner(i)][i] += acc[i];  acc[i] = 0          (exact reference semantics)
// Vector fp32 reductions (red.global.add.v4.f32, .sys scope) travel over NVLink and are applied by the
// owner's L2; the worker never waits for an acknowledgement.
// ---------------------------------------------------------------------------------------------
// Delta form (live != nullptr): no accumulator is kept at all.  `acc` is then a SNAPSHOT of the replica as of the last push, the update to
// send is live - snapshot (exact: both are the same fp32 variable at two times), and the snapshot is advanced to the value just read.
// It is safe to run while training keeps writing `live`: each element is read once, so every local update is sent exactly once.
__global__ void __launch_bounds__(256) push_add_kernel(float* __restrict__ acc, const float* __restrict__ live, PeerTable masters,
                                                       long long shard_size, long long n4) {
    // Runs with a small grid next to the training step: keep kUnroll independent 16-byte loads (x2 in the delta form) in flight per thread.
    constexpr int kUnroll = 4;
    const long long stride = 1ll * gridDim.x * blockDim.x;
    for (long long i0 = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i0 < n4; i0 += stride * kUnroll) {
        float4 a[kUnroll], p[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const long long i = i0 + u * stride;
            if (i < n4) {
                a[u] = reinterpret_cast<float4*>(acc)[i];
                if (live != nullptr) p[u] = ld_volatile_f32x4(live + 4 * i);
            }
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const long long i = i0 + u * stride;
            if (i >= n4) continue;
            const long long e = i * 4;
            const int owner = static_cast<int>(e / shard_size);
            float* dst = static_cast<float*>(masters.ptr[owner]) + (e - owner * shard_size);
            float4 d = a[u];
            if (live != nullptr) {
                reinterpret_cast<float4*>(acc)[i] = p[u];
                d.x = p[u].x - d.x; d.y = p[u].y - d.y; d.z = p[u].z - d.z; d.w = p[u].w - d.w;
            } else {
                reinterpret_cast<float4*>(acc)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
            if (d.x != 0.f || d.y != 0.f || d.z != 0.f || d.w != 0.f) red_add_f32x4_sys(dst, d.x, d.y, d.z, d.w);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// K3  pull: params[i] = master[owner(i)][i], shadow[i] = bf16(params[i])  (+ optional version snapshot for staleness)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) pull_kernel(float* __restrict__ params, __nv_bfloat16* __restrict__ shadow, PeerTable masters,
                                                   PeerTable versions, uint32_t* __restrict__ pulled_ver, long long shard_size, long long n4,
                                                   int chunks_per_shard) {
    // Peer loads cost an NVLink round trip (~2 us): keep kUnroll independent 16-byte loads in flight per thread.
    constexpr int kUnroll = 8;
    const long long stride = 1ll * gridDim.x * blockDim.x;
    for (long long i0 = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i0 < n4; i0 += stride * kUnroll) {
        float4 v[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const long long i = i0 + u * stride;
            if (i < n4) {
                const long long e = i * 4;
                const int owner = static_cast<int>(e / shard_size);
                v[u] = ld_volatile_f32x4(static_cast<const float*>(masters.ptr[owner]) + (e - owner * shard_size));
            }
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const long long i = i0 + u * stride;
            if (i < n4) {
                reinterpret_cast<float4*>(params)[i] = v[u];
                if (shadow != nullptr) {   // fused cast into the worker's bf16 tensor-core weights (same flat offsets)
                    uint2 h;
                    h.x = pack_bf16x2(v[u].x, v[u].y);
                    h.y = pack_bf16x2(v[u].z, v[u].w);
                    reinterpret_cast<uint2*>(shadow)[i] = h;
                }
                if (pulled_ver != nullptr) {
                    const long long e = i * 4;
                    const int owner = static_cast<int>(e / shard_size);
                    const long long off = e - owner * shard_size;
                    if ((off % kChunk) == 0) {
                        const int lc = static_cast<int>(off / kChunk);
                        pulled_ver[owner * chunks_per_shard + lc] = ld_relaxed_sys_u32(static_cast<const uint32_t*>(versions.ptr[owner]) + lc);
                    }
                }
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// K2a push_mailbox: per 16 KiB chunk, if the worker's slot at the owner is free, store the accumulated
// update into the owner's mailbox, zero the accumulator and raise the owner's `full` flag carrying
// (version the worker last pulled + 1).  Busy chunks simply keep accumulating — nothing blocks.
//   busy   : local  uint32 [world * cps]      (set here, cleared remotely by the owner)
//   full   : owner's uint32 [world][cps]      (set remotely here, cleared by the owner)
//   mailbox: owner's float  [world][shard_size]
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) push_mailbox_kernel(float* __restrict__ acc, const float* __restrict__ live, uint32_t* __restrict__ busy,
                                                           const uint32_t* __restrict__ pulled_ver, PeerTable mailboxes, PeerTable fulls,
                                                           int rank, int world, long long shard_size, int cps) {
    __shared__ uint32_t s_busy;
    const int total_chunks = world * cps;
    for (int c = blockIdx.x; c < total_chunks; c += gridDim.x) {
        if (threadIdx.x == 0) s_busy = busy[c];
        __syncthreads();
        const bool is_busy = s_busy != 0;
        __syncthreads();
        if (is_busy) continue;
        const int owner = c / cps, lc = c % cps;
        float* src = acc + static_cast<long long>(owner) * shard_size + static_cast<long long>(lc) * kChunk;
        float* dst = static_cast<float*>(mailboxes.ptr[owner]) + static_cast<long long>(rank) * shard_size + static_cast<long long>(lc) * kChunk;
        const float* lsrc = live != nullptr ? live + static_cast<long long>(owner) * shard_size + static_cast<long long>(lc) * kChunk : nullptr;
        for (int j = threadIdx.x; j < kChunk / 4; j += blockDim.x) {
            float4 a = reinterpret_cast<float4*>(src)[j];
            if (lsrc != nullptr) {            // delta form: send live - snapshot, advance the snapshot (see push_add_kernel)
                const float4 p = ld_volatile_f32x4(lsrc + 4 * j);
                reinterpret_cast<float4*>(src)[j] = p;
                a.x = p.x - a.x; a.y = p.y - a.y; a.z = p.z - a.z; a.w = p.w - a.w;
            } else {
                reinterpret_cast<float4*>(src)[j] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
            reinterpret_cast<float4*>(dst)[j] = a;
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            busy[c] = 1;
            __threadfence_system();
            st_release_sys_u32(static_cast<uint32_t*>(fulls.ptr[owner]) + rank * cps + lc, pulled_ver[c] + 1u);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// K2b owner_apply: the shard owner folds every full mailbox chunk into its master shard:
//   tau = version[chunk] - version_pulled;  s = staleness scale(tau)
//   v = mu * v + s * delta;  theta += v;  version[chunk] += 1
// then frees the slot (local `full` = 0, remote `busy` = 0 at the worker).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float staleness_scale(int mode, float param, uint32_t tau) {
    if (mode == 1) return 1.f / (1.f + static_cast<float>(tau));           // inverse
    if (mode == 2) return __powf(param, static_cast<float>(tau));          // exponential decay, param in (0,1]
    return 1.f;
}

__global__ void __launch_bounds__(256) owner_apply_kernel(float* __restrict__ master, float* __restrict__ mom, uint32_t* __restrict__ version,
                                                          uint32_t* __restrict__ full, const float* __restrict__ mailbox, PeerTable busys,
                                                          uint32_t* __restrict__ stats, int rank, int world, long long shard_size, int cps,
                                                          float mu, int stale_mode, float stale_param, float gain) {
    __shared__ uint32_t s_flag;
    for (int lc = blockIdx.x; lc < cps; lc += gridDim.x) {
        for (int w = 0; w < world; ++w) {
            if (threadIdx.x == 0) s_flag = ld_acquire_sys_u32(full + w * cps + lc);
            __syncthreads();
            const uint32_t flag = s_flag;
            __syncthreads();
            if (flag == 0) continue;
            const uint32_t ver = version[lc];
            const uint32_t tau = ver - (flag - 1u);
            const float s = gain * staleness_scale(stale_mode, stale_param, tau);
            const float* src = mailbox + static_cast<long long>(w) * shard_size + static_cast<long long>(lc) * kChunk;
            float* th = master + static_cast<long long>(lc) * kChunk;
            for (int j = threadIdx.x; j < kChunk / 4; j += blockDim.x) {
                const float4 d = ld_volatile_f32x4(src + 4 * j);
                float4 t = reinterpret_cast<float4*>(th)[j];
                if (mom != nullptr) {
                    float4 v = reinterpret_cast<float4*>(mom + static_cast<long long>(lc) * kChunk)[j];
                    v.x = mu * v.x + s * d.x; v.y = mu * v.y + s * d.y; v.z = mu * v.z + s * d.z; v.w = mu * v.w + s * d.w;
                    reinterpret_cast<float4*>(mom + static_cast<long long>(lc) * kChunk)[j] = v;
                    t.x += v.x; t.y += v.y; t.z += v.z; t.w += v.w;
                } else {
                    t.x += s * d.x; t.y += s * d.y; t.z += s * d.z; t.w += s * d.w;
                }
                reinterpret_cast<float4*>(th)[j] = t;
            }
            __syncthreads();
            if (threadIdx.x == 0) {
                version[lc] = ver + 1u;
                full[w * cps + lc] = 0u;
                if (stats != nullptr) {
                    atomicAdd(&stats[0], 1u);                              // chunks applied
                    atomicAdd(&stats[1], tau);                             // sum of staleness
                    atomicMax(&stats[2], tau);                             // max staleness
                }
                __threadfence_system();
                st_release_sys_u32(static_cast<uint32_t*>(busys.ptr[w]) + rank * cps + lc, 0u);
            }
            __syncthreads();
        }
    }
}

// ---------------------------------------------------------------------------------------------
// K4  sync_rs_update_ag: one kernel per step and GPU does
//   barrier(all grads ready) -> owner reduces its shard over every replica's gradient arena (peer loads)
//   -> SGD / momentum on the fp32 shard -> stores the new weights into every replica's parameter arena
//   (peer stores, the all-gather) -> barrier(all replicas written).
// flags layout per GPU (uint32): [0, kMaxPeers) start flags, [kMaxPeers, 2*kMaxPeers) done flags,
// [2*kMaxPeers] epoch, [2*kMaxPeers + 1] local block counter.   Every flag is written remotely and polled locally.
// ---------------------------------------------------------------------------------------------
// Every rank must reach the same bucket launch within `limit` cycles of the first one (host: DMLB_BARRIER_TIMEOUT_S, default 1800 s —
// the process-group timeout of example/main.py): a peer that stalls in its data loader or a checkpoint write only delays the others;
// a peer that died turns into a CUDA error on the survivors after the limit instead of a hang.
__device__ __forceinline__ void spin_until_ge(const uint32_t* flag, uint32_t want, long long limit) {
    const long long t0 = clock64();
    while (ld_acquire_sys_u32(flag) < want) {
        if (clock64() - t0 > limit) {
            printf("dmlb: cross-GPU barrier timed out (block %d thread %d want %u have %u)\n", blockIdx.x, threadIdx.x, want, *flag);
            __trap();
        }
    }
}

__global__ void __launch_bounds__(512) sync_rs_update_ag_kernel(PeerTable grads, PeerTable params, PeerTable flags, float* __restrict__ mom,
                                                                int rank, int world, long long base, long long slice_len,
                                                                const float* __restrict__ lr_ptr, float mu, float wd, float grad_scale,
                                                                long long spin_limit) {
    uint32_t* my_flags = static_cast<uint32_t*>(flags.ptr[rank]);
    const uint32_t epoch = my_flags[2 * kMaxPeers] + 1u;   // bumped once per launch by the last block, below
    // ---- start barrier: tell every peer that my gradients are final, wait for theirs ----
    if (blockIdx.x == 0 && threadIdx.x < world) {
        st_release_sys_u32(static_cast<uint32_t*>(flags.ptr[threadIdx.x]) + rank, epoch);
    }
    if (threadIdx.x < world) spin_until_ge(my_flags + threadIdx.x, epoch, spin_limit);
    __syncthreads();

    const float lr = *lr_ptr;
    // [base, base + slice_len) is the slice of the flat vector this GPU owns in this launch (the whole shard, or this GPU's
    // part of one gradient bucket when the update is overlapped with the backward pass)
    const long long n4 = slice_len / 4;
    for (long long i = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i < n4; i += 1ll * gridDim.x * blockDim.x) {
        float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
        float4 v[kMaxPeers];
#pragma unroll
        for (int r = 0; r < kMaxPeers; ++r)   // issue every peer's load before consuming any: one NVLink round trip, not `world`
            if (r < world) v[r] = ld_volatile_f32x4(static_cast<const float*>(grads.ptr[r]) + base + 4 * i);
#pragma unroll
        for (int r = 0; r < kMaxPeers; ++r)
            if (r < world) { g.x += v[r].x; g.y += v[r].y; g.z += v[r].z; g.w += v[r].w; }
        g.x *= grad_scale; g.y *= grad_scale; g.z *= grad_scale; g.w *= grad_scale;
        float4 p = reinterpret_cast<const float4*>(static_cast<const float*>(params.ptr[rank]) + base)[i];
        if (wd != 0.f) { g.x += wd * p.x; g.y += wd * p.y; g.z += wd * p.z; g.w += wd * p.w; }
        if (mom != nullptr) {
            float4 v = reinterpret_cast<float4*>(mom)[i];
            v.x = mu * v.x + g.x; v.y = mu * v.y + g.y; v.z = mu * v.z + g.z; v.w = mu * v.w + g.w;
            reinterpret_cast<float4*>(mom)[i] = v;
            g = v;
        }
        p.x -= lr * g.x; p.y -= lr * g.y; p.z -= lr * g.z; p.w -= lr * g.w;
#pragma unroll 4
        for (int r = 0; r < world; ++r) reinterpret_cast<float4*>(static_cast<float*>(params.ptr[r]) + base)[i] = p;
    }

    // ---- end barrier: my stores to every replica are done; wait until every owner has written into mine ----
    __syncthreads();
    __shared__ bool s_last;
    if (threadIdx.x == 0) {
        __threadfence_system();
        const uint32_t done = atomicAdd(&my_flags[2 * kMaxPeers + 1], 1u);
        s_last = (done == gridDim.x - 1);
    }
    __syncthreads();
    if (s_last) {
        if (threadIdx.x < world) {
            st_release_sys_u32(static_cast<uint32_t*>(flags.ptr[threadIdx.x]) + kMaxPeers + rank, epoch);
            spin_until_ge(my_flags + kMaxPeers + threadIdx.x, epoch, spin_limit);
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            my_flags[2 * kMaxPeers + 1] = 0u;
            my_flags[2 * kMaxPeers] = epoch;
            __threadfence();
        }
    }
}

// K4 through NVLS: the reduce-scatter is ONE multimem.ld_reduce per 16 bytes (the NVSwitch adds the replicas) and the all-gather
// ONE multimem.st (the switch replicates it), so each GPU moves n/world in and n/world out instead of (world-1)/world * n each way.
//   mc_grads / mc_params: multicast addresses of the gradient / parameter arenas;  params_local: this GPU's own replica.
__global__ void __launch_bounds__(512) sync_nvls_kernel(const float* __restrict__ mc_grads, float* __restrict__ mc_params,
                                                        const float* __restrict__ params_local, PeerTable flags, float* __restrict__ mom, int rank,
                                                        int world, long long base, long long slice_len, const float* __restrict__ lr_ptr, float mu,
                                                        float wd, float grad_scale, long long spin_limit) {
    uint32_t* my_flags = static_cast<uint32_t*>(flags.ptr[rank]);
    const uint32_t epoch = my_flags[2 * kMaxPeers] + 1u;
    if (blockIdx.x == 0 && threadIdx.x < world) st_release_sys_u32(static_cast<uint32_t*>(flags.ptr[threadIdx.x]) + rank, epoch);
    if (threadIdx.x < world) spin_until_ge(my_flags + threadIdx.x, epoch, spin_limit);
    __syncthreads();

    const float lr = *lr_ptr;
    const long long n4 = slice_len / 4;
    constexpr int kU = 4;     // switch-reduced loads in flight per thread
    const long long stride = 1ll * gridDim.x * blockDim.x;
    for (long long i0 = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i0 < n4; i0 += stride * kU) {
        float4 gs[kU];
#pragma unroll
        for (int u = 0; u < kU; ++u)
            if (i0 + u * stride < n4) gs[u] = multimem_ld_reduce_add_f32x4(mc_grads + base + 4 * (i0 + u * stride));
#pragma unroll
        for (int u = 0; u < kU; ++u) {
            const long long i = i0 + u * stride;
            if (i >= n4) continue;
            float4 g = gs[u];
            g.x *= grad_scale; g.y *= grad_scale; g.z *= grad_scale; g.w *= grad_scale;
            float4 p = reinterpret_cast<const float4*>(params_local + base)[i];
            if (wd != 0.f) { g.x += wd * p.x; g.y += wd * p.y; g.z += wd * p.z; g.w += wd * p.w; }
            if (mom != nullptr) {
                float4 v = reinterpret_cast<float4*>(mom)[i];
                v.x = mu * v.x + g.x; v.y = mu * v.y + g.y; v.z = mu * v.z + g.z; v.w = mu * v.w + g.w;
                reinterpret_cast<float4*>(mom)[i] = v;
                g = v;
            }
            p.x -= lr * g.x; p.y -= lr * g.y; p.z -= lr * g.z; p.w -= lr * g.w;
            multimem_st_f32x4(mc_params + base + 4 * i, p);
        }
    }

    __syncthreads();
    __shared__ bool s_last;
    if (threadIdx.x == 0) {
        __threadfence_system();
        const uint32_t done = atomicAdd(&my_flags[2 * kMaxPeers + 1], 1u);
        s_last = (done == gridDim.x - 1);
    }
    __syncthreads();
    if (s_last) {
        if (threadIdx.x < world) {
            st_release_sys_u32(static_cast<uint32_t*>(flags.ptr[threadIdx.x]) + kMaxPeers + rank, epoch);
            spin_until_ge(my_flags + kMaxPeers + threadIdx.x, epoch, spin_limit);
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            my_flags[2 * kMaxPeers + 1] = 0u;
            my_flags[2 * kMaxPeers] = epoch;
            __threadfence();
        }
    }
}

// ---------------------------------------------------------------------------------------------
// TMA bulk forms of K1 / K3 (1-D `cp.async.bulk`, no tensor map): one thread per CTA moves 16 KiB tiles through a shared-memory ring.
//   pull: peer master --cp.async.bulk--> smem --cp.async.bulk--> local params (+ bf16 mirror written by the CTA's threads)
//   push: local accumulator (or replica + snapshot) --cp.async.bulk--> smem --cp.reduce.async.bulk.add.f32--> peer master
// A tile is one mailbox chunk, so it never straddles two owners.  Each CTA walks a CONTIGUOUS run of tiles: at any moment the grid
// touches ~one 2 MiB page per CTA on the remote side instead of thousands (the per-thread grid-stride movers above spread every warp
// over the whole vector, which is what made them fall off above ~256 MB).  None of these loads uses multicast.
// ---------------------------------------------------------------------------------------------
constexpr int kBulkStages = 4;                  // pull, push: 4 x 16 KiB tiles in flight per CTA
constexpr int kBulkStagesDelta = 2;             // delta push: 2 x (snapshot tile + replica tile)
constexpr uint32_t kTileBytes = kChunk * 4;

__device__ __forceinline__ void bulk_g2s(uint32_t dst_smem, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_smem), "l"(src), "r"(bytes),
                 "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* dst, uint32_t src_smem, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src_smem), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_red_add_f32(void* dst, uint32_t src_smem, uint32_t bytes) {
    asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;" ::"l"(dst), "r"(src_smem), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait() { asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ void tile_run(long long n_tiles, long long* t0, long long* t1) {
    const long long per = (n_tiles + gridDim.x - 1) / gridDim.x;
    *t0 = blockIdx.x * per;
    *t1 = *t0 + per < n_tiles ? *t0 + per : n_tiles;
}

__global__ void __launch_bounds__(128) pull_bulk_kernel(float* __restrict__ params, __nv_bfloat16* __restrict__ shadow, PeerTable masters,
                                                        PeerTable versions, uint32_t* __restrict__ pulled_ver, long long shard_size,
                                                        long long n_tiles, int chunks_per_shard) {
    extern __shared__ __align__(128) uint8_t bulk_smem[];
    __shared__ __align__(8) uint64_t full[kBulkStages];
    long long t0, t1;
    tile_run(n_tiles, &t0, &t1);
    if (t0 >= t1) return;
    const int tid = threadIdx.x;
    const uint32_t tiles = smem_u32(bulk_smem);
    if (tid == 0) {
        for (int s = 0; s < kBulkStages; ++s) mbar_init(smem_u32(&full[s]), 1);
        fence_mbar_init();
    }
    __syncthreads();
    auto issue = [&](long long t) {
        const int stage = static_cast<int>((t - t0) % kBulkStages);
        const long long e = t * kChunk;
        const int owner = static_cast<int>(e / shard_size);
        const float* src = static_cast<const float*>(masters.ptr[owner]) + (e - owner * shard_size);
        mbar_arrive_expect_tx(smem_u32(&full[stage]), kTileBytes);
        bulk_g2s(tiles + stage * kTileBytes, src, kTileBytes, smem_u32(&full[stage]));
    };
    if (tid == 0)
        for (int k = 0; k < kBulkStages && t0 + k < t1; ++k) issue(t0 + k);
    const bool all_threads = shadow != nullptr;
    if (!all_threads && tid != 0) return;
    for (long long t = t0; t < t1; ++t) {
        const long long it = t - t0;
        const int stage = static_cast<int>(it % kBulkStages);
        mbar_wait(smem_u32(&full[stage]), static_cast<uint32_t>((it / kBulkStages) & 1));
        const long long e = t * kChunk;
        if (all_threads) {   // bf16 mirror for the tensor cores, same flat offsets
            const float4* src = reinterpret_cast<const float4*>(bulk_smem + stage * kTileBytes);
            for (int j = tid; j < kChunk / 4; j += blockDim.x) {
                const float4 v = src[j];
                uint2 h;
                h.x = pack_bf16x2(v.x, v.y);
                h.y = pack_bf16x2(v.z, v.w);
                reinterpret_cast<uint2*>(shadow + e)[j] = h;
            }
        }
        if (tid == 0) {
            bulk_s2g(params + e, tiles + stage * kTileBytes, kTileBytes);
            bulk_commit();
            if (pulled_ver != nullptr) {
                const int owner = static_cast<int>(e / shard_size);
                const int lc = static_cast<int>((e - owner * shard_size) / kChunk);
                pulled_ver[owner * chunks_per_shard + lc] = ld_relaxed_sys_u32(static_cast<const uint32_t*>(versions.ptr[owner]) + lc);
            }
        }
        if (all_threads) __syncthreads();            // everybody is done reading the PREVIOUS stage as well
        if (tid == 0 && it >= 1 && t - 1 + kBulkStages < t1) {
            bulk_wait_read<1>();                      // all stores but the newest have read their tile: the previous stage is free
            issue(t - 1 + kBulkStages);
        }
    }
    if (tid == 0) bulk_wait<0>();
}

// kDelta = false: master += acc, acc = 0.   kDelta = true: master += live - snapshot, snapshot = live (acc is the snapshot).
template <bool kDelta>
__global__ void __launch_bounds__(128) push_bulk_kernel(float* __restrict__ acc, const float* __restrict__ live, PeerTable masters,
                                                        long long shard_size, long long n_tiles) {
    extern __shared__ __align__(128) uint8_t bulk_smem[];
    constexpr int kStages = kDelta ? kBulkStagesDelta : kBulkStages;
    __shared__ __align__(8) uint64_t full[kStages];
    constexpr uint32_t kStageBytes = kDelta ? 2 * kTileBytes : kTileBytes;
    long long t0, t1;
    tile_run(n_tiles, &t0, &t1);
    if (t0 >= t1) return;
    const int tid = threadIdx.x;
    const uint32_t tiles = smem_u32(bulk_smem);
    uint8_t* zero_tile = bulk_smem + kStages * kStageBytes;     // kDelta = false only: source of "acc = 0"
    if (!kDelta) {
        for (int j = tid; j < kChunk / 4; j += blockDim.x) reinterpret_cast<float4*>(zero_tile)[j] = make_float4(0.f, 0.f, 0.f, 0.f);
        fence_proxy_async_smem();
    }
    if (tid == 0) {
        for (int s = 0; s < kStages; ++s) mbar_init(smem_u32(&full[s]), 1);
        fence_mbar_init();
    }
    __syncthreads();
    auto issue = [&](long long t) {
        const int stage = static_cast<int>((t - t0) % kStages);
        const long long e = t * kChunk;
        mbar_arrive_expect_tx(smem_u32(&full[stage]), kStageBytes);
        bulk_g2s(tiles + stage * kStageBytes, acc + e, kTileBytes, smem_u32(&full[stage]));
        if (kDelta) bulk_g2s(tiles + stage * kStageBytes + kTileBytes, live + e, kTileBytes, smem_u32(&full[stage]));
    };
    if (tid == 0)
        for (int k = 0; k < kStages && t0 + k < t1; ++k) issue(t0 + k);
    if (!kDelta && tid != 0) return;
    for (long long t = t0; t < t1; ++t) {
        const long long it = t - t0;
        const int stage = static_cast<int>(it % kStages);
        mbar_wait(smem_u32(&full[stage]), static_cast<uint32_t>((it / kStages) & 1));
        const long long e = t * kChunk;
        const int owner = static_cast<int>(e / shard_size);
        float* dst = static_cast<float*>(masters.ptr[owner]) + (e - owner * shard_size);
        if (kDelta) {
            float4* snap = reinterpret_cast<float4*>(bulk_smem + stage * kStageBytes);
            const float4* cur = reinterpret_cast<const float4*>(bulk_smem + stage * kStageBytes + kTileBytes);
            for (int j = tid; j < kChunk / 4; j += blockDim.x) {
                const float4 p = cur[j];
                float4 q = snap[j];
                q.x = p.x - q.x; q.y = p.y - q.y; q.z = p.z - q.z; q.w = p.w - q.w;
                snap[j] = q;
            }
            fence_proxy_async_smem();                 // the difference was written by threads, the reduction reads it through the async proxy
            __syncthreads();
        }
        if (tid == 0) {
            bulk_red_add_f32(dst, tiles + stage * kStageBytes, kTileBytes);
            if (kDelta) bulk_s2g(acc + e, tiles + stage * kStageBytes + kTileBytes, kTileBytes);       // snapshot = the value just sent
            else        bulk_s2g(acc + e, smem_u32(zero_tile), kTileBytes);
            bulk_commit();
            if (it >= 1 && t - 1 + kStages < t1) {
                bulk_wait_read<1>();
                issue(t - 1 + kStages);
            }
        }
    }
    if (tid == 0) bulk_wait<0>();
}

// Grid cap of the push / pull movers (0 = two CTAs per SM).  An optimizer that overlaps them with training sets a small cap: the transfer
// then trickles in the background of the next steps instead of taking every SM away from them for a few microseconds.
static int g_comm_max_blocks = 0;
extern "C" void dmlb_set_comm_max_blocks(int n) { g_comm_max_blocks = n > 0 ? n : 0; }

static inline int comm_grid(long long work_items, int threads) {
    long long blocks = (work_items + threads - 1) / threads;
    long long cap = dmlb_sm_count() * 2ll;
    if (g_comm_max_blocks > 0 && g_comm_max_blocks < cap) cap = g_comm_max_blocks;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return static_cast<int>(blocks);
}

// Cross-GPU barrier time-out in SM clock cycles: DMLB_BARRIER_TIMEOUT_S seconds (default 1800) at the device's peak clock.
static long long barrier_spin_limit() {
    static long long limit = 0;
    if (limit == 0) {
        double secs = 1800.0;
        const char* e = getenv("DMLB_BARRIER_TIMEOUT_S");
        if (e != nullptr && atof(e) > 0.0) secs = atof(e);
        int dev = 0, khz = 2000000;
        if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
        limit = static_cast<long long>(secs * 1e3 * static_cast<double>(khz));
    }
    return limit;
}

static int fill_table(PeerTable* t, void* const* ptrs, int world) {
    DMLB_REQUIRE(world >= 1 && world <= kMaxPeers, "world size must be in [1, 16]");
    for (int i = 0; i < kMaxPeers; ++i) t->ptr[i] = i < world ? ptrs[i] : nullptr;
    return 0;
}

}  // namespace dmlb

using namespace dmlb;

extern "C" int dmlb_comm_chunk_elems() { return kChunk; }
extern "C" int dmlb_comm_flag_words() { return 2 * kMaxPeers + 2; }

// Which mover dmlb_push_add / dmlb_push_delta / dmlb_pull use: 0 = by size (bulk from DMLB_BULK_MIN_BYTES, default 8 MiB, when the grid
// is not capped), 1 = per-thread ld / red, 2 = TMA bulk.
static int g_comm_mover = 0;
extern "C" void dmlb_set_comm_mover(int m) { g_comm_mover = (m >= 0 && m <= 2) ? m : 0; }

static bool use_bulk(long long n, long long shard_size) {
    if (shard_size % kChunk != 0 || n % kChunk != 0) return false;
    if (g_comm_mover != 0) return g_comm_mover == 2;
    static long long min_bytes = -1;
    if (min_bytes < 0) {
        const char* e = getenv("DMLB_BULK_MIN_BYTES");
        min_bytes = (e != nullptr && atoll(e) >= 0) ? atoll(e) : (8ll << 20);
    }
    return g_comm_max_blocks == 0 && n * 4 >= min_bytes;
}

// One CTA per SM: 64 KiB in flight per SM is several times what the link needs (770 GB/s x ~3 us / 148 SMs = 16 KiB); three CTAs per SM
// measured the same below 64 MB and 5-10 % slower above (2 GPUs).
static int bulk_grid(long long n_tiles) {
    long long cap = dmlb_sm_count();
    if (g_comm_max_blocks > 0 && g_comm_max_blocks < cap) cap = g_comm_max_blocks;
    return static_cast<int>(n_tiles < cap ? n_tiles : cap);
}

template <typename Kern>
static int bulk_smem_attr(Kern kern, int bytes) {
    DMLB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    return 0;
}

static int push_add_impl(float* acc, const float* live, void* const* master_ptrs, int world, long long shard_size, long long n, void* stream) {
    DMLB_REQUIRE(n % 4 == 0 && shard_size % 4 == 0, "lengths must be multiples of 4");
    PeerTable t;
    if (fill_table(&t, master_ptrs, world)) return -2;
    if (use_bulk(n, shard_size)) {
        const long long n_tiles = n / kChunk;
        if (live != nullptr) {
            const int smem = kBulkStagesDelta * 2 * kTileBytes;
            if (bulk_smem_attr(push_bulk_kernel<true>, smem)) return -1;
            push_bulk_kernel<true><<<bulk_grid(n_tiles), 128, smem, static_cast<cudaStream_t>(stream)>>>(acc, live, t, shard_size, n_tiles);
        } else {
            const int smem = (kBulkStages + 1) * kTileBytes;
            if (bulk_smem_attr(push_bulk_kernel<false>, smem)) return -1;
            push_bulk_kernel<false><<<bulk_grid(n_tiles), 128, smem, static_cast<cudaStream_t>(stream)>>>(acc, live, t, shard_size, n_tiles);
        }
        DMLB_LAUNCH_CHECK();
        return 0;
    }
    push_add_kernel<<<comm_grid(n / 4, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(acc, live, t, shard_size, n / 4);
    DMLB_LAUNCH_CHECK();
    return 0;
}

extern "C" int dmlb_push_add(float* acc, void* const* master_ptrs, int world, long long shard_size, long long n, void* stream) {
    return push_add_impl(acc, nullptr, master_ptrs, world, shard_size, n, stream);
}

// Delta push: master[owner][i] += live[i] - snapshot[i]; snapshot[i] = live[i].  The worker keeps no accumulator: what it has not pushed yet
// is the difference between its replica and the snapshot (reference: accumulated_gradients, Asynchronous.py:27,55,58-60).
extern "C" int dmlb_push_delta(const float* live, float* snapshot, void* const* master_ptrs, int world, long long shard_size, long long n,
                               void* stream) {
    DMLB_REQUIRE(live != nullptr && snapshot != nullptr, "live and snapshot are required");
    return push_add_impl(snapshot, live, master_ptrs, world, shard_size, n, stream);
}

extern "C" int dmlb_pull(float* params, void* shadow_bf16, void* const* master_ptrs, void* const* version_ptrs, void* pulled_ver, int world,
                         long long shard_size, long long n, void* stream) {
    DMLB_REQUIRE(n % 4 == 0 && shard_size % kChunk == 0, "shard size must be a multiple of the chunk size");
    PeerTable t, v;
    if (fill_table(&t, master_ptrs, world)) return -2;
    if (version_ptrs != nullptr) {
        if (fill_table(&v, version_ptrs, world)) return -2;
    } else {
        for (int i = 0; i < kMaxPeers; ++i) v.ptr[i] = nullptr;
        pulled_ver = nullptr;
    }
    if (use_bulk(n, shard_size)) {
        const long long n_tiles = n / kChunk;
        const int smem = kBulkStages * kTileBytes;
        if (bulk_smem_attr(pull_bulk_kernel, smem)) return -1;
        pull_bulk_kernel<<<bulk_grid(n_tiles), 128, smem, static_cast<cudaStream_t>(stream)>>>(
            params, static_cast<__nv_bfloat16*>(shadow_bf16), t, v, static_cast<uint32_t*>(pulled_ver), shard_size, n_tiles,
            static_cast<int>(shard_size / kChunk));
        DMLB_LAUNCH_CHECK();
        return 0;
    }
    pull_kernel<<<comm_grid(n / 4, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(params, static_cast<__nv_bfloat16*>(shadow_bf16), t, v,
                                                                                     static_cast<uint32_t*>(pulled_ver), shard_size,
                                                                                     n / 4, static_cast<int>(shard_size / kChunk));
    DMLB_LAUNCH_CHECK();
    return 0;
}

static int push_mailbox_impl(float* acc, const float* live, void* busy, const void* pulled_ver, void* const* mailbox_ptrs, void* const* full_ptrs,
                             int rank, int world, long long shard_size, void* stream);

extern "C" int dmlb_push_mailbox(float* acc, void* busy, const void* pulled_ver, void* const* mailbox_ptrs, void* const* full_ptrs, int rank,
                                 int world, long long shard_size, void* stream) {
    return push_mailbox_impl(acc, nullptr, busy, pulled_ver, mailbox_ptrs, full_ptrs, rank, world, shard_size, stream);
}

// Mailbox push in delta form (see dmlb_push_delta): chunks whose slot at the owner is busy are skipped and simply keep growing.
extern "C" int dmlb_push_mailbox_delta(const float* live, float* snapshot, void* busy, const void* pulled_ver, void* const* mailbox_ptrs,
                                       void* const* full_ptrs, int rank, int world, long long shard_size, void* stream) {
    DMLB_REQUIRE(live != nullptr && snapshot != nullptr, "live and snapshot are required");
    return push_mailbox_impl(snapshot, live, busy, pulled_ver, mailbox_ptrs, full_ptrs, rank, world, shard_size, stream);
}

static int push_mailbox_impl(float* acc, const float* live, void* busy, const void* pulled_ver, void* const* mailbox_ptrs, void* const* full_ptrs,
                             int rank, int world, long long shard_size, void* stream) {
    DMLB_REQUIRE(shard_size % kChunk == 0, "shard size must be a multiple of the chunk size");
    PeerTable m, f;
    if (fill_table(&m, mailbox_ptrs, world) || fill_table(&f, full_ptrs, world)) return -2;
    const int cps = static_cast<int>(shard_size / kChunk);
    int blocks = world * cps;
    const int cap = dmlb_sm_count() * 4;
    if (blocks > cap) blocks = cap;
    push_mailbox_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(acc, live, static_cast<uint32_t*>(busy),
                                                                              static_cast<const uint32_t*>(pulled_ver), m, f, rank, world,
                                                                              shard_size, cps);
    DMLB_LAUNCH_CHECK();
    return 0;
}

extern "C" int dmlb_owner_apply(float* master, float* mom, void* version, void* full, const float* mailbox, void* const* busy_ptrs, void* stats,
                                int rank, int world, long long shard_size, float mu, int stale_mode, float stale_param, float gain,
                                void* stream) {
    DMLB_REQUIRE(shard_size % kChunk == 0, "shard size must be a multiple of the chunk size");
    PeerTable b;
    if (fill_table(&b, busy_ptrs, world)) return -2;
    const int cps = static_cast<int>(shard_size / kChunk);
    int blocks = cps;
    const int cap = dmlb_sm_count() * 4;
    if (blocks > cap) blocks = cap;
    owner_apply_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(master, mom, static_cast<uint32_t*>(version),
                                                                             static_cast<uint32_t*>(full), mailbox, b,
                                                                             static_cast<uint32_t*>(stats), rank, world, shard_size, cps, mu,
                                                                             stale_mode, stale_param, gain);
    DMLB_LAUNCH_CHECK();
    return 0;
}

// Range form: this GPU owns [base, base + n) of the flat vector in this launch; mom (may be NULL) is indexed from the start of that
// slice.  max_blocks > 0 caps the grid (a small grid leaves SMs to the backward kernels it overlaps with).
extern "C" int dmlb_sync_rs_update_ag_range(void* const* grad_ptrs, void* const* param_ptrs, void* const* flag_ptrs, float* mom, int rank,
                                            int world, long long base, long long n, const float* lr_ptr, float mu, float wd, float grad_scale,
                                            int max_blocks, void* stream) {
    DMLB_REQUIRE(n % 4 == 0 && base % 4 == 0 && n > 0, "slice must be non-empty and 16-byte aligned");
    PeerTable g, p, f;
    if (fill_table(&g, grad_ptrs, world) || fill_table(&p, param_ptrs, world) || fill_table(&f, flag_ptrs, world)) return -2;
    // One CTA per SM at most: every CTA polls the start flags, so all of them must be able to become resident.
    int blocks = comm_grid(n / 4, 512);
    if (blocks > dmlb_sm_count()) blocks = dmlb_sm_count();
    if (max_blocks > 0 && blocks > max_blocks) blocks = max_blocks;
    sync_rs_update_ag_kernel<<<blocks, 512, 0, static_cast<cudaStream_t>(stream)>>>(g, p, f, mom, rank, world, base, n, lr_ptr, mu, wd,
                                                                                   grad_scale, barrier_spin_limit());
    DMLB_LAUNCH_CHECK();
    return 0;
}

extern "C" int dmlb_sync_rs_update_ag(void* const* grad_ptrs, void* const* param_ptrs, void* const* flag_ptrs, float* mom, int rank, int world,
                                      long long shard_size, const float* lr_ptr, float mu, float wd, float grad_scale, void* stream) {
    return dmlb_sync_rs_update_ag_range(grad_ptrs, param_ptrs, flag_ptrs, mom, rank, world, static_cast<long long>(rank) * shard_size, shard_size,
                                        lr_ptr, mu, wd, grad_scale, 0, stream);
}

// NVLS form of the range kernel: mc_grads / mc_params are multicast addresses (torch symmetric memory or any cuMulticast mapping).
extern "C" int dmlb_sync_nvls_range(const float* mc_grads, float* mc_params, const float* params_local, void* const* flag_ptrs, float* mom,
                                    int rank, int world, long long base, long long n, const float* lr_ptr, float mu, float wd, float grad_scale,
                                    int max_blocks, void* stream) {
    DMLB_REQUIRE(n % 4 == 0 && base % 4 == 0 && n > 0, "slice must be non-empty and 16-byte aligned");
    DMLB_REQUIRE(mc_grads != nullptr && mc_params != nullptr, "multicast addresses are required");
    PeerTable f;
    if (fill_table(&f, flag_ptrs, world)) return -2;
    int blocks = comm_grid(n / 4, 512);
    if (blocks > dmlb_sm_count()) blocks = dmlb_sm_count();
    if (max_blocks > 0 && blocks > max_blocks) blocks = max_blocks;
    sync_nvls_kernel<<<blocks, 512, 0, static_cast<cudaStream_t>(stream)>>>(mc_grads, mc_params, params_local, f, mom, rank, world, base, n, lr_ptr,
                                                                           mu, wd, grad_scale, barrier_spin_limit());
    DMLB_LAUNCH_CHECK();
    return 0;
}

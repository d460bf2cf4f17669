// Peer (symmetric) memory plumbing: raw device allocations that other processes on the same NVSwitch
// domain map with CUDA IPC, so the push / pull / sync kernels in comm.cu can load, store and reduce
// directly into other GPUs' memory over NVLink.
//
// Replaces the reference's transport — dist.init_process_group('gloo') + isend/recv of CPU tensors
// (reference example/main.py:165; asgd/optim/Asynchronous.py:34,49,59).  torch.distributed
// is still used, but only as the control plane that carries the 64-byte IPC handles.
#include "common.cuh"

extern "C" int dmlb_peer_alloc(size_t bytes, void** out) {
    DMLB_REQUIRE(out != nullptr && bytes > 0, "bad allocation request");
    void* p = nullptr;
    DMLB_CHECK_CUDA(cudaMalloc(&p, bytes));
    DMLB_CHECK_CUDA(cudaMemset(p, 0, bytes));
    DMLB_CHECK_CUDA(cudaDeviceSynchronize());
    *out = p;
    return 0;
}

extern "C" int dmlb_peer_free(void* p) {
    if (p) DMLB_CHECK_CUDA(cudaFree(p));
    return 0;
}

// handle_out: 64 bytes (CUDA_IPC_HANDLE_SIZE)
extern "C" int dmlb_ipc_get_handle(void* p, void* handle_out) {
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "unexpected IPC handle size");
    cudaIpcMemHandle_t h;
    DMLB_CHECK_CUDA(cudaIpcGetMemHandle(&h, p));
    memcpy(handle_out, &h, sizeof(h));
    return 0;
}

extern "C" int dmlb_ipc_open_handle(const void* handle, void** out) {
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    void* p = nullptr;
    DMLB_CHECK_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    *out = p;
    return 0;
}

extern "C" int dmlb_ipc_close(void* p) {
    if (p) DMLB_CHECK_CUDA(cudaIpcCloseMemHandle(p));
    return 0;
}

// Same-process multi-device use (tests, single-process launchers): enable direct access to `peer_dev`.
extern "C" int dmlb_enable_peer_access(int peer_dev) {
    int dev = 0;
    DMLB_CHECK_CUDA(cudaGetDevice(&dev));
    if (dev == peer_dev) return 0;
    int can = 0;
    DMLB_CHECK_CUDA(cudaDeviceCanAccessPeer(&can, dev, peer_dev));
    DMLB_REQUIRE(can, "devices cannot access each other's memory");
    cudaError_t e = cudaDeviceEnablePeerAccess(peer_dev, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) {
        cudaGetLastError();
        return 0;
    }
    DMLB_CHECK_CUDA(e);
    return 0;
}

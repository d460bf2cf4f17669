"""The allocation-free split-K forward entry point, called the way a C host would call it.

``dmlb_gemm_bias_act_splitk_noalloc`` exists for users of libdmlb who have no way to obtain device memory for a
workspace: programs that link the library without torch and own no allocator, services that gave the whole GPU to
their own pool at start-up (a late ``cudaMalloc`` may fail there), loops that cannot take the device-wide
synchronisation of ``cudaMalloc`` / ``cudaFree``.  Its workspace is a fixed array inside the library, shared by every
call on the device.

This file is named so that pytest collects it last: if it fails under ``-x``, every other test has already run.
The test goes through the raw C ABI with ``ctypes`` (not through the Python wrappers), checks the result against the
existing entry point that takes a caller-provided workspace — twice in a row, because the second call meets the
array as the first one left it — and checks that an oversized problem is refused with a clear message.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from distributed_ml_pytorch_b200 import _native
    lib = ctypes.CDLL(str(_native.LIB_PATH))
    vp, ci = ctypes.c_void_p, ctypes.c_int
    lib.dmlb_last_error.restype = ctypes.c_char_p
    lib.dmlb_noalloc_scratch_capacity.restype = ctypes.c_longlong
    lib.dmlb_gemm_bias_act_splitk.argtypes = [vp, ci, vp, ci, ci, ci, ci, vp, vp, ci, vp, ci, ci, vp]
    lib.dmlb_gemm_bias_act_splitk.restype = ci
    lib.dmlb_gemm_bias_act_splitk_noalloc.argtypes = [vp, ci, vp, ci, ci, ci, ci, vp, ci, vp, ci, ci, vp]
    lib.dmlb_gemm_bias_act_splitk_noalloc.restype = ci
    return lib


def _rel_err(a, b):
    return ((a.float() - b.float()).norm() / (b.float().norm() + 1e-12)).item()


@pytest.mark.parametrize("M,N,K,split", [(256, 384, 1728, 13), (64, 10, 2048, 8)])
def test_noalloc_matches_workspace_entry_point_twice(M, N, K, split):
    lib = _lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    stream = torch.cuda.current_stream().cuda_stream
    ldy = (N + 7) // 8 * 8
    ldw = (N + 3) // 4 * 4
    assert M * ldw <= lib.dmlb_noalloc_scratch_capacity()
    for call in range(2):
        torch.manual_seed(20 + call)          # fresh operands each round: stale scratch would show up as a wrong sum
        a = torch.randn(M, K, device=dev).to(torch.bfloat16).contiguous()
        w = (torch.randn(N, K, device=dev) * 0.05).to(torch.bfloat16).contiguous()
        bias = torch.randn(N, device=dev)
        ws = torch.zeros(M * ldw, device=dev, dtype=torch.float32)       # only the existing entry point needs this
        y_ref = torch.zeros(M, ldy, device=dev, dtype=torch.bfloat16)
        y_new = torch.zeros(M, ldy, device=dev, dtype=torch.bfloat16)
        rc = lib.dmlb_gemm_bias_act_splitk(a.data_ptr(), K, w.data_ptr(), K, M, N, K, ws.data_ptr(), bias.data_ptr(), 1, y_ref.data_ptr(), ldy,
                                           split, stream)
        assert rc == 0, lib.dmlb_last_error()
        rc = lib.dmlb_gemm_bias_act_splitk_noalloc(a.data_ptr(), K, w.data_ptr(), K, M, N, K, bias.data_ptr(), 1, y_new.data_ptr(), ldy, split,
                                                   stream)
        assert rc == 0, lib.dmlb_last_error()
        torch.cuda.synchronize()
        # same tolerance as the other GEMM tests (fp32 atomics may sum in a different order before the bf16 rounding)
        assert _rel_err(y_new[:, :N], y_ref[:, :N]) < 1e-2
        fp32 = torch.relu(a.float() @ w.float().t() + bias)
        assert _rel_err(y_new[:, :N], fp32) < 1e-2


def test_noalloc_leaves_free_device_memory_unchanged():
    """``cudaMemGetInfo`` before and after: the call must not take (or return) a single byte of device memory — not on the first
    call on the device, which looks the array's address up, and not on later ones."""
    lib = _lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)

    def free_bytes():
        return torch.cuda.mem_get_info(dev)[0]          # cudaMemGetInfo: free bytes of the whole device, all allocators included

    M, N, K, split = 256, 384, 1728, 13
    ldy = (N + 7) // 8 * 8
    a = torch.randn(M, K, device=dev).to(torch.bfloat16).contiguous()
    w = (torch.randn(N, K, device=dev) * 0.05).to(torch.bfloat16).contiguous()
    bias = torch.randn(N, device=dev)
    y = torch.zeros(M, ldy, device=dev, dtype=torch.bfloat16)
    stream = torch.cuda.current_stream().cuda_stream
    torch.cuda.synchronize()
    for _ in range(3):
        before = free_bytes()
        rc = lib.dmlb_gemm_bias_act_splitk_noalloc(a.data_ptr(), K, w.data_ptr(), K, M, N, K, bias.data_ptr(), 1, y.data_ptr(), ldy, split, stream)
        assert rc == 0, lib.dmlb_last_error()
        during = free_bytes()
        torch.cuda.synchronize()
        after = free_bytes()
        assert before == during == after, (before, during, after)


def test_noalloc_rejects_oversized_problem():
    lib = _lib()
    dev = torch.device("cuda", 0)
    cap = lib.dmlb_noalloc_scratch_capacity()
    M, N, K = 4096, 512, 64                       # 4096 * 512 = 2 Mi floats > 1 Mi
    assert M * N > cap
    a = torch.zeros(M, K, device=dev, dtype=torch.bfloat16)
    w = torch.zeros(N, K, device=dev, dtype=torch.bfloat16)
    y = torch.zeros(M, N, device=dev, dtype=torch.bfloat16)
    rc = lib.dmlb_gemm_bias_act_splitk_noalloc(a.data_ptr(), K, w.data_ptr(), K, M, N, K, None, 0, y.data_ptr(), N, 1,
                                               torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    assert b"exceeds the built-in scratch" in lib.dmlb_last_error()

"""NCCL point-to-point harness (work-list row K14; reference pytorch_p2p_ex.py:12,15): send/recv overlapped with
compute on a side stream, the 1 KB - 1 GB bandwidth sweep of BASELINE.json config 5, and the measurement of how much of a send hides
behind this package's own tensor-core GEMM — all device-timed."""
from __future__ import annotations

from typing import Dict, List

import torch
import torch.distributed as dist


def overlapped_send_recv(tensor: torch.Tensor, src: int, dst: int, side: torch.cuda.Stream):
    """Issue the transfer on ``side`` (after the producer stream) and return immediately; the caller joins with
    ``torch.cuda.current_stream().wait_stream(side)`` when it needs the data."""
    rank = dist.get_rank()
    side.wait_stream(torch.cuda.current_stream(tensor.device))
    with torch.cuda.stream(side):
        if rank == src:
            dist.send(tensor, dst=dst)
        elif rank == dst:
            dist.recv(tensor, src=src)
    return side


def p2p_bandwidth_sweep(rank: int, size: int, backend: str = "nccl", min_bytes: int = 1 << 10, max_bytes: int = 1 << 30,
                        iters: int = 20) -> List[Dict]:
    """rank 0 -> rank 1 send/recv; returns rows {bytes, ms, gbs} (max over ranks of the device time per message)."""
    cuda = backend == "nccl"
    device = torch.device("cuda", rank) if cuda else torch.device("cpu")
    rows = []
    nbytes = min_bytes
    while nbytes <= max_bytes:
        t = torch.empty(nbytes, dtype=torch.uint8, device=device)
        n = iters if nbytes <= (64 << 20) else 5
        for _ in range(3):
            dist.send(t, dst=1) if rank == 0 else dist.recv(t, src=0)
        if cuda:
            torch.cuda.synchronize(device)
            dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(n):
                dist.send(t, dst=1) if rank == 0 else dist.recv(t, src=0)
            e1.record()
            e1.synchronize()
            ms = e0.elapsed_time(e1) / n
            m = torch.tensor([ms], device=device, dtype=torch.float64)
            dist.all_reduce(m, op=dist.ReduceOp.MAX)
            ms = float(m.item())
        else:
            import time
            dist.barrier()
            t0 = time.perf_counter()
            for _ in range(n):
                dist.send(t, dst=1) if rank == 0 else dist.recv(t, src=0)
            ms = (time.perf_counter() - t0) * 1e3 / n
        rows.append({"bytes": nbytes, "ms": ms, "gbs": nbytes / (ms * 1e-3) / 1e9})
        nbytes *= 4
    return rows


def p2p_overlap_sweep(rank: int, size: int, min_bytes: int = 1 << 20, max_bytes: int = 1 << 30, iters: int = 8, gemm_dim: int = 4096) -> List[Dict]:
    """How much of an NCCL send / recv (rank 0 -> rank 1, on a side stream) hides behind compute on the main stream.

    The compute is this package's tcgen05 GEMM (``ops.kernels.gemm``, bf16 ``gemm_dim``^3), repeated so that it lasts about as long as the
    transfer.  Three device times per message size, each the max over the two ranks: the transfer alone, the GEMMs alone, both issued
    together.  ``hidden = (t_send + t_gemm - t_both) / min(t_send, t_gemm)``: 1.0 = the shorter of the two costs nothing, 0.0 = they ran
    one after the other."""
    from ..ops import kernels as K
    device = torch.device("cuda", rank)
    side = torch.cuda.Stream(device=device)
    a = torch.randn(gemm_dim, gemm_dim, device=device).to(torch.bfloat16)
    b = torch.randn(gemm_dim, gemm_dim, device=device).to(torch.bfloat16)
    out = torch.empty(gemm_dim, gemm_dim, device=device, dtype=torch.bfloat16)

    def gemms(k):
        for _ in range(k):
            K.gemm(a, gemm_dim, False, b, gemm_dim, False, gemm_dim, gemm_dim, gemm_dim, out, gemm_dim)

    def xfer(t):
        if rank == 0:
            dist.send(t, dst=1)
        elif rank == 1:
            dist.recv(t, src=0)

    def timed(fn):
        torch.cuda.synchronize(device)
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        e1.synchronize()
        m = torch.tensor([e0.elapsed_time(e1) / iters], device=device, dtype=torch.float64)
        dist.all_reduce(m, op=dist.ReduceOp.MAX)
        return float(m.item())

    gemms(3)
    t1 = timed(lambda: gemms(1))
    rows = []
    nbytes = min_bytes
    while nbytes <= max_bytes:
        t = torch.empty(nbytes, dtype=torch.uint8, device=device)
        for _ in range(2):
            xfer(t)
        t_send = timed(lambda: xfer(t))
        k = max(1, round(t_send / t1))
        t_gemm = timed(lambda: gemms(k))

        def both():
            side.wait_stream(torch.cuda.current_stream(device))
            with torch.cuda.stream(side):
                xfer(t)
            gemms(k)
            torch.cuda.current_stream(device).wait_stream(side)
        both()
        t_both = timed(both)
        rows.append({"bytes": nbytes, "send_ms": t_send, "gemm_ms": t_gemm, "gemms": k, "both_ms": t_both,
                     "hidden": (t_send + t_gemm - t_both) / min(t_send, t_gemm), "send_gbs": nbytes / (t_send * 1e-3) / 1e9,
                     "gemm_tflops": 2.0 * gemm_dim ** 3 * k / (t_gemm * 1e-3) / 1e12})
        del t
        nbytes *= 4
    return rows

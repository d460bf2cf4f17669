"""NCCL point-to-point harness (work-list row K14; reference pytorch_p2p_ex.py:12,15): send/recv overlapped with
compute on a side stream, and the 1 KB - 1 GB bandwidth sweep of BASELINE.json config 5, device-timed."""
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

"""Process launchers: "multi-node without a cluster" in the pattern of the reference's pytorch_p2p_ex.py:18-36
(``mp.set_start_method('spawn')`` + one ``init_process`` per rank on 127.0.0.1) and a torchrun-free multi-GPU
launcher for ``example/main.py`` / ``run-pytorch.py --local``.
"""
from __future__ import annotations

import os
import socket
from typing import Callable, Optional, Sequence

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def init_process(rank: int, size: int, fn: Callable, backend: str = "gloo", master_addr: str = "127.0.0.1", master_port: int = 29500,
                 args: Sequence = ()):
    """Initialise the distributed environment, then run ``fn(rank, size, *args)`` (reference pytorch_p2p_ex.py:18-23)."""
    os.environ["MASTER_ADDR"] = master_addr
    os.environ["MASTER_PORT"] = str(master_port)
    os.environ["RANK"] = str(rank)
    os.environ["WORLD_SIZE"] = str(size)
    os.environ.setdefault("LOCAL_RANK", str(rank))
    kwargs = {}
    if backend == "nccl":
        torch.cuda.set_device(rank % max(1, torch.cuda.device_count()))
        kwargs["device_id"] = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group(backend, rank=rank, world_size=size, **kwargs)
    try:
        fn(rank, size, *args)
    finally:
        if dist.is_initialized():
            dist.destroy_process_group()


def _entry(rank, size, fn, backend, port, args):
    init_process(rank, size, fn, backend=backend, master_port=port, args=args)


def spawn(fn: Callable, world_size: int, backend: str = "gloo", args: Sequence = (), port: Optional[int] = None, join: bool = True):
    """Run ``fn(rank, world_size, *args)`` in ``world_size`` fresh processes on this machine."""
    port = port or free_port()
    return mp.spawn(_entry, args=(world_size, fn, backend, port, tuple(args)), nprocs=world_size, join=join)

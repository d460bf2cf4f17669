#!/usr/bin/env python
"""Point-to-point send/recv demo — same behaviour as the reference's pytorch_p2p_ex.py (rank 0 sends tensor(1.) to
rank 1, both print it), spawned as two local processes on 127.0.0.1:29500.

    python pytorch_p2p_ex.py                    # gloo / CPU tensors, exactly the reference
    python pytorch_p2p_ex.py --backend nccl     # 2 GPUs: ncclSend/ncclRecv over NVLink, issued on a side stream and
                                                # overlapped with compute on the main stream
    python pytorch_p2p_ex.py --backend nccl --sweep    # bandwidth sweep 1 KB - 1 GB (BASELINE.json config 5)
    python pytorch_p2p_ex.py --backend nccl --overlap  # send || tcgen05 GEMM vs send + GEMM, 1 MB - 1 GB

Point-to-point stays on NCCL by design (BASELINE.json north star); everything bulk goes through the fused
peer-memory kernels in distributed_ml_pytorch_b200.parallel instead.
"""
import argparse
import os
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def run(rank, size, backend="gloo"):
    device = torch.device("cuda", rank) if backend == "nccl" else torch.device("cpu")
    tensor = torch.zeros(1, device=device)
    if backend == "nccl":
        # overlap: the transfer runs on a side stream while the main stream keeps computing
        side = torch.cuda.Stream(device=device)
        a = torch.randn(2048, 2048, device=device)
        side.wait_stream(torch.cuda.current_stream(device))
        with torch.cuda.stream(side):
            if rank == 0:
                tensor += 1
                dist.send(tensor=tensor, dst=1)
            else:
                dist.recv(tensor=tensor, src=0)
        busy = a @ a                                    # independent compute on the main stream
        torch.cuda.current_stream(device).wait_stream(side)
        torch.cuda.synchronize(device)
        del busy
    elif rank == 0:
        tensor += 1
        dist.send(tensor=tensor, dst=1)                 # send the tensor to process 1
    else:
        dist.recv(tensor=tensor, src=0)                 # receive tensor from process 0
    print("Rank ", rank, " has data ", tensor[0])


def sweep(rank, size, backend="nccl"):
    """send/recv bandwidth 1 KB .. 1 GB, device-timed with CUDA events, max over the two ranks."""
    from distributed_ml_pytorch_b200.parallel.p2p import p2p_bandwidth_sweep
    rows = p2p_bandwidth_sweep(rank, size, backend)
    if rank == 0:
        for r in rows:
            print("{bytes:>12d} B  {ms:10.4f} ms  {gbs:8.2f} GB/s".format(**r))


def overlap(rank, size, backend="nccl"):
    """How much of a 1 MB .. 1 GB send hides behind the package's own tensor-core GEMM on the main stream."""
    from distributed_ml_pytorch_b200.parallel.p2p import p2p_overlap_sweep
    rows = p2p_overlap_sweep(rank, size)
    if rank == 0:
        import json
        for r in rows:
            print("{bytes:>12d} B  send {send_ms:8.4f} ms ({send_gbs:6.1f} GB/s)  {gemms:3d} GEMM {gemm_ms:8.4f} ms ({gemm_tflops:6.1f} TFLOP/s)  "
                  "together {both_ms:8.4f} ms  hidden {hidden:5.2f}".format(**r))
        out = os.environ.get("P2P_OVERLAP_OUT")
        if out:
            with open(out, "w") as f:
                json.dump(rows, f, indent=1)


def init_process(rank, size, fn, backend="gloo"):
    """Initialize the distributed environment."""
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = os.environ.get("P2P_PORT", "29500")
    kw = {}
    if backend == "nccl":
        torch.cuda.set_device(rank)
        kw["device_id"] = torch.device("cuda", rank)
    dist.init_process_group(backend, rank=rank, world_size=size, **kw)
    fn(rank, size, backend)
    dist.destroy_process_group()


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--backend", default="gloo", choices=["gloo", "nccl"])
    ap.add_argument("--sweep", action="store_true")
    ap.add_argument("--overlap", action="store_true", help="nccl: measure how much of a send hides behind the tcgen05 GEMM (1 MB - 1 GB)")
    a = ap.parse_args()
    size = 2
    processes = []
    mp.set_start_method("spawn")
    for rank in range(size):
        p = mp.Process(target=init_process, args=(rank, size, overlap if a.overlap else sweep if a.sweep else run, a.backend))
        p.start()
        processes.append(p)
    for p in processes:
        p.join()
    sys.exit(max(p.exitcode or 0 for p in processes))

#!/usr/bin/env python
"""B2-nccl: the reference's parameter-server protocol on `nccl` isend / irecv with CUDA tensors.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P baseline/nccl_protocol.py --steps 100

The stock reference arm of bench.py is what the reference hard-codes: gloo over loopback TCP with CPU messages
(example/main.py:165).  This harness answers "what if the same protocol ran on the box's best library transport?":
rank 0 is the server (one GPU holds the flat parameter vector), ranks 1..N-1 are workers that train the REFERENCE's own
AlexNet (baseline/_ref/example/models.py, cuDNN/cuBLAS, fp32 tensors) and exchange the reference's messages — a flat fp32
CUDA tensor [sender, code, payload(N)] — with dist.isend / dist.irecv on the NCCL backend:

    GradientUpdate (every n_push steps)   worker -> server: payload = -lr * sum(grad); server: theta += payload
    ParameterRequest (every n_pull steps) worker -> server (the reference uploads N dummy floats with it, Asynchronous.py:48-49)
    ParameterUpdate                       server -> worker reply; the worker overwrites its model when the reply has arrived

NCCL has no receive-from-anyone and is not thread safe per communicator, so the reference's listener THREAD becomes a posted
irecv that is polled between steps (worker) and one posted irecv per worker that the serve loop polls (server).  Each
server <-> worker pair talks over its own two-rank communicator: point-to-point operations of ONE communicator execute in
issue order, so with a single communicator a reply to worker 2 would sit behind the posted receive from worker 1.  Everything
else — message sizes, cadence, per-tensor SGD launches, ravel / unravel copies — is the reference's.
Prints one JSON line: images/sec = batch * workers * steps / max-over-ranks device time (CUDA events).
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "baseline", "_ref"), os.path.join(ROOT, "baseline", "_ref", "example")]

import torch
import torch.distributed as dist
import torch.nn.functional as F

REQ, GRAD, UPD, BYE = 0.0, 1.0, 2.0, 3.0


def ravel(model, grads=False):
    return torch.cat([(p.grad if grads else p.data).reshape(-1) for p in model.parameters()])


def unravel(model, flat):
    o = 0
    for p in model.parameters():
        n = p.numel()
        p.data.copy_(flat[o:o + n].view_as(p))
        o += n


def serve(n, world, dev, pairs):
    theta = torch.zeros(n, device=dev)
    bufs = [torch.zeros(n + 2, device=dev) for _ in range(world)]
    works = [None] + [dist.irecv(bufs[w], src=w, group=pairs[w]) for w in range(1, world)]
    sends = []
    alive = world - 1
    have_init = False
    while alive:
        for w in range(1, world):
            wk = works[w]
            if wk is None or not wk.is_completed():
                continue
            wk.wait()
            code = float(bufs[w][1].item())
            if code == BYE:
                works[w] = None
                alive -= 1
                continue
            if code == UPD:
                if not have_init:                                 # the first worker's upload sets the initial parameters
                    theta.copy_(bufs[w][2:])
                    have_init = True
            elif code == GRAD:
                theta.add_(bufs[w][2:])
            elif code == REQ:
                m = torch.empty(n + 2, device=dev)
                m[0], m[1] = 0.0, UPD
                m[2:].copy_(theta)
                sends.append((dist.isend(m, dst=w, group=pairs[w]), m))
                sends = [(s, t) for (s, t) in sends if not s.is_completed()]
            works[w] = dist.irecv(bufs[w], src=w, group=pairs[w])
    for s, _ in sends:
        s.wait()


def work(args, rank, world, dev, pair):
    from models import AlexNet                                     # the reference's model, unmodified
    torch.manual_seed(1234)
    net = AlexNet().to(dev)
    net.train()
    n = sum(p.numel() for p in net.parameters())
    lr, n_push, n_pull = args.lr, args.n_push, args.n_pull
    acc = torch.zeros(n, device=dev)
    pending_sends = []

    def send(code, payload):
        m = torch.empty(n + 2, device=dev)
        m[0], m[1] = float(rank), code
        m[2:].copy_(payload)
        pending_sends.append((dist.isend(m, dst=0, group=pair), m))
        while len(pending_sends) > 8:
            pending_sends.pop(0)[0].wait()

    send(UPD, ravel(net))                                           # "this sets the initial model parameters" (Asynchronous.py:33-34)
    reply, reply_work = torch.zeros(n + 2, device=dev), None
    B = args.batch_size
    g = torch.Generator().manual_seed(100 + rank)
    pool = 64
    xs = torch.randn(pool, B, 3, 32, 32, generator=g)
    ys = torch.randint(0, 10, (pool, B), generator=g)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    idx = 0
    for it in range(args.warmup + args.steps):
        if it == args.warmup:
            torch.cuda.synchronize()
            e0.record()
        x, y = xs[it % pool].to(dev), ys[it % pool].to(dev)         # .cuda() of every batch (example/main.py:64-65)
        for p in net.parameters():
            p.grad = None
        loss = F.cross_entropy(net(x), y)
        loss.backward()
        # ---- the reference's optimizer step (asgd/optim/Asynchronous.py:42-71) on CUDA tensors ----
        if reply_work is not None and reply_work.is_completed():   # listener: apply a reply that has arrived
            reply_work.wait()
            unravel(net, reply[2:])
            reply_work = None
        if idx % n_pull == 0:
            send(REQ, acc)                                          # dummy payload, like the reference
            if reply_work is None:
                reply_work = dist.irecv(reply, src=0, group=pair)
        acc.add_(ravel(net, grads=True), alpha=-lr)
        if idx % n_push == 0:
            send(GRAD, acc)
            acc.zero_()
        for p in net.parameters():
            p.data.add_(p.grad, alpha=-lr)
        idx += 1
        loss.item()                                                 # example/main.py:79
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    if reply_work is not None:
        reply_work.wait()
    send(BYE, acc)
    for s, _ in pending_sends:
        s.wait()
    return ms


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--batch-size", type=int, default=64)
    ap.add_argument("--lr", type=float, default=0.008)
    ap.add_argument("--n-push", type=int, default=10)
    ap.add_argument("--n-pull", type=int, default=10)
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)))
    torch.cuda.set_device(dev)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    from models import AlexNet
    n = sum(p.numel() for p in AlexNet().parameters())
    pairs = [None] + [dist.new_group(ranks=[0, w], backend="nccl") for w in range(1, world)]   # collective: every rank creates all
    t0 = time.time()
    ms = 0.0
    if rank == 0:
        serve(n, world, dev, pairs)
    else:
        ms = work(args, rank, world, dev, pairs[rank])
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        ms = float(t.item())
        workers = world - 1
        print(json.dumps({"impl": "reference protocol on nccl isend/irecv (B2-nccl)", "n_gpus": world, "workers": workers, "steps": args.steps,
                          "ms_per_step": round(ms / args.steps, 4), "value": round(args.batch_size * workers * args.steps / (ms / 1e3), 1),
                          "unit": "images/sec", "dtype": "fp32 (cuDNN TF32 convolutions)", "message_floats": n + 2, "n_push": args.n_push,
                          "n_pull": args.n_pull, "wall_s": round(time.time() - t0, 1)}), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

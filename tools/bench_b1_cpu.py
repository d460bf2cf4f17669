#!/usr/bin/env python
"""LeNet, asynchronous SGD, 1 server + 2 workers on CPU over gloo (BASELINE.json configs[0]).

    python tools/bench_b1_cpu.py [--steps 150] [--impl ours|reference|both]

ours      : distributed_ml_pytorch_b200 (asgd.optim.Asynchronous + asgd.server.ParameterServer, gloo transport), LeNet on
            MNIST-shaped 1x28x28 synthetic data (the north-star shape) and on 3x32x32.
reference : baseline/_ref — the reference's own Asynchronous optimizer, LeNet class and (completed) server, driven by a
            loop that mirrors example/main.py:57-73 (its main() hard-codes AlexNet, main.py:41, so LeNet cannot be selected
            through its CLI); 3x32x32 only (its LeNet does not accept 1x28x28).
Images/sec = workers * batch * steps / wall time of the slowest worker (CPU run: wall clock).
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch
import torch.distributed as dist
import torch.nn.functional as F

BATCH, LR, NP = 64, 0.008, 10


def _ours(rank, size, steps, channels, hw, out):
    from asgd.optim import Asynchronous
    from asgd.server import ParameterServer
    from distributed_ml_pytorch_b200.models import build_model
    torch.manual_seed(0)
    torch.set_num_threads(2)
    net = build_model("lenet", in_channels=channels, input_hw=hw)
    if rank == 0:
        ParameterServer(model=net).run()
        return
    opt = Asynchronous(net.parameters(), lr=LR, n_push=NP, n_pull=NP, model=net)
    g = torch.Generator().manual_seed(rank)
    xs = torch.randn(16, BATCH, channels, hw, hw, generator=g)
    ys = torch.randint(0, 10, (16, BATCH), generator=g)
    for i in range(5):
        opt.zero_grad(); F.cross_entropy(net(xs[i % 16]), ys[i % 16]).backward(); opt.step()
    t0 = time.perf_counter()
    for i in range(steps):
        opt.zero_grad()
        loss = F.cross_entropy(net(xs[i % 16]), ys[i % 16])
        loss.backward()
        opt.step()
        loss.item()
    dt = time.perf_counter() - t0
    opt.finish()
    with open(os.path.join(out, f"ours{rank}.json"), "w") as f:
        json.dump({"s": dt, "loss": float(loss.detach())}, f)


def _ref(rank, size, steps, out):
    sys.path[:0] = [os.path.join(ROOT, "baseline", "_ref"), os.path.join(ROOT, "baseline", "_ref", "example")]
    import warnings
    warnings.filterwarnings("ignore")
    from asgd.optim import Asynchronous            # the reference's class (baseline/_ref/asgd/optim/Asynchronous.py)
    from asgd.utils.messaging import MessageCode, send_message
    from models import LeNet                       # the reference's example/models.py
    torch.manual_seed(0)
    torch.set_num_threads(2)
    net = LeNet()
    n_params = sum(p.numel() for p in net.parameters())
    if rank == 0:
        import asgd.server as srv
        server = srv.ParameterServer(model=net)
        done = 0
        while done < size - 1:
            dist.recv(tensor=server.m_parameter)
            code = MessageCode(int(server.m_parameter[1].item()))
            if code == MessageCode.Shutdown:
                done += 1
                continue
            server.receive(int(server.m_parameter[0].item()), code, server.m_parameter[2:])
        for w in range(1, size):
            send_message(MessageCode.Shutdown, torch.zeros(n_params), dst=w)
        from asgd.utils import messaging as m
        for wk, _ in m._keep:
            wk.wait()
        return
    opt = Asynchronous(net.parameters(), lr=LR, n_push=NP, n_pull=NP, model=net)
    g = torch.Generator().manual_seed(rank)
    xs = torch.randn(16, BATCH, 3, 32, 32, generator=g)
    ys = torch.randint(0, 10, (16, BATCH), generator=g)
    for i in range(5):
        opt.zero_grad(); F.cross_entropy(net(xs[i % 16]), ys[i % 16]).backward(); opt.step()
    t0 = time.perf_counter()
    for i in range(steps):                          # example/main.py:68-79
        opt.zero_grad()
        outputs = net(xs[i % 16])
        loss = F.cross_entropy(outputs, ys[i % 16])
        loss.backward()
        opt.step()
        loss.item()
    dt = time.perf_counter() - t0
    send_message(MessageCode.Shutdown, torch.zeros(n_params))
    from asgd.utils import messaging as m
    for wk, _ in m._keep:
        wk.wait()
    # The reference keeps its listener in a local variable (Asynchronous.py:37-38).  It must have consumed the server's Shutdown
    # (sent once EVERY worker is done) before this process tears the group down, or the thread dies inside dist.recv and aborts.
    import threading
    for th in threading.enumerate():
        if th is not threading.current_thread() and type(th).__name__ == "Listener":
            th.join(timeout=120)
    with open(os.path.join(out, f"ref{rank}.json"), "w") as f:
        json.dump({"s": dt, "loss": float(loss.detach())}, f)


def main():
    import tempfile
    from distributed_ml_pytorch_b200.parallel.launcher import spawn
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=150)
    ap.add_argument("--impl", default="both")
    a = ap.parse_args()
    res = {}
    with tempfile.TemporaryDirectory() as out:
        if a.impl in ("ours", "both"):
            for name, (c, hw) in {"ours_lenet_mnist_1x28x28": (1, 28), "ours_lenet_3x32x32": (3, 32)}.items():
                spawn(_ours, 3, args=(a.steps, c, hw, out))
                ts = [json.load(open(os.path.join(out, f"ours{r}.json"))) for r in (1, 2)]
                res[name] = {"images_per_sec": round(2 * BATCH * a.steps / max(t["s"] for t in ts), 1), "final_loss": ts[0]["loss"]}
        if a.impl in ("reference", "both"):
            spawn(_ref, 3, args=(a.steps, out))
            ts = [json.load(open(os.path.join(out, f"ref{r}.json"))) for r in (1, 2)]
            res["reference_lenet_3x32x32"] = {"images_per_sec": round(2 * BATCH * a.steps / max(t["s"] for t in ts), 1), "final_loss": ts[0]["loss"]}
    print(json.dumps({"config": "LeNet async, 1 server + 2 workers, CPU/gloo, batch 64, n_push=n_pull=10", "steps": a.steps, **res}))


if __name__ == "__main__":
    main()

"""Is the flagship loop GPU-bound?  (a) back-to-back graph replays, (b) the full TrainStep call loop, (c) host time per call."""
import os, sys, time, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from distributed_ml_pytorch_b200.engine import TrainStep
from distributed_ml_pytorch_b200.models import build_model
from distributed_ml_pytorch_b200.parallel.async_optim import Asynchronous
from distributed_ml_pytorch_b200.parallel.fused_sgd import FusedSGD

dev = torch.device("cuda", 0)
for name, mk in (("fused-sgd", lambda n: FusedSGD(n.parameters(), lr=0.008, model=n)),
                 ("async", lambda n: Asynchronous(n.parameters(), lr=0.008, n_push=10, n_pull=10, model=n))):
    torch.manual_seed(0)
    net = build_model("alexnet").to(dev)
    opt = mk(net)
    ts = TrainStep(net, opt, 64, (3, 32, 32), device=dev)
    x = torch.randn(64, 3, 32, 32, device=dev); y = torch.randint(0, 10, (64,), device=dev)
    for _ in range(20): ts(x, y)
    ts.flush(); torch.cuda.synchronize()
    N = 300
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(ts.stream):
        e0.record(ts.stream)
        for _ in range(N): ts.graph.replay()
        e1.record(ts.stream)
    e1.synchronize()
    a = e0.elapsed_time(e1) / N
    if ts._ring is not None:                 # direct replays advanced the loss kernel's launch counter behind TrainStep's back
        ts._ring_base = int(ts._ring[1].item()) - ts.step_count
    e0.record(ts.stream)
    t0 = time.perf_counter()
    for _ in range(N): ts(x, y)
    t1 = time.perf_counter()
    e1.record(ts.stream); ts.flush(); e1.synchronize()
    b = e0.elapsed_time(e1) / N
    print(f"{name}: graph replay only {a*1e3:.1f} us/step | TrainStep loop {b*1e3:.1f} us/step (device) | host loop wall {(t1-t0)/N*1e6:.1f} us/call | kernels {ts.kernels_per_step}")

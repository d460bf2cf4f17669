"""Per-call device time of AlexNet conv1 forward / forward+backward inside a CUDA graph (10 calls per replay)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from distributed_ml_pytorch_b200 import nn as dnn
from distributed_ml_pytorch_b200.ops import functional as OF
from distributed_ml_pytorch_b200.parallel.fused_sgd import FusedSGD
from distributed_ml_pytorch_b200.models import build_model

def timeit(fn, reps=10, iters=30):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(3): fn()
        s.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for _ in range(reps): fn()
        for _ in range(3): g.replay()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(iters): g.replay()
        e1.record(s)
        e1.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (reps * iters)

for dtype in ("tf32", "bf16"):
    for smallc in (True, False):
        OF._conv_opts.smallc = smallc
        torch.manual_seed(0)
        net = build_model("alexnet", dtype=dtype).cuda()
        opt = FusedSGD(net.parameters(), lr=0.0, model=net)
        conv1 = net.features[0]
        x = torch.randn(64, 3, 32, 32, device="cuda")
        if smallc:
            xin = OF.smallc_input(64, 3, 32, 32, conv1.compute_dtype, x.device, conv1._rowwin or __import__("distributed_ml_pytorch_b200.ops.kernels", fromlist=["x"]).rowwin_geometry(32, 32, 11, 11, 4, 4, 5, 5, conv1.compute_dtype), (5, 5))
            OF.fill_smallc_input(xin, x)
        else:
            xin = x
        y = conv1(xin)
        g = OF.new_act4(*y.shape, y.device, y.dtype); g.copy_(torch.randn_like(y.float()))
        OF.set_wgrad_side_stream(False)
        try:
            tf = timeit(lambda: conv1(xin))
            print(f"{dtype} smallc={smallc}: conv1 fwd {tf:.2f} us", flush=True)
        except Exception as e:
            print("fwd capture failed", dtype, smallc, str(e)[:200], flush=True)
            continue
        def fb():
            out = conv1(xin)
            out.backward(g)
        try:
            tfb = timeit(fb)
            print(f"{dtype} smallc={smallc}: fwd+bwd {tfb:.2f} us  (bwd = act_bwd + wgrad: {tfb - tf:.2f} us)", flush=True)
        except Exception as e:
            print("fwd+bwd capture failed", dtype, smallc, str(e)[:200], flush=True)

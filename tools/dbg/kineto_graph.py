import sys, os, json, collections, tempfile
sys.path.insert(0, ".")
import torch
from torch.profiler import profile, ProfilerActivity
from distributed_ml_pytorch_b200.engine import TrainStep
from distributed_ml_pytorch_b200.models import build_model
from distributed_ml_pytorch_b200.parallel.fused_sgd import FusedSGD
model = sys.argv[1] if len(sys.argv) > 1 else "resnet50"
hw = int(sys.argv[2]) if len(sys.argv) > 2 else 224
B = int(sys.argv[3]) if len(sys.argv) > 3 else 64
nc = 10 if hw <= 64 else 1000
dev = torch.device("cuda", 0)
torch.manual_seed(0)
net = build_model(model, input_hw=hw, num_classes=nc).to(dev).train()
opt = FusedSGD(net.parameters(), lr=0.008, model=net)
ts = TrainStep(net, opt, B, (3, hw, hw), device=dev)
x = torch.randn(4, B, 3, hw, hw, device=dev); y = torch.randint(0, nc, (4, B), device=dev)
for i in range(6): ts(x[i % 4], y[i % 4])
ts.flush(); torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for i in range(3): ts(x[i % 4], y[i % 4])
    ts.flush(); torch.cuda.synchronize()
path = os.path.join(tempfile.mkdtemp(), "t.json")
prof.export_chrome_trace(path)
ev = [e for e in json.load(open(path))["traceEvents"] if e.get("cat") == "kernel"]
ev.sort(key=lambda e: e["ts"])
t0, t1 = ev[0]["ts"], max(e["ts"] + e["dur"] for e in ev)
print("kernels", len(ev), "span ms per step", (t1 - t0) / 3e3)
per_stream = collections.defaultdict(float)
agg = collections.defaultdict(lambda: [0, 0.0])
for e in ev:
    per_stream[e["args"].get("stream")] += e["dur"]
    n = e["name"][:100]; agg[n][0] += 1; agg[n][1] += e["dur"]
print("busy ms per step by stream", {k: round(v / 3e3, 3) for k, v in per_stream.items()})
# idle gaps on the union of all streams
iv = sorted((e["ts"], e["ts"] + e["dur"]) for e in ev)
busy = 0.0; cs, ce = iv[0]
for s, e in iv[1:]:
    if s > ce: busy += ce - cs; cs, ce = s, e
    else: ce = max(ce, e)
busy += ce - cs
print("union busy ms per step", busy / 3e3)
tot = sum(v[1] for v in agg.values())
for n, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:24]:
    print(f"{v[1] / 3:9.0f} us {100 * v[1] / tot:5.1f}% n={v[0] // 3:4d}  {n}")

import sys, os, re, collections
sys.path.insert(0, ".")
import torch
from torch.profiler import profile, ProfilerActivity
from distributed_ml_pytorch_b200.models import build_model
from distributed_ml_pytorch_b200.ops import functional as OF
from distributed_ml_pytorch_b200.parallel.fused_sgd import FusedSGD
model = sys.argv[1] if len(sys.argv) > 1 else "resnet50"
hw = int(sys.argv[2]) if len(sys.argv) > 2 else 224
B = int(sys.argv[3]) if len(sys.argv) > 3 else 64
dev = torch.device("cuda", 0)
torch.manual_seed(0)
net = build_model(model, input_hw=hw, num_classes=10 if hw <= 64 else 1000).to(dev)
opt = FusedSGD(net.parameters(), lr=0.008, momentum=0.9, model=net)
x = torch.randn(B, 3, hw, hw, device=dev); y = torch.randint(0, 10, (B,), device=dev)
def step():
    opt.zero_grad(); loss = OF.cross_entropy(net(x), y); loss.backward(); opt.step()
for _ in range(3): step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5): step()
e1.record(); torch.cuda.synchronize()
print("eager ms/step", e0.elapsed_time(e1) / 5)
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    step(); torch.cuda.synchronize()
agg = collections.defaultdict(lambda: [0, 0.0])
tot = 0.0
for ev in prof.events():
    if ev.device_type == torch.autograd.DeviceType.CUDA:
        n = ev.name[:110]
        agg[n][0] += 1; agg[n][1] += ev.device_time_total if hasattr(ev, "device_time_total") else ev.cuda_time_total
        tot += agg[n][1] * 0
tot = sum(v[1] for v in agg.values())
print("sum of kernel time us", round(tot))
for n, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:40]:
    print(f"{v[1]:9.0f} us {100 * v[1] / tot:5.1f}% n={v[0]:4d}  {n}")

"""Repeat the AlexNet trajectory test body and print the loss curves (used to set robust thresholds for the GPU test)."""
import sys
import torch
import torch.nn.functional as F
from distributed_ml_pytorch_b200.engine import TrainStep
from distributed_ml_pytorch_b200.models import build_model
from distributed_ml_pytorch_b200.parallel import Asynchronous

lr = float(sys.argv[1]) if len(sys.argv) > 1 else 0.1
torch.manual_seed(5)
ref0 = build_model("alexnet").train()
sd = {k: v.clone() for k, v in ref0.state_dict().items()}
x = torch.randn(64, 3, 32, 32)
y = torch.randint(0, 10, (64,))
sgd = torch.optim.SGD(ref0.parameters(), lr=lr)
cpu = []
for _ in range(200):
    sgd.zero_grad()
    loss = F.cross_entropy(ref0(x), y)
    loss.backward()
    sgd.step()
    cpu.append(loss.item())
pick = [0, 5, 10, 20, 50, 100, 150, 199]
print("cpu ", [round(cpu[i], 4) for i in pick], "min", round(min(cpu), 4))
for rep in range(6):
    net = build_model("alexnet")
    net.load_state_dict(sd)
    net = net.cuda().train()
    opt = Asynchronous(net.parameters(), lr=lr, n_push=1, n_pull=1, model=net)
    ts = TrainStep(net, opt, 64, (3, 32, 32))
    xg, yg = x.cuda(), y.cuda()
    gpu = [ts(xg, yg, sync=True).loss for _ in range(200)]
    print("gpu%d" % rep, [round(gpu[i], 4) for i in pick], "min", round(min(gpu), 4), "maxdiff20", round(max(abs(gpu[i] - cpu[i]) for i in range(20)), 5))
    opt.finish()

#!/usr/bin/env python
"""A few eager (non-graph) training steps of one model — the command ncu wraps for per-kernel timing.

    python tools/profile_step.py --model alexnet --steps 4 --batch 64 [--mode local]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from distributed_ml_pytorch_b200.models import build_model
from distributed_ml_pytorch_b200.ops import functional as OF
from distributed_ml_pytorch_b200.parallel.fused_sgd import FusedSGD

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="alexnet")
ap.add_argument("--steps", type=int, default=4)
ap.add_argument("--batch", type=int, default=64)
ap.add_argument("--hw", type=int, default=32)
a = ap.parse_args()
dev = torch.device("cuda", 0)
torch.manual_seed(0)
net = build_model(a.model, input_hw=a.hw, num_classes=10 if a.hw <= 64 else 1000).to(dev)
opt = FusedSGD(net.parameters(), lr=0.008, model=net)
x = torch.randn(a.batch, 3, a.hw, a.hw, device=dev)
y = torch.randint(0, 10, (a.batch,), device=dev)
OF.set_unit_loss_grad(True)
for i in range(a.steps):
    torch.cuda.nvtx.range_push(f"step{i}")
    opt.zero_grad()
    loss = OF.cross_entropy(net(x), y)
    loss.backward()
    opt.step()
    torch.cuda.nvtx.range_pop()
torch.cuda.synchronize()
print("ok", float(loss.detach()))

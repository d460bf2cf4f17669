#!/usr/bin/env python
"""Stock-PyTorch baselines for the NORTH-STAR configurations the reference itself cannot run (it has no ResNets and
no synchronous mode): torchvision model + cuDNN/cuBLAS + NCCL ``all_reduce`` (or DDP) + ``torch.optim.SGD``.

    python baseline/torch_baseline.py --model resnet18 --input-hw 32 --batch-size 64 [--amp] [--channels-last]
    python -m torch.distributed.run --nproc-per-node N ... baseline/torch_baseline.py --model resnet50 --input-hw 224 --amp

Same timing rules as bench.py (device-timed, max over ranks, pinned-host inputs copied every step, loss read back every
step) and the same JSON shape, with "impl": "torch-baseline".  This is BASELINE.md rows B3 / B4; none of this
framework's code is on the path.
"""
import argparse
import json
import os

import torch
import torch.distributed as dist
import torch.nn.functional as F
import torchvision

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="resnet18")
ap.add_argument("--input-hw", type=int, default=32)
ap.add_argument("--classes", type=int, default=10)
ap.add_argument("--batch-size", type=int, default=64)
ap.add_argument("--steps", type=int, default=100)
ap.add_argument("--warmup", type=int, default=10)
ap.add_argument("--amp", action="store_true", help="bf16 autocast")
ap.add_argument("--channels-last", action="store_true")
ap.add_argument("--momentum", type=float, default=0.0)
a = ap.parse_args()

rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
torch.manual_seed(1234)
if a.model == "alexnet":
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "example"))
    from models import AlexNet          # the reference's own model file
    net = AlexNet()
else:
    net = getattr(torchvision.models, a.model)(num_classes=a.classes)
    if a.input_hw <= 64:                # the usual CIFAR stem, same as this framework's small_input variant
        net.conv1 = torch.nn.Conv2d(3, 64, 3, 1, 1, bias=False)
        net.maxpool = torch.nn.Identity()
net = net.to(dev).train()
if a.channels_last:
    net = net.to(memory_format=torch.channels_last)
if world > 1:
    net = torch.nn.parallel.DistributedDataParallel(net, device_ids=[local])
opt = torch.optim.SGD(net.parameters(), lr=0.008, momentum=a.momentum)
B, HW = a.batch_size, a.input_hw
pool = max(8, min(256, (160 << 20) // (B * 3 * HW * HW * 4) + 1))
hx = torch.randn(pool, B, 3, HW, HW).pin_memory()
hy = torch.randint(0, a.classes, (pool, B)).pin_memory()


def step(i):
    x = hx[i % pool].to(dev, non_blocking=True)
    y = hy[i % pool].to(dev, non_blocking=True)
    if a.channels_last:
        x = x.contiguous(memory_format=torch.channels_last)
    opt.zero_grad(set_to_none=True)
    with torch.autocast("cuda", dtype=torch.bfloat16, enabled=a.amp):
        loss = F.cross_entropy(net(x), y)
    loss.backward()
    opt.step()
    return loss.item()          # the reference's loop reads the loss every step (main.py:79)


for i in range(max(a.warmup, 3)):
    step(i)
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(a.steps):
    last = step(i)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
if world > 1:
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
if rank == 0:
    val = B * world * a.steps / (ms / 1e3)
    print(json.dumps({"metric": "images/sec (device-timed, max over ranks)", "value": round(val, 1), "unit": "images/sec", "n_gpus": world,
                      "steps": a.steps, "warmup": max(a.warmup, 3), "ms_per_step": round(ms / a.steps, 5), "impl": "torch-baseline",
                      "dtype": "bf16-autocast" if a.amp else "fp32",
                      "config": {"model": a.model, "input": f"3x{HW}x{HW}", "per_gpu_batch": B, "global_batch": B * world,
                                 "parallelism": "single GPU" if world == 1 else f"DDP (NCCL all-reduce) dp{world}", "channels_last": a.channels_last},
                      "final_loss": last}))
if world > 1:
    dist.destroy_process_group()

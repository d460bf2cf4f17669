#!/usr/bin/env python
"""Every communication kernel once, on ONE GPU (world 1: the GPU is worker and owner of the whole vector) — the command ncu wraps.
With one GPU nothing crosses NVLink; the capture shows the kernels' own shape (registers, shared memory, DRAM throughput, issue mix)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from distributed_ml_pytorch_b200.parallel import peer

mb = int(sys.argv[1]) if len(sys.argv) > 1 else 64
dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
pg = peer.PeerGroup(device=dev)
chunk = peer.chunk_elems()
n = (mb << 20) // 4 // chunk * chunk
cps = n // chunk
master = pg.symmetric(n * 4)
version = pg.symmetric(cps * 4)
mailbox = pg.symmetric(n * 4)
full = pg.symmetric(cps * 4)
busy = pg.symmetric(cps * 4)
params_s = pg.symmetric(n * 4)
grads_s = pg.symmetric(n * 4)
flags = pg.symmetric(peer.flag_words() * 4)
for b in (master, mailbox, params_s, grads_s):
    b.view(torch.float32, n).fill_(1.0)
for b in (version, full, busy):
    b.view(torch.int32, cps).zero_()
flags.view(torch.int32, peer.flag_words()).zero_()
acc = torch.ones(n, device=dev)
live = torch.randn(n, device=dev)
snap = torch.zeros(n, device=dev)
params = torch.empty(n, device=dev)
shadow = torch.empty(n, device=dev, dtype=torch.bfloat16)
pv = torch.zeros(cps, dtype=torch.int32, device=dev)
stats = torch.zeros(4, dtype=torch.int32, device=dev)
lr = torch.tensor([0.01], device=dev)
for rep in range(2):
    for mover in (1, 2):
        peer.set_comm_mover(mover)
        peer.push_add(acc, master.table, 1, n, n)
        peer.push_delta(live, snap, master.table, 1, n, n)
        peer.pull(params, master.table, 1, n, n, version.table, pv, shadow=shadow)
    peer.set_comm_mover(0)
    acc.fill_(1.0)
    peer.push_mailbox(acc, busy.view(torch.int32, cps), pv, mailbox.table, full.table, 0, 1, n)
    peer.owner_apply(master.view(torch.float32, n), None, version.view(torch.int32, cps), full.view(torch.int32, cps), mailbox.view(torch.float32, n),
                     busy.table, stats, 0, 1, n, 0.0, 1, 0.9)
    peer.sync_rs_update_ag(grads_s.table, params_s.table, flags.table, None, 0, 1, n, lr, 0.0, 0.0, 1.0)
torch.cuda.synchronize()
print("ok", mb, "MB", float(master.view(torch.float32, n)[:4].sum()))

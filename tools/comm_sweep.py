#!/usr/bin/env python
"""Bandwidth sweep of the fused NVLink kernels against NCCL (BASELINE.json config 5: 1 KB - 1 GB at 2/4/8 GPUs).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/comm_sweep.py [--max-mb 1024]

Per message size (fp32 elements, whole vector = one "model"):
  push   dmlb_push_add        every GPU adds its N-vector into the sharded masters (red.add.v4 over NVLink)
  pull   dmlb_pull            every GPU reads the N-vector back from the shard owners (peer loads)
  sync   dmlb_sync_rs_update_ag   barrier + reduce-scatter + SGD + all-gather + barrier, one kernel (peer loads / stores)
  nvls   dmlb_sync_nvls_range     the same step with multimem.ld_reduce / multimem.st (reduction and multicast inside the NVSwitch)
  nccl   all_reduce + flat axpy   the baseline formulation of the synchronous step
  p2p    NCCL send/recv rank0 -> rank1 (the reference's pytorch_p2p_ex.py pattern)
All times are CUDA-event device times, max over ranks.  "wire" = bytes each GPU must move over NVLink in one
direction: N*4*(W-1)/W for push, pull and each half of sync; roofline = wire / 770 GB/s (measured peer copy).
Writes one JSON line per size (rank 0) and, with --out, a JSON file.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.distributed as dist

from distributed_ml_pytorch_b200.parallel import peer

LINK_GBS = 770.0


def timed(fn, iters, pre=None):
    """Device time per call: `iters` back-to-back launches between one pair of events (so the host's launch latency is
    hidden behind queued work); `pre` (e.g. refilling the accumulator) is timed separately and subtracted."""
    def loop(body):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(iters):
            body()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / iters
    if pre is None:
        return loop(fn)
    pre()            # first use of `pre` (lazy module load of its kernel) must not land inside the timed loop

    def both():
        pre()
        fn()
    return max(loop(both) - loop(pre), 1e-4)


def max_ranks(ms, dev):
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-mb", type=int, default=1024)
    ap.add_argument("--out", default=None)
    ap.add_argument("--min-kb", type=int, default=1)
    ap.add_argument("--movers-only", action="store_true", help="only the push / pull movers (per-thread and TMA bulk forms), no sync / NCCL columns")
    a = ap.parse_args()
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    pg = peer.PeerGroup(device=dev)
    chunk = peer.chunk_elems()
    lr = torch.tensor([0.01], device=dev)
    rows = []
    nbytes = a.min_kb << 10
    while nbytes <= a.max_mb << 20:
        n = max(nbytes // 4, 4)
        shard = ((n + world - 1) // world + chunk - 1) // chunk * chunk
        padded = shard * world
        master = pg.symmetric(shard * 4)
        params_s = pg.symmetric(padded * 4)
        grads_s = pg.symmetric(padded * 4)
        flags = pg.symmetric(peer.flag_words() * 4)
        acc = torch.ones(padded, device=dev)
        params = params_s.view(torch.float32, padded)
        grads = grads_s.view(torch.float32, padded)
        grads.fill_(1.0)
        nccl_buf = torch.ones(n, device=dev)
        nccl_p = torch.zeros(n, device=dev)
        iters = 20 if nbytes <= (64 << 20) else 6
        pg.barrier()
        # warm-up
        for _ in range(3):
            peer.push_add(acc, master.table, world, shard, padded)
            peer.pull(params, master.table, world, shard, padded)
            peer.sync_rs_update_ag(grads_s.table, params_s.table, flags.table, None, rank, world, shard, lr, 0.0, 0.0, 1.0 / world)
            dist.all_reduce(nccl_buf)
        pg.barrier()
        movers = {}
        snap = torch.zeros(padded, device=dev)
        wire = n * 4 * (world - 1) / world if world > 1 else n * 4
        for name, mv in (("thread", 1), ("bulk", 2)):
            peer.set_comm_mover(mv)
            for _ in range(2):
                peer.push_add(acc, master.table, world, shard, padded)
                peer.push_delta(params, snap, master.table, world, shard, padded)
                peer.pull(params, master.table, world, shard, padded)
            pg.barrier()
            tp = max_ranks(timed(lambda: peer.push_add(acc, master.table, world, shard, padded), iters, pre=lambda: acc.fill_(1.0)), dev)
            pg.barrier()
            td = max_ranks(timed(lambda: peer.push_delta(params, snap, master.table, world, shard, padded), iters,
                                 pre=lambda: params.add_(1.0)), dev)      # a delta of zero is not sent by the per-thread form
            pg.barrier()
            tl = max_ranks(timed(lambda: peer.pull(params, master.table, world, shard, padded), iters), dev)
            pg.barrier()
            movers[name] = {"push_us": tp * 1e3, "push_delta_us": td * 1e3, "pull_us": tl * 1e3, "push_gbs": wire / (tp * 1e-3) / 1e9,
                            "push_delta_gbs": wire / (td * 1e-3) / 1e9, "pull_gbs": wire / (tl * 1e-3) / 1e9}
        peer.set_comm_mover(0)
        del snap
        if a.movers_only:
            row = {"bytes": n * 4, "world": world, "movers": movers,
                   "best_push_frac_of_link": max(m["push_gbs"] for m in movers.values()) / LINK_GBS if world > 1 else None,
                   "best_pull_frac_of_link": max(m["pull_gbs"] for m in movers.values()) / LINK_GBS if world > 1 else None}
            rows.append(row)
            if rank == 0:
                print(json.dumps(row), flush=True)
            del acc, params, grads, nccl_buf, nccl_p
            pg.close()
            pg = peer.PeerGroup(device=dev)
            nbytes *= 4
            continue
        t_push = max_ranks(timed(lambda: peer.push_add(acc, master.table, world, shard, padded), iters, pre=lambda: acc.fill_(1.0)), dev)
        pg.barrier()
        t_pull = max_ranks(timed(lambda: peer.pull(params, master.table, world, shard, padded), iters), dev)
        pg.barrier()
        t_sync = max_ranks(timed(lambda: peer.sync_rs_update_ag(grads_s.table, params_s.table, flags.table, None, rank, world, shard, lr, 0.0, 0.0,
                                                                1.0 / world), iters), dev)
        pg.barrier()
        t_nvls = None
        if world > 1:
            try:     # the same step through the NVSwitch: multimem.ld_reduce + multimem.st on multicast mappings
                sg = peer.SymmMemBuffer(padded * 4, dev)
                sp = peer.SymmMemBuffer(padded * 4, dev)
                sf = peer.SymmMemBuffer(peer.flag_words() * 4, dev)
                if sg.mc_ptr and sp.mc_ptr:
                    pl = sp.view(torch.float32, padded)
                    sg.view(torch.float32, padded).fill_(1.0)
                    pg.barrier()
                    f = lambda: peer.sync_nvls_range(sg.mc_ptr, sp.mc_ptr, pl, sf.table, 0, rank, world, rank * shard, shard, lr, 0.0, 0.0, 1.0 / world)
                    for _ in range(3):
                        f()
                    pg.barrier()
                    t_nvls = max_ranks(timed(f, iters), dev)
                    pg.barrier()
                del sg, sp, sf
            except Exception as exc:
                if rank == 0:
                    print("nvls unavailable:", exc, flush=True)

        def nccl_step():
            dist.all_reduce(nccl_buf)
            nccl_p.add_(nccl_buf, alpha=-0.01 / world)
        t_nccl = max_ranks(timed(nccl_step, iters), dev)
        t_p2p = None
        if world >= 2:
            t = torch.empty(max(nbytes, 4), dtype=torch.uint8, device=dev)
            if rank < 2:
                def p2p():
                    dist.send(t, dst=1) if rank == 0 else dist.recv(t, src=0)
                for _ in range(2):
                    p2p()
                ms = timed(p2p, iters)
            else:
                ms = 0.0
            t_p2p = max_ranks(ms, dev)
        wire = n * 4 * (world - 1) / world
        row = {"bytes": n * 4, "world": world, "movers": movers,
               "push_us": t_push * 1e3, "pull_us": t_pull * 1e3, "sync_us": t_sync * 1e3, "sync_nvls_us": (t_nvls * 1e3 if t_nvls else None),
               "nccl_allreduce_sgd_us": t_nccl * 1e3,
               "p2p_us": (t_p2p * 1e3 if t_p2p is not None else None),
               "push_gbs": wire / (t_push * 1e-3) / 1e9 if world > 1 else n * 4 / (t_push * 1e-3) / 1e9,
               "pull_gbs": wire / (t_pull * 1e-3) / 1e9 if world > 1 else n * 4 / (t_pull * 1e-3) / 1e9,
               "sync_gbs_per_dir": wire / (t_sync * 1e-3) / 1e9 if world > 1 else None,
               "p2p_gbs": (max(nbytes, 4) / (t_p2p * 1e-3) / 1e9 if t_p2p else None),
               "push_frac_of_link": (wire / (t_push * 1e-3) / 1e9 / LINK_GBS if world > 1 else None),
               "pull_frac_of_link": (wire / (t_pull * 1e-3) / 1e9 / LINK_GBS if world > 1 else None),
               "sync_vs_nccl_speedup": t_nccl / t_sync, "nvls_vs_nccl_speedup": (t_nccl / t_nvls if t_nvls else None)}
        rows.append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)
        del acc, params, grads, nccl_buf, nccl_p
        pg.close()
        pg = peer.PeerGroup(device=dev)
        nbytes *= 4
    if rank == 0 and a.out:
        os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
        with open(a.out, "w") as f:
            json.dump(rows, f, indent=1)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

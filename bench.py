#!/usr/bin/env python
"""Headline benchmark: training throughput (images/sec) of the reference's default workload —
AlexNet on CIFAR-10-shaped synthetic data (3x32x32), batch 64 per GPU, lr 0.008, asynchronous SGD with
``n_push = n_pull = 10`` (reference example/main.py:41,142-147; BASELINE.json configs[1]).

    python bench.py --gpus 1 --steps 200 --warmup 20
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...
    python bench.py --impl reference ...      # the unmodified reference (baseline/_ref), same metric/config

Own arm: one process per GPU; every GPU is a worker *and* owns 1/N of the sharded parameter server; the
training step is one CUDA-graph replay of sm_100a kernels, pushes/pulls are fused NVLink peer-memory kernels.
``value`` is device-timed (CUDA events, max over ranks) with device-resident inputs; ``e2e`` drives the public
``TrainStep`` API from pinned host batches (H2D copy every step) and reads the loss back (D2H) every step.
Inputs rotate through a pool of distinct batches larger than the 126 MB L2.

Reference arm: imports baseline/_ref/example/main.py *unmodified* and calls its ``main(args)`` — the
reference's own hot loop (main.py:57-91), optimizer and models — on a synthetic loader; timing events are
recorded by the loader at the boundaries of the timed steps.  N=1 is the reference's ``--no-distributed --cuda``
mode; N>1 is its 1 server (rank 0) + N-1 workers topology over gloo (main.py:151,165-167).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

BATCH = 64
LR = 0.008
N_PUSH = N_PULL = 10
L2_BYTES = 126 * 1024 * 1024


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--warmup", type=int, default=30)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--model", default="alexnet")
    ap.add_argument("--batch-size", type=int, default=BATCH, help="per-GPU batch (reference default 64)")
    ap.add_argument("--mode", choices=["async", "sync", "local"], default="async")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--input-hw", type=int, default=32, help="image side (32 = CIFAR shape of the headline config; 224 for ResNet-50)")
    ap.add_argument("--classes", type=int, default=10)
    ap.add_argument("--momentum", type=float, default=0.0)
    ap.add_argument("--staleness", default=None)
    ap.add_argument("--server-gain", default="auto", help="async + --momentum: auto = (1 - mu) / workers, none = 1, or a number")
    ap.add_argument("--pull-lag", type=int, default=None, help="async mode: steps between issuing a pull and landing it (default: the optimizer's)")
    ap.add_argument("--dtype", choices=["tf32", "bf16"], default=None,
                    help="CUDA compute dtype (default: tf32 for the reference's LeNet / AlexNet = the reference's own precision, bf16 for the ResNets)")
    ap.add_argument("--no-second-dtype", action="store_true", help="skip the short extra measurement in the other compute dtype (1 GPU only)")
    a = ap.parse_args()
    a.pool_cap_bytes = 2 << 30      # host pool of distinct batches never exceeds 2 GiB
    return a


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


def published_baseline():
    try:
        with open(os.path.join(ROOT, "BASELINE.json")) as f:
            pub = json.load(f).get("published", {})
        v = pub.get("images_per_sec") if isinstance(pub, dict) else None
        return float(v) if v else None
    except Exception:
        return None


def max_over_ranks(ms: float, world: int, device) -> float:
    import torch
    import torch.distributed as dist
    if world == 1:
        return ms
    t = torch.tensor([ms], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def barrier(world: int):
    import torch
    import torch.distributed as dist
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


# ------------------------------------------------------------------------------------------------
# own arm
# ------------------------------------------------------------------------------------------------
def _measure(args, dtype, rank, local_rank, world, device, steps, warmup, with_e2e=True) -> dict:
    """Build model + optimizer + TrainStep in compute dtype ``dtype`` and time ``steps`` steps (device-timed, then end to end)."""
    import torch

    from distributed_ml_pytorch_b200 import _native
    from distributed_ml_pytorch_b200.engine import TrainStep
    from distributed_ml_pytorch_b200.models import build_model
    from distributed_ml_pytorch_b200.utils.timing import ClockSampler

    torch.manual_seed(1234)  # identical initial replicas on every rank
    B = args.batch_size
    HW, NC = args.input_hw, args.classes
    model = build_model(args.model, num_classes=NC, input_hw=HW, dtype=dtype).to(device)
    model.train()
    optimizer, parallelism = make_optimizer(args, model, rank, world, device)
    ts = TrainStep(model, optimizer, B, (3, HW, HW), device=device, use_graph=not args.no_graph)

    # rotating input pool larger than L2 (host pinned for e2e, device copy for the device-timed number)
    torch.manual_seed(100 + rank)
    bytes_per_batch = B * 3 * HW * HW * 4 + B * 8
    pool = max(8, (L2_BYTES * 5 // 4) // bytes_per_batch + 1)
    hx = torch.randn(pool, B, 3, HW, HW).pin_memory()
    hy = torch.randint(0, NC, (pool, B), dtype=torch.int64).pin_memory()
    dx, dy = hx.to(device), hy.to(device)

    K_, W_ = steps, max(warmup, 3)
    for i in range(W_):
        ts(hx[i % pool], hy[i % pool])
    ts.flush()

    sampler = ClockSampler(gpu_index=torch.cuda.current_device() if "CUDA_VISIBLE_DEVICES" not in os.environ else local_rank).start()
    # ---- device-timed, device-resident inputs ----
    barrier(world)
    launches0 = _native.launch_count()
    pushes0 = getattr(optimizer, "pushes", 0)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ts.stream)
    for i in range(K_):
        ts(dx[i % pool], dy[i % pool])
    e1.record(ts.stream)
    last = ts.flush()
    barrier(world)
    dev_ms = max_over_ranks(e0.elapsed_time(e1), world, device)
    eager_launches = _native.launch_count() - launches0
    timed_pushes = getattr(optimizer, "pushes", 0) - pushes0
    res = {"dev_ms": dev_ms, "kernels_per_step": int(ts.kernels_per_step), "gpu_launches": int(ts.kernels_per_step * K_ + eager_launches),
           "final_loss": (last.loss if last is not None else None), "parallelism": parallelism, "pool": pool, "bytes_per_batch": bytes_per_batch,
           "cuda_graph": ts.graph is not None, "timed_push_intervals": int(timed_pushes), "steps": K_, "warmup": W_}
    if with_e2e:
        # ---- end to end: pinned host -> H2D each step, loss read back each step ----
        barrier(world)
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record(ts.stream)
        seen = 0
        loss_sum = 0.0
        for i in range(K_):
            r = ts(hx[i % pool], hy[i % pool])
            if r is not None:
                loss_sum += r.loss
                seen += 1
        r = ts.flush()
        if r is not None:
            loss_sum += r.loss
            seen += 1
        f1.record(ts.stream)
        f1.synchronize()
        barrier(world)
        res["e2e_ms"] = max_over_ranks(f0.elapsed_time(f1), world, device)
        res["mean_loss"] = loss_sum / seen if seen else None
    res["clocks"] = sampler.stop()
    if hasattr(optimizer, "finish"):
        optimizer.finish()
    if hasattr(optimizer, "close"):
        optimizer.close()
    del ts, optimizer, model, dx, dy, hx, hy
    torch.cuda.empty_cache()
    return res


def run_ours(args) -> dict:
    import torch
    import torch.distributed as dist

    from distributed_ml_pytorch_b200.models import default_dtype

    rank, local_rank, world = dist_env()
    if world != args.gpus:
        if args.gpus > 1 and world == 1:
            raise SystemExit("launch multi-GPU runs with torch.distributed.run (see module docstring)")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=device)
    dtype = args.dtype or default_dtype(args.model)
    B = args.batch_size
    HW = args.input_hw
    m = _measure(args, dtype, rank, local_rank, world, device, args.steps, args.warmup)
    K_, W_ = m["steps"], m["warmup"]
    dev_ms, e2e_ms, clocks = m["dev_ms"], m["e2e_ms"], m["clocks"]
    global_batch = B * world
    value = global_batch * K_ / (dev_ms / 1e3)
    e2e = global_batch * K_ / (e2e_ms / 1e3)
    pub = published_baseline()
    async_mode = "async" in m["parallelism"]
    out = {
        "metric": "images/sec (device-timed, max over ranks)",
        "value": round(value, 1), "unit": "images/sec", "n_gpus": world, "steps": K_, "warmup": W_,
        "ms_per_step": round(dev_ms / K_, 5), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": (round(value / pub, 3) if pub else None), "dtype": dtype, "data": "synthetic",
        "impl": "ours",
        "config": {"model": args.model, "global_batch": global_batch, "per_gpu_batch": B, "input": f"3x{HW}x{HW}", "seq_len": None,
                   "parallelism": m["parallelism"], "lr": LR, "n_push": N_PUSH, "n_pull": N_PULL,
                   "precision": ("fp32 parameters / activations / gradients, tensor-core products in tf32 with fp32 accumulation "
                                 "(the reference: fp32 tensors, cuDNN TF32 convolutions)" if dtype == "tf32" else
                                 "fp32 master parameters and gradients, bf16 GEMM operands and activations, fp32 accumulation"),
                   "l2_policy": f"inputs rotate through a pool of {m['pool']} distinct batches ({m['pool'] * m['bytes_per_batch'] >> 20} MiB > 126 MiB L2)",
                   "cuda_graph": m["cuda_graph"], "loss_readback": "every step, read one step late from a pinned ring the loss kernel writes (no D2H copy on the step stream)"},
        "clocks": {"sm_mhz": clocks.get("sm_mhz"), "sm_max_mhz": clocks.get("sm_max_mhz"), "reasons": clocks.get("reasons", [])},
        "e2e": {"value": round(e2e, 1), "unit": "images/sec", "h2d_bytes_per_step": m["bytes_per_batch"], "d2h_bytes_per_step": 8,
                "ms_per_step": round(e2e_ms / K_, 5), "mean_loss": m["mean_loss"]},
        "gpu_launches": m["gpu_launches"],
        "kernels_per_step": m["kernels_per_step"],
        "final_loss": m["final_loss"],
    }
    if async_mode:
        out["timed_push_intervals"] = m["timed_push_intervals"]
        if m["timed_push_intervals"] < 3:
            out["note"] = (f"the device-timed region of {K_} steps holds only {m['timed_push_intervals']} push/pull interval(s) (n_push = {N_PUSH}); "
                           "run with --steps >= 30 for a value that averages over the communication cadence")
    if world == 1 and not args.no_second_dtype and args.model in ("alexnet", "lenet"):
        # context, not the headline: the same step in the other compute dtype (short, device-timed only)
        other = "bf16" if dtype == "tf32" else "tf32"
        try:
            o = _measure(args, other, rank, local_rank, world, device, min(args.steps, 100), args.warmup, with_e2e=False)
            out["other_dtype"] = {"dtype": other, "value": round(global_batch * o["steps"] / (o["dev_ms"] / 1e3), 1), "unit": "images/sec",
                                  "ms_per_step": round(o["dev_ms"] / o["steps"], 5), "steps": o["steps"], "kernels_per_step": o["kernels_per_step"]}
        except Exception as exc:  # the headline number must not depend on the context measurement
            out["other_dtype"] = {"dtype": other, "error": f"{type(exc).__name__}: {exc}"[:200]}
    if world > 1:
        dist.destroy_process_group()
    return out if rank == 0 else {}


def make_optimizer(args, model, rank, world, device):
    from distributed_ml_pytorch_b200.parallel.fused_sgd import FusedSGD
    if args.mode == "local" or (world == 1 and args.mode != "async"):
        return FusedSGD(model.parameters(), lr=LR, model=model), "single-gpu fused SGD"
    try:
        from distributed_ml_pytorch_b200.parallel.async_optim import Asynchronous
        from distributed_ml_pytorch_b200.parallel.sync_optim import Synchronous
    except ImportError:
        return FusedSGD(model.parameters(), lr=LR, model=model), "single-gpu fused SGD"
    if args.mode == "sync":
        opt = Synchronous(model.parameters(), lr=LR, model=model, momentum=args.momentum)
        how = "NVLS multimem.ld_reduce/multimem.st" if getattr(opt, "nvls", False) else "NVLink peer loads/stores"
        return opt, f"sync-dp{world} fused reduce-scatter+update+all-gather ({how}), 4 buckets overlapped with backward"
    opt = Asynchronous(model.parameters(), lr=LR, n_push=N_PUSH, n_pull=N_PULL, model=model, momentum=args.momentum, staleness=args.staleness,
                       server_gain=(None if args.server_gain == "none" else args.server_gain if args.server_gain == "auto" else float(args.server_gain)),
                       **({} if args.pull_lag is None else {"pull_lag": args.pull_lag}))
    return opt, f"async-ps dp{world}, parameter server sharded over {world} GPU(s), fused NVLink push/pull"


# ------------------------------------------------------------------------------------------------
# reference arm
# ------------------------------------------------------------------------------------------------
class _TimedSyntheticLoader:
    """Synthetic CIFAR-shaped loader handed to the reference's main(); records CUDA events exactly around the
    timed steps (it is data, not code: the reference's loop, optimizer and models run untouched)."""

    def __init__(self, batch, warmup, steps, seed):
        import torch
        g = torch.Generator().manual_seed(seed)
        self.n = warmup + steps
        self.warmup, self.steps = warmup, steps
        pool = max(8, (L2_BYTES * 5 // 4) // (batch * 3 * 32 * 32 * 4) + 1)
        self.x = torch.randn(pool, batch, 3, 32, 32, generator=g)
        self.y = torch.randint(0, 10, (pool, batch), generator=g)
        self.pool = pool
        self.e0 = torch.cuda.Event(enable_timing=True)
        self.e1 = torch.cuda.Event(enable_timing=True)
        self.ms = None
        self.before_timed = None

    def __len__(self):
        return self.n

    def __iter__(self):
        import torch
        for i in range(self.n):
            if i == self.warmup:
                if self.before_timed is not None:
                    self.before_timed()
                torch.cuda.synchronize()
                self.e0.record()
            yield self.x[i % self.pool], self.y[i % self.pool]
        self.e1.record()
        torch.cuda.synchronize()
        self.ms = self.e0.elapsed_time(self.e1)


def run_reference(args) -> dict:
    ref_root = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.exists(os.path.join(ref_root, "example", "main.py")):
        return {"impl": "reference", "unavailable": "baseline/_ref is not populated (run python baseline/install_ref.py where /root/reference exists)"}
    try:
        import torch
        import torch.distributed as dist
        sys.path[:0] = [ref_root, os.path.join(ref_root, "example")]
        import warnings
        warnings.filterwarnings("ignore")
        import main as refmain  # the reference's example/main.py, unmodified
    except Exception as exc:  # pragma: no cover
        return {"impl": "reference", "unavailable": f"reference import failed: {type(exc).__name__}: {exc}"[:300]}

    from distributed_ml_pytorch_b200.utils.timing import ClockSampler
    rank, local_rank, world = dist_env()
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    K_, W_ = args.steps, max(args.warmup, 3)
    B = args.batch_size
    ns = argparse.Namespace(batch_size=B, test_batch_size=B, epochs=1, lr=LR, num_pull=N_PULL, num_push=N_PUSH, cuda=True,
                            log_interval=10 ** 9, no_distributed=(world == 1), rank=rank, world_size=world, server=(world > 1 and rank == 0),
                            master="127.0.0.1", port=os.environ.get("MASTER_PORT", "29500"))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("gloo", rank=rank, world_size=world)  # the backend the reference hard-codes (main.py:165)
    loader = _TimedSyntheticLoader(B, W_, K_, seed=100 + rank)
    test_loader = [(loader.x[0][:8], loader.y[0][:8])]
    refmain.get_dataset = lambda a, transform: (loader, test_loader)
    n_workers = max(1, world - 1)
    sampler = None
    note = ""
    n_params = sum(p.numel() for p in refmain.AlexNet().parameters())
    wg = dist.new_group(ranks=list(range(1, world))) if world > 2 else None  # collective: every rank creates it
    if world > 1 and rank == 0:
        # reference server role (main.py:166-167 -> init_server() -> ParameterServer(model).run()).  The serve loop
        # is run here instead of server.run() only so that it can end: the reference server never exits (§3.2).
        from asgd.utils.messaging import MessageCode, send_message
        import asgd.server as srv
        server = srv.ParameterServer(model=refmain.AlexNet())
        done = 0
        while done < n_workers:
            dist.recv(tensor=server.m_parameter)
            code = MessageCode(int(server.m_parameter[1].item()))
            if code == MessageCode.Shutdown:
                done += 1
                continue
            server.receive(int(server.m_parameter[0].item()), code, server.m_parameter[2:])
        for w in range(1, world):  # stop the workers' listener threads before any collective runs
            send_message(MessageCode.Shutdown, torch.zeros(n_params), dst=w)
        from asgd.utils import messaging as _m
        for wk, _ in _m._keep:
            wk.wait()
        ms = 0.0
    else:
        if rank == (1 if world > 1 else 0):
            sampler = ClockSampler(gpu_index=local_rank).start()
        if wg is not None:
            loader.before_timed = lambda: dist.barrier(group=wg)  # workers enter the timed region together
        try:
            refmain.main(ns)
        except Exception as exc:
            # the reference's end-of-epoch evaluate() hands CUDA tensors to sklearn (main.py:127) and raises;
            # that happens after the timed region has been closed by the loader.
            note = f"after the timed region the reference raised {type(exc).__name__} (its evaluate()/CSV tail); ignored"
            if loader.ms is None:
                return {"impl": "reference", "unavailable": f"reference hot loop failed: {type(exc).__name__}: {exc}"[:300]}
        ms = loader.ms
        if world > 1:
            from asgd.utils.messaging import MessageCode, send_message
            from asgd.utils import messaging as _m
            send_message(MessageCode.Shutdown, torch.zeros(n_params))
            for wk, _ in _m._keep:
                wk.wait()
            time.sleep(1.0)  # let the listener thread consume the server's Shutdown before collectives start
    clocks = sampler.stop() if sampler is not None else {}
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        obj = [clocks]
        dist.broadcast_object_list(obj, src=1)  # clocks were sampled next to the first worker
        clocks = obj[0]
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return {}
    imgs = B * n_workers * K_
    value = imgs / (ms / 1e3)
    pub = published_baseline()
    bytes_per_batch = B * 3 * 32 * 32 * 4 + B * 8
    return {
        "metric": "images/sec (device-timed, max over ranks)", "value": round(value, 1), "unit": "images/sec", "n_gpus": world,
        "steps": K_, "warmup": W_, "ms_per_step": round(ms / K_, 5), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": (round(value / pub, 3) if pub else None), "dtype": "fp32", "data": "synthetic", "impl": "reference",
        "config": {"model": "alexnet (reference example/models.py, unmodified)", "global_batch": B * n_workers, "per_gpu_batch": B,
                   "input": "3x32x32", "seq_len": None,
                   "parallelism": ("reference --no-distributed --cuda (optim.SGD)" if world == 1 else
                                   f"reference async: 1 server (rank 0) + {n_workers} worker(s), gloo, n_push=n_pull={N_PUSH}"),
                   "lr": LR, "l2_policy": f"inputs rotate through a pool of {loader.pool} distinct host batches",
                   "completion": "shipped files byte-identical; ABSENT modules supplied (baseline/ref_completion)", "note": note},
        "clocks": {"sm_mhz": clocks.get("sm_mhz"), "sm_max_mhz": clocks.get("sm_max_mhz"), "reasons": clocks.get("reasons", [])},
        "e2e": {"value": round(value, 1), "unit": "images/sec", "h2d_bytes_per_step": bytes_per_batch, "d2h_bytes_per_step": 4,
                "note": "the reference's loop is end to end by construction: .cuda() of each batch and loss.item() every step (main.py:65,79)"},
        "gpu_launches": None,
    }


def _relaunch_under_torchrun(args) -> None:
    """`python bench.py --gpus N` (N > 1) without a launcher: re-execute under torch.distributed.run, one rank per GPU."""
    import socket
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.abspath(__file__)] + sys.argv[1:]
    os.execv(sys.executable, cmd)


def main():
    args = parse()
    import torch
    if not torch.cuda.is_available():
        # nothing is measured without a B200: say so in the output contract instead of dying inside CUDA initialisation
        print(json.dumps({"impl": args.impl, "unavailable": "no CUDA device visible; bench.py measures on B200 GPUs (CPU numbers: tools/bench_b1_cpu.py)"}),
              flush=True)
        sys.exit(0 if args.impl == "reference" else 1)
    if args.gpus > 1 and "RANK" not in os.environ and "WORLD_SIZE" not in os.environ:
        _relaunch_under_torchrun(args)
    if args.impl == "reference":
        try:
            out = run_reference(args)
        except Exception as exc:
            rank = dist_env()[0]
            out = {"impl": "reference", "unavailable": f"{type(exc).__name__}: {exc}"[:300]} if rank == 0 else {}
    else:
        out = run_ours(args)
    if out:
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Training example — same command line as the reference's example/main.py (every flag and default of
main.py:142-155 is kept), running on distributed_ml_pytorch_b200.

    python example/main.py --no-distributed                       # make single   (plain SGD, CPU)
    python example/main.py --no-distributed --cuda                # make gpu      (fused SGD, one B200)
    python example/main.py --rank 0 --world-size 3 --server       # make server   (CPU/gloo parameter server)
    python example/main.py --rank 1 --world-size 3                # make first / second (workers)
    torchrun --nproc-per-node 8 example/main.py --cuda            # 8 x B200: sharded PS over NVLink (rank/world from env)

Additive flags: --model, --mode {async,sync}, --momentum, --staleness, --synthetic, --ckpt, --resume,
--steps-per-epoch, --no-graph, --input-hw, --in-channels, --seed.

Differences from the reference's script, all listed in SURVEY.md §7.5: the model is selectable (reference hard-codes
AlexNet, main.py:41,136); ``log/`` is created; ``net.train()`` is restored after evaluation; accuracy covers every test
batch; the server exits when all workers are done and does not fall through into ``main(args)``; the optimizer is
built after the model has moved to the GPU.
"""
from __future__ import annotations

import argparse
import os
import sys
import time
from datetime import datetime

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.distributed as dist
import torch.optim as optim

from asgd.optim import Asynchronous, FusedSGD, Synchronous
from asgd.server import ParameterServer
from distributed_ml_pytorch_b200.models import MODEL_NAMES, build_model
from distributed_ml_pytorch_b200.ops import cross_entropy
from distributed_ml_pytorch_b200.utils.checkpoint import load_model_state, load_optimizer_state, read_checkpoint, save_checkpoint
from distributed_ml_pytorch_b200.utils import trace
from distributed_ml_pytorch_b200.utils.csvlog import CSVLogger, log_path

CLASSES = ("plane", "car", "bird", "cat", "deer", "dog", "frog", "horse", "ship", "truck")


class SyntheticLoader:
    """Deterministic CIFAR-shaped batches (no network on this machine); pinned on request for async H2D."""

    def __init__(self, n_batches: int, batch_size: int, channels: int, hw: int, seed: int, pin: bool = False, classes: int = 10):
        g = torch.Generator().manual_seed(seed)
        pool = min(n_batches, 64)
        self.x = torch.randn(pool, batch_size, channels, hw, hw, generator=g)
        self.y = torch.randint(0, classes, (pool, batch_size), generator=g)
        if pin and torch.cuda.is_available():
            self.x, self.y = self.x.pin_memory(), self.y.pin_memory()
        self.n = n_batches

    def __len__(self):
        return self.n

    def __iter__(self):
        for i in range(self.n):
            yield self.x[i % self.x.shape[0]], self.y[i % self.y.shape[0]]


def get_dataset(args, transform=None):
    """CIFAR-10 through torchvision when it is already on disk (reference main.py:23-29), else synthetic."""
    on_disk = not args.synthetic and os.path.isdir(os.path.join(args.data_root, "cifar-10-batches-py"))
    if on_disk and args.cuda and not args.host_data:
        # B200 path: the 150 MB uint8 image set lives in HBM; shuffling, ToTensor and Normalize are one kernel per batch.
        from distributed_ml_pytorch_b200.utils.data import GpuResidentLoader, cifar10_arrays
        device = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
        xs, ys = cifar10_arrays(args.data_root, train=True)
        xt, yt = cifar10_arrays(args.data_root, train=False)
        train = GpuResidentLoader(xs, ys, args.batch_size, device, shuffle=True, drop_last=True, seed=args.seed + 1000 * (args.rank or 0))
        test = GpuResidentLoader(xt, yt, args.test_batch_size, device, shuffle=False, drop_last=False)
        return train, test
    if on_disk:
        import torchvision
        import torchvision.transforms as transforms
        transform = transform or transforms.Compose([transforms.ToTensor(), transforms.Normalize((0.5, 0.5, 0.5), (0.5, 0.5, 0.5))])
        trainset = torchvision.datasets.CIFAR10(root=args.data_root, train=True, download=False, transform=transform)
        testset = torchvision.datasets.CIFAR10(root=args.data_root, train=False, download=False, transform=transform)
        trainloader = torch.utils.data.DataLoader(trainset, batch_size=args.batch_size, shuffle=True, num_workers=1, drop_last=True,
                                                  pin_memory=args.cuda)
        testloader = torch.utils.data.DataLoader(testset, batch_size=args.test_batch_size, shuffle=False, num_workers=1)
        return trainloader, testloader
    steps = args.steps_per_epoch or 782            # 50000 / 64, the reference's iterations per epoch
    seed = args.seed + 1000 * (args.rank or 0)      # every worker draws its own stream, like independent shuffling
    train = SyntheticLoader(steps, args.batch_size, args.in_channels, args.input_hw, seed, pin=args.cuda)
    test = SyntheticLoader(max(1, min(4, 10000 // max(1, args.test_batch_size)) or 1), min(args.test_batch_size, 1000), args.in_channels,
                           args.input_hw, seed + 7)
    return train, test


def _gain(args):
    g = getattr(args, "server_gain", "auto")
    return None if g in (None, "none") else g if g == "auto" else float(g)


def make_optimizer(args, net):
    if args.no_distributed:
        if args.cuda:
            return FusedSGD(net.parameters(), lr=args.lr, momentum=args.momentum, model=net)
        return optim.SGD(net.parameters(), lr=args.lr, momentum=args.momentum)     # reference main.py:44
    if args.mode == "sync":
        return Synchronous(net.parameters(), lr=args.lr, model=net, momentum=args.momentum)
    servers = list(range(args.num_servers)) if (args.num_servers > 1 and not args.cuda) else None
    return Asynchronous(net.parameters(), lr=args.lr, n_push=args.num_push, n_pull=args.num_pull, model=net, momentum=args.momentum,
                        staleness=args.staleness, server_ranks=servers, server_gain=_gain(args))


def main(args):
    device = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0))) if args.cuda else torch.device("cpu")
    if args.cuda:
        torch.cuda.set_device(device)
    trainloader, testloader = get_dataset(args)
    torch.manual_seed(args.seed)
    net = build_model(args.model, in_channels=args.in_channels, input_hw=args.input_hw, dtype=args.dtype).to(device)
    net.train()
    start_epoch, resumed_opt = 0, None
    if args.resume:
        # The model is restored BEFORE the optimizer exists: a distributed optimizer seeds the parameter server (the sharded masters
        # on NVLink, the server rank's vector on gloo) from the model it is constructed with.
        sd, meta, resumed_opt = read_checkpoint(args.resume)
        load_model_state(net, sd)
        start_epoch = int(meta.get("epoch", -1)) + 1
    optimizer = make_optimizer(args, net)
    load_optimizer_state(optimizer, resumed_opt)
    scheduler = optim.lr_scheduler.LambdaLR(optimizer, lr_lambda=lambda epoch: 1 / (epoch + 1))   # reference main.py:47-48

    rank = dist.get_rank() if (dist.is_available() and dist.is_initialized()) else None
    logger = CSVLogger(log_path(args.no_distributed, args.cuda, rank))
    step_fn = None
    if args.cuda and not args.no_graph and hasattr(optimizer, "arena"):
        from distributed_ml_pytorch_b200.engine import TrainStep
        step_fn = TrainStep(net, optimizer, args.batch_size, (args.in_channels, args.input_hw, args.input_hw), device=device)

    t_train = time.time()
    for epoch in range(start_epoch, args.epochs):
        print("Training for epoch {}".format(epoch))
        t_epoch = time.time()
        seen = 0
        for i, (inputs, labels) in enumerate(trainloader, 0):
            if step_fn is not None:
                with trace.span("step (enqueue)", "train", epoch=epoch, iteration=i):
                    res = step_fn(inputs, labels)            # H2D copy + graph replay; result of the previous step
                loss_value = res.loss if res is not None else float("nan")
            else:
                with trace.span("step", "train", epoch=epoch, iteration=i):
                    inputs, labels = inputs.to(device, non_blocking=True), labels.to(device, non_blocking=True)
                    optimizer.zero_grad()
                    with trace.span("forward+loss", "train"):
                        outputs = net(inputs)
                        loss = cross_entropy(outputs, labels)
                    with trace.span("backward", "train"):
                        loss.backward()
                    with trace.span("optimizer.step", "train"):
                        optimizer.step()
                    loss_value = loss.item()
            seen += inputs.shape[0]
            test_loss = test_acc = None
            if i % args.log_interval == 0 and i > 0:
                _sync_step(step_fn, device, before=True)     # the graph step runs on its own stream: let it finish before evaluating
                with trace.span("evaluate", "eval", iteration=i):
                    test_loss, test_acc = evaluate(net, testloader, args, device=device)
                _sync_step(step_fn, device, before=False)
                print("Timestamp: {} | Iteration: {:6} | Loss: {:6.4f} | Test Loss: {:6.4f} | Test Accuracy: {:6.4f}".format(
                    datetime.now(), i, loss_value, test_loss, test_acc))
            logger.append(i, loss_value, test_loss, test_acc, images_per_sec=seen / max(1e-9, time.time() - t_epoch))
        if step_fn is not None:
            step_fn.flush()
        _sync_step(step_fn, device, before=True)
        print("epoch {}: {} images in {:.2f} s = {:.0f} images/sec".format(epoch, seen, time.time() - t_epoch, seen / max(1e-9, time.time() - t_epoch)))
        print("learning rate is : {}".format(optimizer.param_groups[0]["lr"]))
        evaluate(net, testloader, args, verbose=True, device=device)
        _sync_step(step_fn, device, before=False)
        if args.step_scheduler:
            scheduler.step()
        ck = _ckpt_path(args, rank, optimizer)
        if ck:
            save_checkpoint(ck, net, optimizer, epoch=epoch)
    print("wrote", logger.to_csv())
    if hasattr(optimizer, "finish"):
        optimizer.finish()
    if args.trace:
        print("wrote", trace.save(_trace_path(args.trace, rank)))
    if hasattr(optimizer, "pushes"):
        stats = optimizer.staleness_stats() if hasattr(optimizer, "staleness_stats") else {}
        n_bytes = 4 * sum(p.numel() for p in net.parameters())
        wall = max(1e-9, time.time() - t_train)
        print("pushes: {} | pulls: {} | {:.1f} MB pushed, {:.1f} MB pulled ({:.1f} MB/s of parameter traffic over {:.1f} s){}".format(
            optimizer.pushes, optimizer.pulls, optimizer.pushes * n_bytes / 1e6, optimizer.pulls * n_bytes / 1e6,
            (optimizer.pushes + optimizer.pulls) * n_bytes / 1e6 / wall, wall, " | staleness: {}".format(stats) if stats else ""))
    if hasattr(optimizer, "close"):
        optimizer.close()                 # NVLink-shared buffers and their IPC mappings
    print("Finished Training")


def _ckpt_path(args, rank, optimizer):
    """Who writes --ckpt.  One process: the path itself.  NVLink transport / synchronous mode: rank 0 alone (it writes the server's
    parameters, every replica of a synchronous run is identical).  CPU/gloo asynchronous: the server rank owns ``<path>`` (its final
    parameters), every worker writes its replica + optimizer state to ``<path>.rank<r>`` — no two processes share a file."""
    if not args.ckpt:
        return None
    if rank is None:
        return args.ckpt
    if getattr(optimizer, "transport", None) == "gloo":
        return f"{args.ckpt}.rank{rank}"
    return args.ckpt if rank == 0 else None


def _trace_path(path, rank):
    """One trace file per process: out.json -> out.rank3.json in a distributed run."""
    if rank is None:
        return path
    root, ext = os.path.splitext(path)
    return f"{root}.rank{rank}{ext or '.json'}"


def _sync_step(step_fn, device, before: bool) -> None:
    """Order the TrainStep stream against the default stream around an evaluation (both touch the weights and their shadows)."""
    if step_fn is None or device.type != "cuda":
        return
    if before:
        step_fn.synchronize()
    else:
        step_fn.stream.wait_stream(torch.cuda.current_stream(device))


def evaluate(net, testloader, args, verbose=False, device=None):
    device = device or next(net.parameters()).device
    was_training = net.training
    net.eval()
    test_loss, correct, total, n_batches = 0.0, 0, 0, 0
    preds, trues = [], []
    with torch.no_grad():
        for images, labels in testloader:
            images, labels = images.to(device), labels.to(device)
            outputs = net(images)
            predicted = outputs.float().argmax(1)
            test_loss += float(cross_entropy(outputs, labels))
            correct += int((predicted == labels).sum())
            total += labels.numel()
            n_batches += 1
            if verbose:
                preds.append(predicted.cpu())
                trues.append(labels.cpu())
    net.train(was_training)
    test_accuracy = correct / max(1, total)
    if verbose:
        print("Loss: {:.3f}".format(test_loss))
        print("Accuracy: {:.3f}".format(test_accuracy))
        try:
            from sklearn.metrics import classification_report
            print(classification_report(torch.cat(trues).numpy(), torch.cat(preds).numpy(), labels=list(range(10)), target_names=CLASSES,
                                        zero_division=0))
        except Exception as exc:  # sklearn is optional
            print("(classification report unavailable: %s)" % exc)
    return test_loss, test_accuracy


def init_server(args):
    """Server role (reference main.py:135-138).  CPU/gloo: blocking message loop; --cuda: NVLink shard owner."""
    model = build_model(args.model, in_channels=args.in_channels, input_hw=args.input_hw)
    if args.cuda:
        device = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
        torch.cuda.set_device(device)
        torch.manual_seed(args.seed)
        ParameterServer.run_peer(model.to(device), momentum=args.momentum, staleness=args.staleness, gain=_gain(args))
        return
    server = ParameterServer(model=model, momentum=args.momentum, staleness=args.staleness, gain=_gain(args), shard_index=args.rank if args.num_servers > 1 else 0,
                             n_shards=args.num_servers, idle_timeout_s=args.server_idle_timeout_s)
    server.run()
    if args.trace:
        print("wrote", trace.save(_trace_path(args.trace, args.rank)))
    if args.ckpt:
        server.save(args.ckpt)
    print("server done:", server.counters, "| staleness histogram (updates applied at version lag tau):", dict(sorted(server.staleness_hist.items())))


def build_parser():
    parser = argparse.ArgumentParser(description="Distbelief training example")
    parser.add_argument("--batch-size", type=int, default=64, metavar="N", help="input batch size for training (default: 64)")
    parser.add_argument("--test-batch-size", type=int, default=10000, metavar="N", help="input batch size for testing (default: 10000)")
    parser.add_argument("--epochs", type=int, default=20, metavar="N", help="number of epochs to train (default: 20)")
    parser.add_argument("--lr", type=float, default=0.008, metavar="LR", help="learning rate (default: 0.008)")
    parser.add_argument("--num-pull", type=int, default=10, metavar="N", help="how often to pull params (default: 10)")
    parser.add_argument("--num-push", type=int, default=10, metavar="N", help="how often to push grads (default: 10)")
    parser.add_argument("--cuda", action="store_true", default=False, help="use CUDA for training")
    parser.add_argument("--log-interval", type=int, default=100, metavar="N", help="how often to evaluate and print out")
    parser.add_argument("--no-distributed", action="store_true", default=False, help="whether to use AsynchronousSGD or normal SGD")
    parser.add_argument("--rank", type=int, metavar="N", help="rank of current process (0 is server, 1+ is training node)")
    parser.add_argument("--world-size", type=int, default=3, metavar="N", help="size of the world")
    parser.add_argument("--server", action="store_true", default=False, help="server node?")
    parser.add_argument("--master", type=str, default="localhost", help="ip address of the master (server) node")
    parser.add_argument("--port", type=str, default="29500", help="port on master node to communicate with")
    # additive flags
    parser.add_argument("--model", default="alexnet", choices=MODEL_NAMES)
    parser.add_argument("--mode", default="async", choices=["async", "sync"])
    parser.add_argument("--dtype", default=None, choices=["tf32", "bf16"],
                        help="CUDA compute dtype: tf32 = fp32 tensors with TF32 tensor-core products, the reference's own precision (default for "
                        "lenet / alexnet); bf16 = bf16 operands and activations over fp32 masters (default for the resnets)")
    parser.add_argument("--momentum", type=float, default=0.0)
    parser.add_argument("--staleness", default=None, choices=[None, "none", "inverse", "exp"])
    parser.add_argument("--server-gain", default="auto", help="gain of the server's momentum update: auto = (1 - momentum) / workers (the "
                        "step size of plain SGD whatever the worker count), none = 1 (classic), or a number; no effect without --momentum")
    parser.add_argument("--backend", default=None, help="torch.distributed backend (default: nccl with --cuda, gloo otherwise)")
    parser.add_argument("--trace", default=None, help="write a Chrome / Perfetto trace of this process (steps, push / pull, server "
                        "messages) to this path; distributed runs add .rank<N>")
    parser.add_argument("--server-idle-timeout-s", type=float, default=None, help="CPU/gloo server: exit with an error when no worker "
                        "message arrives for this long (default: wait forever, like the reference)")
    parser.add_argument("--num-servers", type=int, default=1, help="CPU/gloo: shard the parameter server over ranks 0..K-1 (each started "
                        "with --server); on GPUs the server is always sharded over all ranks")
    parser.add_argument("--host-data", action="store_true", help="keep the torchvision DataLoader on the host even with --cuda "
                        "(default with --cuda: the dataset is resident on the GPU, utils/data.py)")
    parser.add_argument("--synthetic", action="store_true", help="synthetic CIFAR-shaped data even if ./data holds CIFAR-10")
    parser.add_argument("--data-root", default="./data")
    parser.add_argument("--steps-per-epoch", type=int, default=None)
    parser.add_argument("--input-hw", type=int, default=32)
    parser.add_argument("--in-channels", type=int, default=3)
    parser.add_argument("--seed", type=int, default=1234)
    parser.add_argument("--ckpt", default=None, help="save a state_dict checkpoint here after every epoch")
    parser.add_argument("--resume", default=None)
    parser.add_argument("--no-graph", action="store_true", help="eager kernels instead of the CUDA-graph TrainStep")
    parser.add_argument("--timeout-s", type=int, default=1800,
                        help="process-group timeout: a dead peer makes send/recv raise after this long instead of hanging (the reference has none)")
    parser.add_argument("--step-scheduler", action="store_true", help="step the 1/(epoch+1) LambdaLR each epoch (the reference never does)")
    return parser


if __name__ == "__main__":
    args = build_parser().parse_args()
    print(args)
    if args.trace:
        trace.start(process_name=("server" if args.server else "worker") + ("" if args.rank is None else f" rank {args.rank}"))
    if not args.no_distributed:
        # Initialise the distributed environment; server and clients call this as an entry point (reference main.py:159-167).
        if args.rank is None:
            args.rank = int(os.environ.get("RANK", 0))
            args.world_size = int(os.environ.get("WORLD_SIZE", args.world_size))
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1" if args.master == "localhost" else args.master)
        os.environ.setdefault("MASTER_PORT", args.port)
        backend = args.backend or ("nccl" if args.cuda else "gloo")
        kw = {}
        if backend == "nccl":
            dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", args.rank % max(1, torch.cuda.device_count()))))
            os.environ.setdefault("LOCAL_RANK", str(dev.index))
            torch.cuda.set_device(dev)
            kw["device_id"] = dev
        from datetime import timedelta
        dist.init_process_group(backend, rank=args.rank, world_size=args.world_size, timeout=timedelta(seconds=args.timeout_s), **kw)
        if args.server and args.mode == "sync":
            raise SystemExit("--mode sync has no server rank (every rank trains and the update is an all-reduce): drop --server")
        if args.server:
            init_server(args)
            dist.destroy_process_group()
            sys.exit(0)
    main(args)
    if dist.is_available() and dist.is_initialized():
        dist.destroy_process_group()

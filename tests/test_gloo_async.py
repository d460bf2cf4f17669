"""Distributed tests without GPUs: mp.spawn + gloo on localhost, exactly the reference's own way of running
"multi-node" (pytorch_p2p_ex.py:26-36).  pytest-timeout guards against a never-terminating server."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.nn.functional as F

from distributed_ml_pytorch_b200.models import build_model
from distributed_ml_pytorch_b200.parallel import Asynchronous, ParameterServer, Synchronous
from distributed_ml_pytorch_b200.parallel.launcher import spawn
from distributed_ml_pytorch_b200.utils.serialization import ravel_model_params

pytestmark = pytest.mark.timeout(180)


def _batch(seed, n=8, c=1, hw=28):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(n, c, hw, hw, generator=g), torch.randint(0, 10, (n,), generator=g)


def _p2p(rank, size, out_dir):
    tensor = torch.zeros(1)
    if rank == 0:
        tensor += 1
        dist.send(tensor=tensor, dst=1)
    else:
        dist.recv(tensor=tensor, src=0)
    torch.save(tensor, os.path.join(out_dir, f"p2p{rank}.pt"))


def test_p2p_send_recv(tmp_path):
    spawn(_p2p, 2, args=(str(tmp_path),))
    assert torch.load(tmp_path / "p2p0.pt").item() == 1.0 and torch.load(tmp_path / "p2p1.pt").item() == 1.0


def _one_worker(rank, size, out_dir, steps):
    torch.manual_seed(0)
    net = build_model("lenet", in_channels=1, input_hw=28).eval()
    if rank == 0:
        server = ParameterServer(model=net)
        server.run()
        torch.save({"theta": server.parameter_shard, "counters": server.counters}, os.path.join(out_dir, "server.pt"))
        return
    # blocking_pull: every pull reply is applied before the next step, so the run is reproducible (replies otherwise land whenever
    # the listener thread receives them, and a late one legitimately rewinds the replica by a step)
    opt = Asynchronous(net.parameters(), lr=0.05, n_push=1, n_pull=1, model=net, blocking_pull=True)
    for i in range(steps):
        x, y = _batch(i)
        opt.zero_grad()
        F.cross_entropy(net(x), y).backward()
        opt.step()
    from distributed_ml_pytorch_b200.utils import messaging
    assert len(messaging._pending) <= messaging._MAX_INFLIGHT      # 20 work items were queued: the send queue stays bounded
    opt.finish()
    torch.save(ravel_model_params(net).clone(), os.path.join(out_dir, "worker.pt"))


def test_one_worker_push_pull_every_step_equals_sgd(tmp_path):
    steps = 6
    spawn(_one_worker, 2, args=(str(tmp_path), steps))
    srv = torch.load(tmp_path / "server.pt", weights_only=False)
    worker = torch.load(tmp_path / "worker.pt")
    torch.manual_seed(0)
    ref = build_model("lenet", in_channels=1, input_hw=28).eval()
    sgd = torch.optim.SGD(ref.parameters(), lr=0.05)
    for i in range(steps):
        x, y = _batch(i)
        sgd.zero_grad()
        F.cross_entropy(ref(x), y).backward()
        sgd.step()
    expect = ravel_model_params(ref)
    torch.testing.assert_close(srv["theta"], expect, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(worker, expect, rtol=1e-4, atol=1e-5)
    assert srv["counters"]["GradientUpdate"] == steps and srv["counters"]["ParameterRequest"] == steps
    assert srv["counters"]["ParameterUpdate"] == 1


def _ps_training(rank, size, out_dir, steps, n_push, n_pull):
    """BASELINE.json config 1: LeNet on MNIST-shaped synthetic data, 1 server + 2 workers, CPU/gloo."""
    torch.manual_seed(0)
    net = build_model("lenet", in_channels=1, input_hw=28)
    if rank == 0:
        server = ParameterServer(model=net, momentum=0.0)
        server.run()                      # returns: every worker sent WorkerDone
        server.save(os.path.join(out_dir, "server_ck.pt"))
        torch.save(server.counters, os.path.join(out_dir, "counters.pt"))
        return
    net.train()
    opt = Asynchronous(net.parameters(), lr=0.02, n_push=n_push, n_pull=n_pull, model=net)
    losses = []
    for i in range(steps):
        x, y = _batch(1000 * rank + i % 4, n=16)   # 4 recurring batches per worker: the loss must fall
        opt.zero_grad()
        loss = F.cross_entropy(net(x), y)
        loss.backward()
        opt.step()
        losses.append(loss.item())
    opt.finish()
    torch.save(losses, os.path.join(out_dir, f"losses{rank}.pt"))


def test_one_server_two_workers_lenet_mnist_shape(tmp_path):
    steps, n_push, n_pull = 40, 5, 5
    spawn(_ps_training, 3, args=(str(tmp_path), steps, n_push, n_pull))
    counters = torch.load(tmp_path / "counters.pt", weights_only=False)
    assert counters["GradientUpdate"] == 2 * (steps // n_push)
    assert counters["ParameterRequest"] == 2 * (steps // n_pull)
    for r in (1, 2):
        losses = torch.load(tmp_path / f"losses{r}.pt", weights_only=False)
        assert all(l == l for l in losses)
        assert sum(losses[-8:]) / 8 < sum(losses[:8]) / 8
    sd = torch.load(tmp_path / "server_ck.pt", weights_only=False)
    assert set(sd.keys()) == set(build_model("lenet", in_channels=1, input_hw=28).state_dict().keys())


def _sync(rank, size, out_dir, steps):
    torch.manual_seed(0)
    net = build_model("lenet", in_channels=1, input_hw=28).eval()
    opt = Synchronous(net.parameters(), lr=0.05, model=net, momentum=0.9)
    for i in range(steps):
        x, y = _batch(10 * i + rank)
        opt.zero_grad()
        F.cross_entropy(net(x), y).backward()
        opt.step()
    torch.save(ravel_model_params(net).clone(), os.path.join(out_dir, f"sync{rank}.pt"))


def test_synchronous_mode_equals_large_batch_sgd(tmp_path):
    steps = 4
    spawn(_sync, 2, args=(str(tmp_path), steps))
    a, b = torch.load(tmp_path / "sync0.pt"), torch.load(tmp_path / "sync1.pt")
    torch.testing.assert_close(a, b, rtol=0, atol=0)
    torch.manual_seed(0)
    ref = build_model("lenet", in_channels=1, input_hw=28).eval()
    sgd = torch.optim.SGD(ref.parameters(), lr=0.05, momentum=0.9)
    for i in range(steps):
        (x0, y0), (x1, y1) = _batch(10 * i), _batch(10 * i + 1)
        sgd.zero_grad()
        F.cross_entropy(ref(torch.cat([x0, x1])), torch.cat([y0, y1])).backward()
        sgd.step()
    torch.testing.assert_close(a, ravel_model_params(ref), rtol=1e-4, atol=1e-5)


def _sharded_servers(rank, size, out_dir, steps, n_servers):
    """NORTH-STAR sharded parameter server on the gloo transport: ranks 0..n_servers-1 each own one slice of the flat vector."""
    torch.manual_seed(0)
    net = build_model("lenet", in_channels=1, input_hw=28).eval()
    if rank < n_servers:
        server = ParameterServer(model=net, shard_index=rank, n_shards=n_servers)
        server.run()
        server.save(os.path.join(out_dir, "server.pt"))                     # -> server.pt.shard<rank>
        torch.save({"bounds": server.bounds, "counters": server.counters}, os.path.join(out_dir, f"srv{rank}.pt"))
        return
    opt = Asynchronous(net.parameters(), lr=0.05, n_push=1, n_pull=1, model=net, server_ranks=list(range(n_servers)), blocking_pull=True)
    for i in range(steps):
        x, y = _batch(i)
        opt.zero_grad()
        F.cross_entropy(net(x), y).backward()
        opt.step()
    opt.finish()
    torch.save(ravel_model_params(net).clone(), os.path.join(out_dir, "worker.pt"))


def test_two_server_shards_one_worker_equals_sgd(tmp_path):
    steps, n_servers = 5, 2
    spawn(_sharded_servers, n_servers + 1, args=(str(tmp_path), steps, n_servers))
    torch.manual_seed(0)
    ref = build_model("lenet", in_channels=1, input_hw=28).eval()
    sgd = torch.optim.SGD(ref.parameters(), lr=0.05)
    for i in range(steps):
        x, y = _batch(i)
        sgd.zero_grad()
        F.cross_entropy(ref(x), y).backward()
        sgd.step()
    expect = ravel_model_params(ref)
    torch.testing.assert_close(torch.load(tmp_path / "worker.pt"), expect, rtol=1e-4, atol=1e-5)
    covered = 0
    for r in range(n_servers):
        info = torch.load(tmp_path / f"srv{r}.pt", weights_only=False)
        part = torch.load(tmp_path / f"server.pt.shard{r}", weights_only=False)
        b, e = info["bounds"]
        covered += e - b
        torch.testing.assert_close(part["parameter_shard"], expect[b:e], rtol=1e-4, atol=1e-5)
        assert info["counters"]["GradientUpdate"] == steps and info["counters"]["ParameterRequest"] == steps
    assert covered == expect.numel()
    merged = ParameterServer.merge_shards([tmp_path / f"server.pt.shard{r}" for r in range(n_servers)],
                                          build_model("lenet", in_channels=1, input_hw=28))
    torch.testing.assert_close(merged["fc3.bias"], dict(ref.named_parameters())["fc3.bias"].detach(), rtol=1e-4, atol=1e-5)


def _dying_worker(rank, size, out_dir):
    torch.manual_seed(0)
    net = build_model("lenet", in_channels=1, input_hw=28)
    if rank == 0:
        server = ParameterServer(model=net, idle_timeout_s=3.0)
        try:
            server.run()
        except Exception as exc:
            torch.save({"type": type(exc).__name__, "msg": str(exc), "counters": server.counters}, os.path.join(out_dir, "server_exc.pt"))
        return
    opt = Asynchronous(net.parameters(), lr=0.05, n_push=1, n_pull=1, model=net)
    x, y = _batch(0)
    opt.zero_grad()
    F.cross_entropy(net(x), y).backward()
    opt.step()
    from distributed_ml_pytorch_b200.utils import messaging
    messaging.flush_sends()
    import time
    time.sleep(1.0)
    # the worker "crashes" here: no WorkerDone is ever sent (the reference's server would block forever) and the
    # process ends without tearing the process group down
    os._exit(0)


def test_server_reports_lost_workers_instead_of_hanging(tmp_path):
    spawn(_dying_worker, 2, args=(str(tmp_path),))
    exc = torch.load(tmp_path / "server_exc.pt", weights_only=False)
    assert exc["type"] == "WorkerLost" and "0 of 1 workers had finished" in exc["msg"]       # idle timeout or closed connection
    assert exc["counters"]["GradientUpdate"] == 1

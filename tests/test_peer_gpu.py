"""NVLink peer-memory kernels and the GPU transports of the optimizers.

World-size-1 cases run on a single GPU (the kernels treat the local GPU as its only peer); the multi-GPU cases spawn
one process per GPU over NCCL + CUDA IPC and are skipped when fewer than two GPUs are visible.
"""
import copy
import os

import pytest
import torch
import torch.nn.functional as F

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300)]

from distributed_ml_pytorch_b200.models import build_model
from distributed_ml_pytorch_b200.ops import functional as OF
from distributed_ml_pytorch_b200.parallel import Asynchronous, FusedSGD, Synchronous
from distributed_ml_pytorch_b200.parallel.launcher import spawn
from distributed_ml_pytorch_b200.utils.serialization import ravel_model_params


def _batch(seed, n=16, dev="cuda"):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(n, 3, 32, 32, generator=g).to(dev), torch.randint(0, 10, (n,), generator=g).to(dev)


def _steps(net, opt, n, seed0=0):
    losses = []
    for i in range(n):
        x, y = _batch(seed0 + i)
        opt.zero_grad()
        loss = OF.cross_entropy(net(x), y)
        loss.backward()
        opt.step()
        losses.append(loss.item())
    return losses


def test_async_peer_world1_equals_fused_sgd():
    torch.manual_seed(0)
    a = build_model("lenet").cuda().eval()
    b = copy.deepcopy(a)
    oa = Asynchronous(a.parameters(), lr=0.05, n_push=1, n_pull=1, model=a)
    ob = FusedSGD(b.parameters(), lr=0.05, model=b)
    la, lb = _steps(a, oa, 5), _steps(b, ob, 5)
    assert la == pytest.approx(lb, rel=1e-3, abs=1e-3)
    torch.testing.assert_close(ravel_model_params(a), ravel_model_params(b), rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(oa.server_parameters(), oa.arena.flat_params(), rtol=1e-5, atol=1e-6)
    oa.finish()


def test_async_peer_accumulates_between_pushes():
    torch.manual_seed(0)
    net = build_model("lenet").cuda().eval()
    opt = Asynchronous(net.parameters(), lr=0.05, n_push=3, n_pull=1000, model=net)
    theta0 = opt.server_parameters().clone()
    _steps(net, opt, 4)   # pushes at idx 0 and 3
    assert opt.pushes == 2
    torch.testing.assert_close(opt.server_parameters(), opt.arena.flat_params(), rtol=1e-5, atol=1e-6)
    assert not torch.equal(theta0, opt.server_parameters())
    assert opt.accumulated_gradients.abs().sum().item() == 0.0


def test_mailbox_momentum_and_staleness_world1():
    """K2: with one worker pushing every step the owner applies v = mu v + delta; theta += v (tau = 0)."""
    torch.manual_seed(0)
    net = build_model("lenet").cuda().eval()
    opt = Asynchronous(net.parameters(), lr=0.05, n_push=1, n_pull=1000, model=net, momentum=0.5, staleness="inverse")
    theta = opt.server_parameters().clone()
    v = torch.zeros_like(theta)
    for i in range(4):
        _steps(net, opt, 1, seed0=i)
        delta = opt.arena.flat_grads() * (-0.05)   # this step's update, arena order == server order (the replica itself also moves
                                                   # when the step-0 pull lands, so it is not measured on the replica)
        tau = max(0, i - 1)                  # step 0 pushes then pulls (version 1); afterwards the snapshot ages by one per push
        v = 0.5 * v + delta / (1.0 + tau)
        theta = theta + v
        torch.testing.assert_close(opt.server_parameters(), theta, rtol=1e-4, atol=1e-6)
    st = opt.staleness_stats()
    assert st["chunks_applied"] > 0 and st["max_staleness"] == 2
    opt.finish()


def test_sync_peer_world1_equals_fused_sgd_with_momentum():
    torch.manual_seed(0)
    a = build_model("lenet").cuda().eval()
    b = copy.deepcopy(a)
    oa = Synchronous(a.parameters(), lr=0.05, model=a, momentum=0.9, weight_decay=1e-3)
    ob = FusedSGD(b.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-3, model=b)
    la, lb = _steps(a, oa, 5), _steps(b, ob, 5)
    assert la == pytest.approx(lb, rel=1e-3, abs=1e-3)
    torch.testing.assert_close(ravel_model_params(a), ravel_model_params(b), rtol=1e-4, atol=1e-5)


def test_train_step_graph_matches_eager():
    from distributed_ml_pytorch_b200.engine import TrainStep
    torch.manual_seed(0)
    a = build_model("alexnet").cuda().eval()
    b = copy.deepcopy(a)
    oa = Asynchronous(a.parameters(), lr=0.01, n_push=2, n_pull=2, model=a)
    ob = Asynchronous(b.parameters(), lr=0.01, n_push=2, n_pull=2, model=b)
    ts = TrainStep(a, oa, 16, (3, 32, 32), use_graph=True)
    assert ts.graph is not None and ts.kernels_per_step > 10
    graph_losses = []
    for i in range(6):
        x, y = _batch(i)
        graph_losses.append(ts(x, y, sync=True).loss)
    eager_losses = _steps(b, ob, 6)
    assert graph_losses == pytest.approx(eager_losses, rel=2e-3, abs=2e-3)
    torch.testing.assert_close(ravel_model_params(a), ravel_model_params(b), rtol=1e-3, atol=1e-4)
    # pipelined (one-step-late) result delivery
    r0 = ts(*_batch(10))
    r1 = ts(*_batch(11))
    assert r1 is not None and r1.step == ts.step_count - 2
    assert ts.flush().step == ts.step_count - 1


# ------------------------------------------------------------------------------------------------
# multi-GPU
# ------------------------------------------------------------------------------------------------
def _need(n):
    if torch.cuda.device_count() < n:
        pytest.skip(f"needs {n} GPUs")


def _mgpu_sync(rank, size, out_dir, steps, nvls=None):
    dev = torch.device("cuda", rank)
    torch.manual_seed(0)
    net = build_model("lenet").to(dev).eval()
    opt = Synchronous(net.parameters(), lr=0.05, model=net, momentum=0.9, nvls=nvls)
    if rank == 0:
        with open(os.path.join(out_dir, "nvls.txt"), "w") as f:
            f.write(str(opt.nvls))
    for i in range(steps):
        x, y = _batch(10 * i + rank, dev=dev)
        opt.zero_grad()
        OF.cross_entropy(net(x), y).backward()
        opt.step()
    opt.finish()
    torch.save(ravel_model_params(net).cpu().clone(), os.path.join(out_dir, f"sync{rank}.pt"))


@pytest.mark.parametrize("world,nvls", [(2, True), (2, False), (4, None), (8, None), (8, False)])
def test_sync_multi_gpu_equals_large_batch_sgd(tmp_path, world, nvls):
    """nvls=True / None: NVSwitch multimem path (None: where it was measured faster, i.e. from 4 GPUs); nvls=False: peer loads / stores."""
    _need(world)
    steps = 4
    spawn(_mgpu_sync, world, backend="nccl", args=(str(tmp_path), steps, nvls))
    print("NVLS used:", open(tmp_path / "nvls.txt").read())
    outs = [torch.load(tmp_path / f"sync{r}.pt") for r in range(world)]
    for o in outs[1:]:
        assert torch.equal(o, outs[0])           # every replica holds bit-identical weights
    torch.manual_seed(0)
    ref = build_model("lenet").eval()
    sgd = torch.optim.SGD(ref.parameters(), lr=0.05, momentum=0.9)
    for i in range(steps):
        xs, ys = zip(*[_batch(10 * i + r, dev="cpu") for r in range(world)])
        sgd.zero_grad()
        F.cross_entropy(ref(torch.cat(xs)), torch.cat(ys)).backward()
        sgd.step()
    err = (outs[0] - ravel_model_params(ref)).norm() / ravel_model_params(ref).norm()
    assert err < 2e-3, err     # bf16 tensor-core forward/backward vs fp32 CPU


def _mgpu_async(rank, size, out_dir, steps, momentum):
    dev = torch.device("cuda", rank)
    torch.manual_seed(0)
    net = build_model("lenet").to(dev).eval()
    opt = Asynchronous(net.parameters(), lr=0.02, n_push=1, n_pull=1, model=net, momentum=momentum,
                       staleness="inverse" if momentum else None)
    theta0 = opt.server_parameters().cpu().clone()
    local_sum = torch.zeros_like(opt.arena.flat_params())
    for i in range(steps):
        x, y = _batch(100 * rank + i, dev=dev)
        opt.zero_grad()
        OF.cross_entropy(net(x), y).backward()
        opt._apply_pending()
        local_sum += opt.arena.flat_grads() * (-0.02)      # this worker's -lr*g contribution (the replica itself also moves when a
        opt.step_local()                                   # fetched pull lands)
        opt.communicate()
    opt.finish()
    final = opt.server_parameters().cpu()
    torch.save({"theta0": theta0, "sum": local_sum.cpu(), "final": final, "stats": opt.staleness_stats()}, os.path.join(out_dir, f"async{rank}.pt"))


@pytest.mark.parametrize("world", [2, 8])
def test_async_multi_gpu_server_sums_all_pushes(tmp_path, world):
    """Plain-add push (reference semantics): server == theta0 + sum over workers and steps of (-lr * g), whatever the interleaving."""
    _need(world)
    spawn(_mgpu_async, world, backend="nccl", args=(str(tmp_path), 5, 0.0))
    outs = [torch.load(tmp_path / f"async{r}.pt", weights_only=False) for r in range(world)]
    expect = outs[0]["theta0"] + sum(o["sum"] for o in outs)
    for o in outs:
        torch.testing.assert_close(o["final"], expect, rtol=1e-4, atol=1e-5)


def test_async_multi_gpu_mailbox_runs_and_tracks_staleness(tmp_path):
    _need(2)
    spawn(_mgpu_async, 2, backend="nccl", args=(str(tmp_path), 6, 0.5))
    outs = [torch.load(tmp_path / f"async{r}.pt", weights_only=False) for r in range(2)]
    assert torch.equal(outs[0]["final"], outs[1]["final"])
    assert not torch.equal(outs[0]["final"], outs[0]["theta0"])
    assert torch.isfinite(outs[0]["final"]).all()
    assert sum(o["stats"]["chunks_applied"] for o in outs) > 0


def _add_update(opt, delta):
    """A local update that has not been pushed: into the accumulator, or (overlap mode: no accumulator) into the replica."""
    if opt.overlap:
        opt.arena.params.add_(delta)
    else:
        opt.accumulated_gradients.add_(delta)


@pytest.mark.parametrize("overlap", [False, True])
def test_mailbox_protocol_stress_world1(overlap):
    """Flag / version protocol under load (the pool's compute-sanitizer is closed, so the race check is an invariant check):
    with scale 1 and no momentum the mailbox path must deliver every pushed update exactly once, whatever the interleaving."""
    torch.manual_seed(0)
    net = build_model("lenet").cuda().eval()
    opt = Asynchronous(net.parameters(), lr=0.05, n_push=1, n_pull=7, model=net, staleness="exp", staleness_param=1.0, overlap=overlap)
    assert opt._use_mailbox and opt.overlap == overlap
    theta0 = opt.server_parameters().clone()
    total = torch.zeros_like(theta0)
    g = torch.Generator(device="cuda").manual_seed(1)
    for i in range(200):
        delta = torch.randn(opt.arena.padded, device="cuda", generator=g) * 1e-3
        delta[opt.arena.numel:] = 0
        _add_update(opt, delta)
        total += delta[: opt.arena.numel]
        opt._push()
        if i % 3 != 0:                     # the owner does not serve every step: slots stay busy, updates keep accumulating
            opt._owner_apply()
        if i % 7 == 0:
            opt._pull()
    for _ in range(3):
        opt._owner_apply()
        opt._push()
    opt._owner_apply()
    torch.cuda.synchronize()
    assert opt.accumulated_gradients.abs().sum().item() == 0.0
    torch.testing.assert_close(opt.server_parameters(), theta0 + total, rtol=1e-4, atol=2e-5 if overlap else 1e-5)
    assert opt.staleness_stats()["chunks_applied"] > 0


def _mgpu_mailbox_stress(rank, size, out_dir, steps):
    dev = torch.device("cuda", rank)
    torch.manual_seed(0)
    net = build_model("lenet").to(dev).eval()
    opt = Asynchronous(net.parameters(), lr=0.02, n_push=1, n_pull=5, model=net, staleness="exp", staleness_param=1.0)
    theta0 = opt.server_parameters().cpu().clone()
    total = torch.zeros(opt.arena.numel, device=dev)
    g = torch.Generator(device=dev).manual_seed(100 + rank)
    for i in range(steps):
        delta = torch.randn(opt.arena.padded, device=dev, generator=g) * 1e-3
        delta[opt.arena.numel:] = 0
        _add_update(opt, delta)
        total += delta[: opt.arena.numel]
        opt._push()
        if (i + rank) % 2 == 0:
            opt._owner_apply()
        if i % 5 == 0:
            opt._pull()
    for _ in range(4):                      # drain: every rank serves, then re-pushes what was still busy
        opt.peer.barrier()
        opt._owner_apply()
        opt.peer.barrier()
        opt._push()
    opt.peer.barrier()
    opt._owner_apply()
    opt.peer.barrier()
    left = opt.accumulated_gradients.abs().sum().item()
    torch.save({"theta0": theta0, "sum": total.cpu(), "final": opt.server_parameters().cpu(), "left": left}, os.path.join(out_dir, f"stress{rank}.pt"))


@pytest.mark.parametrize("world", [2, 8])
def test_mailbox_protocol_stress_multi_gpu(tmp_path, world):
    _need(world)
    spawn(_mgpu_mailbox_stress, world, backend="nccl", args=(str(tmp_path), 60))
    outs = [torch.load(tmp_path / f"stress{r}.pt", weights_only=False) for r in range(world)]
    assert all(o["left"] == 0.0 for o in outs)
    expect = outs[0]["theta0"] + sum(o["sum"] for o in outs)
    for o in outs:
        torch.testing.assert_close(o["final"], expect, rtol=1e-4, atol=2e-5)


def test_train_step_graph_with_dropout_and_lr_schedule():
    """LeNet in train mode inside the CUDA graph: dropout masks must change from replay to replay (device-side counter) and a
    learning-rate change between replays must take effect (lr lives in device memory)."""
    from distributed_ml_pytorch_b200.engine import TrainStep
    torch.manual_seed(0)
    net = build_model("lenet").cuda().train()
    opt = FusedSGD(net.parameters(), lr=0.05, model=net)
    ts = TrainStep(net, opt, 32, (3, 32, 32))
    x, y = _batch(0, n=32)
    l0 = [ts(x, y, sync=True).loss for _ in range(6)]
    assert len(set(round(v, 6) for v in l0)) > 1 and all(v == v for v in l0)
    before = opt.arena.flat_params().clone()
    for g in opt.param_groups:
        g["lr"] = 0.0                       # a scheduler would do this between steps
    ts(x, y, sync=True)
    assert torch.equal(opt.arena.flat_params(), before)


def _bulk_vs_thread(dev, pg, world, rank, tiles_per_shard=5):
    """TMA bulk movers against the per-thread movers on the same inputs: push_add, push_delta, pull (+ bf16 mirror, version snapshot)."""
    from distributed_ml_pytorch_b200.parallel import peer
    chunk = peer.chunk_elems()
    shard = tiles_per_shard * chunk
    n = shard * world
    g = torch.Generator(device=dev).manual_seed(7 + rank)
    out = {}
    for name, mv in (("thread", 1), ("bulk", 2)):
        peer.set_comm_mover(mv)
        master = pg.symmetric(shard * 4)
        version = pg.symmetric(tiles_per_shard * 4)
        master.view(torch.float32, shard).copy_(torch.arange(shard, device=dev, dtype=torch.float32) * 1e-3 + rank)
        version.view(torch.int32, tiles_per_shard).copy_(torch.arange(tiles_per_shard, device=dev, dtype=torch.int32) + 10 * rank)
        g.manual_seed(7 + rank)
        acc = torch.randn(n, device=dev, generator=g)
        live = torch.randn(n, device=dev, generator=g)
        snap = live + torch.randn(n, device=dev, generator=g) * 1e-2
        live0 = live.clone()
        pg.barrier()
        peer.push_add(acc, master.table, world, shard, n)
        peer.push_delta(live, snap, master.table, world, shard, n)
        torch.cuda.synchronize(dev)
        pg.barrier()
        params = torch.empty(n, device=dev)
        shadow = torch.empty(n, device=dev, dtype=torch.bfloat16)
        pv = torch.zeros(world * tiles_per_shard, dtype=torch.int32, device=dev)
        peer.pull(params, master.table, world, shard, n, version.table, pv, shadow=shadow)
        torch.cuda.synchronize(dev)
        pg.barrier()
        assert acc.abs().sum().item() == 0.0 and torch.equal(snap, live) and torch.equal(live, live0)
        assert torch.equal(shadow.float(), params.to(torch.bfloat16).float())
        out[name] = (params.cpu(), pv.cpu())
        del master, version
    peer.set_comm_mover(0)
    torch.testing.assert_close(out["bulk"][0], out["thread"][0], rtol=1e-5, atol=1e-5)   # fp32 adds from several GPUs land in any order
    assert torch.equal(out["bulk"][1], out["thread"][1])
    expect_ver = torch.cat([torch.arange(tiles_per_shard, dtype=torch.int32) + 10 * r for r in range(world)])
    assert torch.equal(out["bulk"][1], expect_ver)


def test_bulk_movers_match_thread_movers_world1():
    from distributed_ml_pytorch_b200.parallel import peer
    dev = torch.device("cuda", 0)
    pg = peer.PeerGroup(device=dev)
    _bulk_vs_thread(dev, pg, 1, 0, tiles_per_shard=331)       # more tiles than CTAs, uneven runs, several ring wraps
    _bulk_vs_thread(dev, pg, 1, 0, tiles_per_shard=3)
    pg.close()


def _mgpu_bulk(rank, size, out_dir):
    from distributed_ml_pytorch_b200.parallel import peer
    dev = torch.device("cuda", rank)
    pg = peer.PeerGroup(device=dev)
    _bulk_vs_thread(dev, pg, size, rank, tiles_per_shard=37)
    pg.close()
    torch.save({"ok": True}, os.path.join(out_dir, f"bulk{rank}.pt"))


@pytest.mark.parametrize("world", [2, 8])
def test_bulk_movers_match_thread_movers_multi_gpu(tmp_path, world):
    _need(world)
    spawn(_mgpu_bulk, world, backend="nccl", args=(str(tmp_path),))
    assert all(os.path.exists(tmp_path / f"bulk{r}.pt") for r in range(world))


def _mgpu_momentum_converges(rank, size, out_dir, steps):
    """BASELINE config 4 in small: ResNet-18, asynchronous, server momentum 0.9 + inverse-staleness scaling, server_gain='auto'."""
    dev = torch.device("cuda", rank)
    torch.manual_seed(0)
    net = build_model("resnet18").to(dev).train()
    opt = Asynchronous(net.parameters(), lr=0.05, n_push=2, n_pull=2, model=net, momentum=0.9, staleness="inverse", server_gain="auto")
    g = torch.Generator().manual_seed(50 + rank)
    x = torch.randn(32, 3, 32, 32, generator=g).to(dev)          # one fixed batch per worker: the run has to memorise it
    y = torch.randint(0, 10, (32,), generator=g).to(dev)
    losses = []
    for _ in range(steps):
        opt.zero_grad()
        loss = OF.cross_entropy(net(x), y)
        loss.backward()
        opt.step()
        losses.append(loss.item())
    opt.finish()
    torch.save({"losses": losses, "stats": opt.staleness_stats(), "gain": opt.server_gain}, os.path.join(out_dir, f"conv{rank}.pt"))


@pytest.mark.parametrize("world", [2, 8])
def test_async_momentum_staleness_trains_multi_gpu(tmp_path, world):
    """Round 1: this configuration (several workers feeding one server momentum buffer) climbed to loss 10.  With the gain
    (1 - mu) / workers the loss has to fall on every worker."""
    _need(world)
    spawn(_mgpu_momentum_converges, world, backend="nccl", args=(str(tmp_path), 80))
    for r in range(world):
        o = torch.load(tmp_path / f"conv{r}.pt", weights_only=False)
        first, last = sum(o["losses"][:5]) / 5, sum(o["losses"][-5:]) / 5
        print(f"rank {r}: loss {first:.3f} -> {last:.3f}, gain {o['gain']:.4f}, staleness {o['stats']}")
        assert o["gain"] == pytest.approx(0.1 / world)
        assert all(l == l and l < 20 for l in o["losses"])
        assert last < 0.7 * first, (first, last)

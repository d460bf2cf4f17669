"""CPU unit tests: flat arena, shard map, serialization, staleness scale, state_dict layouts, local async == SGD."""
import copy

import pytest
import torch
import torch.nn.functional as F

from distributed_ml_pytorch_b200.models import AlexNet, LeNet, build_model
from distributed_ml_pytorch_b200.parallel import Asynchronous, DownpourSGD, FlatArena, FusedSGD, ShardMap, Synchronous, staleness_scale
from distributed_ml_pytorch_b200.utils.serialization import numel, ravel_model_params, unravel_model_params


def test_shard_map_partition():
    m = ShardMap(2472266, 8)
    assert m.shard_size % 256 == 0 and m.padded >= m.numel and m.padded == 8 * m.shard_size
    covered = 0
    for s in range(8):
        b, e = m.valid_bounds(s)
        covered += e - b
        assert m.owner_of(b) == s
    assert covered == m.numel
    assert ShardMap(10, 4).shard_size == 256
    with pytest.raises(IndexError):
        m.owner_of(m.padded)


def test_ravel_unravel_roundtrip_and_arena_views():
    net = LeNet()
    flat = ravel_model_params(net)
    assert flat.numel() == numel(net) == 62006
    ref = copy.deepcopy(net)
    unravel_model_params(net, flat * 2)
    for p, q in zip(net.parameters(), ref.parameters()):
        assert torch.equal(p, q * 2)
    arena = FlatArena(net, n_shards=4)
    assert arena.is_bound() and arena.padded % 4 == 0
    # ravel is now a zero-copy view of the arena, in model.parameters() order
    assert ravel_model_params(net).data_ptr() == arena.params.data_ptr()
    assert torch.equal(ravel_model_params(net), torch.cat([p.detach().reshape(-1) for p in net.parameters()]))
    # gradients land in the arena
    out = net(torch.randn(2, 3, 32, 32))
    out.sum().backward()
    assert arena.grads[: arena.numel].abs().sum() > 0
    assert torch.equal(ravel_model_params(net, grads=True), arena.grads[: arena.numel])
    unravel_model_params(net, torch.zeros(arena.numel))
    assert arena.params.abs().sum() == 0


def test_state_dict_layouts_match_reference_keys():
    le = {k: tuple(v.shape) for k, v in LeNet().state_dict().items()}
    assert le == {"conv1.weight": (6, 3, 5, 5), "conv1.bias": (6,), "conv2.weight": (16, 6, 5, 5), "conv2.bias": (16,),
                  "fc1.weight": (120, 400), "fc1.bias": (120,), "fc2.weight": (84, 120), "fc2.bias": (84,),
                  "fc3.weight": (10, 84), "fc3.bias": (10,)}
    al = {k: tuple(v.shape) for k, v in AlexNet().state_dict().items()}
    assert al["features.0.weight"] == (64, 3, 11, 11) and al["features.3.weight"] == (192, 64, 5, 5)
    assert al["features.6.weight"] == (384, 192, 3, 3) and al["features.8.weight"] == (256, 384, 3, 3)
    assert al["features.10.weight"] == (256, 256, 3, 3) and al["classifier.weight"] == (10, 256)
    assert len(al) == 12
    assert sum(p.numel() for p in AlexNet().parameters()) == 2472266


def test_lenet_mnist_shape():
    net = build_model("lenet", in_channels=1, input_hw=28)
    assert net(torch.randn(3, 1, 28, 28)).shape == (3, 10)


def test_staleness_scale():
    assert staleness_scale(None, 0.9, 5) == 1.0
    assert staleness_scale("inverse", 0.9, 0) == 1.0 and staleness_scale("inverse", 0.9, 3) == 0.25
    assert abs(staleness_scale("exp", 0.5, 3) - 0.125) < 1e-12


def _train(net, opt, steps=5, seed=0):
    g = torch.Generator().manual_seed(seed)
    losses = []
    for _ in range(steps):
        x = torch.randn(8, 3, 32, 32, generator=g)
        y = torch.randint(0, 10, (8,), generator=g)
        opt.zero_grad()
        loss = F.cross_entropy(net(x), y)
        loss.backward()
        opt.step()
        losses.append(loss.item())
    return losses


def test_single_worker_async_equals_plain_sgd():
    """One worker, push and pull every step: asynchronous SGD degenerates to SGD."""
    torch.manual_seed(0)
    a = LeNet().eval()
    b = copy.deepcopy(a)
    la = _train(a, Asynchronous(a.parameters(), lr=0.05, n_push=1, n_pull=1, model=a))
    lb = _train(b, torch.optim.SGD(b.parameters(), lr=0.05, momentum=0.0))
    assert la == pytest.approx(lb, rel=1e-5, abs=1e-6)
    for p, q in zip(a.parameters(), b.parameters()):
        torch.testing.assert_close(p, q, rtol=1e-5, atol=1e-6)


def test_async_accumulates_between_pushes():
    torch.manual_seed(0)
    net = LeNet().eval()
    opt = Asynchronous(net.parameters(), lr=0.05, n_push=3, n_pull=100, model=net)
    assert DownpourSGD is Asynchronous
    theta0 = opt.server_parameters().clone()
    _train(net, opt, steps=4)   # pushes at idx 0 and 3
    assert opt.pushes == 2 and opt.pulls == 1
    # server = theta0 + everything pushed so far; local replica additionally holds nothing unpushed after idx 3
    torch.testing.assert_close(opt.server_parameters(), ravel_model_params(net), rtol=1e-5, atol=1e-6)
    assert not torch.equal(theta0, opt.server_parameters())
    assert opt.accumulated_gradients.abs().sum() == 0


def test_fused_sgd_matches_torch_sgd_with_momentum():
    torch.manual_seed(1)
    a = LeNet().eval()
    b = copy.deepcopy(a)
    la = _train(a, FusedSGD(a.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-3, model=a))
    lb = _train(b, torch.optim.SGD(b.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-3))
    assert la == pytest.approx(lb, rel=1e-4, abs=1e-5)


def test_synchronous_single_process_is_sgd_and_lr_scheduler_works():
    torch.manual_seed(2)
    a = LeNet().eval()
    b = copy.deepcopy(a)
    oa = Synchronous(a.parameters(), lr=0.05, model=a)
    ob = torch.optim.SGD(b.parameters(), lr=0.05)
    sa = torch.optim.lr_scheduler.LambdaLR(oa, lr_lambda=lambda e: 1 / (e + 1))   # reference main.py:47-48
    sb = torch.optim.lr_scheduler.LambdaLR(ob, lr_lambda=lambda e: 1 / (e + 1))
    for _ in range(2):
        la, lb = _train(a, oa, steps=2), _train(b, ob, steps=2)
        sa.step(); sb.step()
        assert la == pytest.approx(lb, rel=1e-5, abs=1e-6)
    assert oa.param_groups[0]["lr"] == pytest.approx(0.05 / 3)


def test_checkpoint_roundtrip_with_reference_layout(tmp_path):
    from distributed_ml_pytorch_b200.utils.checkpoint import load_checkpoint, save_checkpoint
    net = AlexNet()
    opt = Asynchronous(net.parameters(), lr=0.01, n_push=2, n_pull=2, model=net)
    _train(net, opt, steps=3)
    path = tmp_path / "ck.pt"
    save_checkpoint(str(path), net, opt, epoch=1)
    sd = torch.load(path, weights_only=False)
    assert set(sd["model"].keys()) == set(AlexNet().state_dict().keys())
    net2 = AlexNet()
    opt2 = Asynchronous(net2.parameters(), lr=0.5, n_push=2, n_pull=2, model=net2)
    meta = load_checkpoint(str(path), net2, opt2)
    assert meta["epoch"] == 1 and opt2.idx == opt.idx
    for p, q in zip(net.parameters(), net2.parameters()):
        assert torch.equal(p, q)
    # a bare reference-style file (torch.save(model.state_dict())) loads too
    torch.save(net.state_dict(), tmp_path / "bare.pt")
    load_checkpoint(str(tmp_path / "bare.pt"), net2)


def test_moving_the_model_after_building_the_optimizer_is_reported():
    """The reference's order (optimizer first, then net.cuda(), main.py:46,53-54) cannot work with a flat arena: say so loudly."""
    net = LeNet()
    opt = FusedSGD(net.parameters(), lr=0.1, model=net)
    opt.zero_grad()                                  # fine while the parameters are still the arena views
    net.double()                                     # stands in for .cuda() on this CPU box: parameters are re-created
    net.float()
    opt.zero_grad()                                  # same device -> silently re-bound
    assert next(net.parameters()).data_ptr() == opt.arena.params.data_ptr()
    net.to("meta")
    with pytest.raises(RuntimeError, match="after the optimizer was created"):
        opt.zero_grad()


def test_parameter_server_message_semantics_without_a_network(monkeypatch):
    """Server contract: ParameterUpdate sets, GradientUpdate adds the already -lr-scaled payload,
    ParameterRequest replies with the current vector; plus the north-star momentum / staleness-scaled update."""
    from distributed_ml_pytorch_b200.parallel import server as srv
    from distributed_ml_pytorch_b200.utils.messaging import MessageCode
    sent = []
    monkeypatch.setattr(srv, "send_message", lambda code, payload=None, dst=0, group=None: sent.append((code, None if payload is None else payload.clone(), dst)))
    net = LeNet()
    n = numel(net)
    ps = srv.ParameterServer(net, n_workers=2)
    theta = torch.arange(n, dtype=torch.float32) / n
    ps.receive(1, MessageCode.ParameterUpdate, theta)
    ps.receive(1, MessageCode.ParameterRequest, torch.zeros(0))
    assert sent[-1][0] == MessageCode.ParameterUpdate and sent[-1][2] == 1 and torch.equal(sent[-1][1], theta)
    delta = torch.full((n,), -0.5)
    ps.receive(2, MessageCode.GradientUpdate, delta)
    assert torch.allclose(ps.parameter_shard, theta - 0.5)
    ps.receive(1, MessageCode.WorkerDone, torch.zeros(0))
    ps.receive(2, MessageCode.WorkerDone, torch.zeros(0))
    assert ps.done_workers == {1, 2} and ps.counters == {"ParameterUpdate": 1, "GradientUpdate": 1, "ParameterRequest": 1}
    # momentum 0.5 + inverse staleness: worker 1 pulled at version 0, two foreign updates later its push is scaled by 1/3
    ps2 = srv.ParameterServer(net, n_workers=2, momentum=0.5, staleness="inverse")
    ps2.receive(1, MessageCode.ParameterUpdate, torch.zeros(n))
    ps2.receive(1, MessageCode.ParameterRequest, torch.zeros(0))
    one = torch.ones(n)
    ps2.receive(2, MessageCode.GradientUpdate, one)       # worker 2 never pulled -> tau = 0: v = 1, theta = 1
    ps2.receive(2, MessageCode.GradientUpdate, one)       # v = 0.5 + 1 = 1.5, theta = 2.5
    ps2.receive(1, MessageCode.GradientUpdate, one)       # tau = 2: v = 0.75 + 1/3, theta = 2.5 + 0.75 + 1/3
    assert torch.allclose(ps2.parameter_shard, torch.full((n,), 2.5 + 0.75 + 1 / 3))
    assert ps2.staleness_hist == {0: 2, 2: 1}
    snap = ps2.snapshot()
    assert set(snap.keys()) == set(LeNet().state_dict().keys())


def test_server_gain_keeps_momentum_server_stable_with_many_workers():
    """Round 1's ResNet-50 run with server momentum 0.9 and 8 workers climbed to loss 10: W workers feeding one momentum buffer step
    W * lr / (1 - mu) per unit of gradient.  On f = |theta|^2 / 2 with lr = 0.1 that is 8 (diverges); gain='auto' brings it back to 0.1."""
    from distributed_ml_pytorch_b200.parallel.async_optim import resolve_server_gain
    from distributed_ml_pytorch_b200.parallel.server import ParameterServer
    from distributed_ml_pytorch_b200.utils.messaging import MessageCode
    assert resolve_server_gain(None, 0.9, 8) == 1.0
    assert resolve_server_gain("auto", 0.9, 8) == pytest.approx(0.0125)
    assert resolve_server_gain("auto", 0.0, 8) == 1.0          # no momentum: the reference's plain sum of updates
    assert resolve_server_gain(0.5, 0.9, 8) == 0.5

    def run(gain):
        net = torch.nn.Linear(4, 1, bias=False)
        with torch.no_grad():
            net.weight.fill_(1.0)
        ps = ParameterServer(net, n_workers=8, momentum=0.9, gain=gain)
        for w in range(1, 9):
            ps.pulled_version[w] = 0
        for _ in range(60):
            theta = ps.parameter_shard.clone()
            for w in range(1, 9):
                ps.receive(w, MessageCode.GradientUpdate, -0.1 * theta)        # every worker's gradient of |theta|^2 / 2 at the pulled point
                ps.pulled_version[w] = ps.version
        return float(ps.parameter_shard.norm())

    assert run("auto") < 0.05                # converging: 2.0 -> ~1e-3 in 60 rounds
    assert not run(None) < 1e3               # the classic update with 8 workers blows up

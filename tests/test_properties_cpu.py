"""Property tests (hypothesis) for the host-side bookkeeping that every transport relies on: the shard cut, the flat
(de)serialisation the reference calls ravel / unravel (Asynchronous.py:18,27,34,54) and the staleness scale."""
import torch
from hypothesis import given, settings, strategies as st

from distributed_ml_pytorch_b200.parallel.arena import ShardMap
from distributed_ml_pytorch_b200.parallel.async_optim import staleness_scale
from distributed_ml_pytorch_b200.utils.serialization import ravel_model_params, unravel_model_params


@settings(max_examples=200, deadline=None)
@given(numel=st.integers(1, 5_000_000), n_shards=st.integers(1, 16), align=st.sampled_from([1, 4, 256, 4096]))
def test_shard_map_is_a_partition(numel, n_shards, align):
    sm = ShardMap(numel, n_shards, align=align)
    assert sm.shard_size % align == 0 and sm.padded >= numel and sm.padded == sm.shard_size * n_shards
    assert sm.padded - numel < n_shards * align + n_shards          # padding is bounded by the rounding, not by the size
    prev_end, covered = 0, 0
    for s in range(n_shards):
        b, e = sm.bounds(s)
        vb, ve = sm.valid_bounds(s)
        assert b == prev_end and e - b == sm.shard_size and vb == min(b, numel) and ve == min(e, numel) and vb <= ve   # empty past the end
        covered += ve - vb
        prev_end = e
    assert covered == numel and prev_end == sm.padded
    for idx in {0, numel - 1, numel // 2}:
        o = sm.owner_of(idx)
        assert sm.bounds(o)[0] <= idx < sm.bounds(o)[1]


@settings(max_examples=40, deadline=None)
@given(shapes=st.lists(st.lists(st.integers(1, 5), min_size=1, max_size=3), min_size=1, max_size=5), seed=st.integers(0, 1000))
def test_ravel_unravel_roundtrip_random_models(shapes, seed):
    torch.manual_seed(seed)

    class M(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.ps = torch.nn.ParameterList([torch.nn.Parameter(torch.randn(*s)) for s in shapes])

    m = M()
    flat = ravel_model_params(m)
    assert flat.numel() == sum(p.numel() for p in m.parameters())
    assert torch.equal(flat, torch.cat([p.detach().reshape(-1) for p in m.parameters()]))     # model.parameters() order, row-major
    new = torch.randn_like(flat)
    unravel_model_params(m, new)
    assert torch.equal(ravel_model_params(m), new)
    for p in m.parameters():
        p.grad = torch.randn_like(p)
    g = ravel_model_params(m, grads=True)
    assert torch.equal(g, torch.cat([p.grad.reshape(-1) for p in m.parameters()]))


@given(tau=st.integers(0, 1000), param=st.floats(0.01, 0.999))
def test_staleness_scale_is_a_decreasing_weight(tau, param):
    for mode in ("inverse", "exp"):
        s0, s1 = staleness_scale(mode, param, tau), staleness_scale(mode, param, tau + 1)
        assert 0.0 <= s1 <= s0 <= 1.0
    assert staleness_scale(None, param, tau) == 1.0 and staleness_scale("none", param, tau) == 1.0
    assert staleness_scale("inverse", param, 0) == 1.0 and staleness_scale("exp", param, 0) == 1.0

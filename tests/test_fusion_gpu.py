"""Launch-count reductions of the training step must not change its numbers.

K8: the split-K data-gradient finish of a conv block also runs the previous block's pool / ReLU backward and bias gradient
(`dmlb_splitk_finish_actbwd`), replacing a finish kernel + an activation-backward kernel per layer boundary.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

from distributed_ml_pytorch_b200 import _native
from distributed_ml_pytorch_b200.models import build_model
from distributed_ml_pytorch_b200.ops import functional as OF
from distributed_ml_pytorch_b200.parallel.fused_sgd import FusedSGD


def _grads(dtype, fused, seed=0):
    torch.manual_seed(seed)
    net = build_model("alexnet", dtype=dtype).cuda()
    opt = FusedSGD(net.parameters(), lr=0.0, model=net)
    torch.manual_seed(seed + 1)
    x = torch.randn(64, 3, 32, 32, device="cuda")
    y = torch.randint(0, 10, (64,), device="cuda")
    old = OF.set_fused_act_bwd(fused)
    try:
        net(x)                                     # warm-up: shadows, workspaces
        opt.zero_grad()
        torch.cuda.synchronize()
        n0 = _native.launch_count()
        loss = OF.cross_entropy(net(x), y)
        loss.backward()
        torch.cuda.synchronize()
        launches = _native.launch_count() - n0
    finally:
        OF.set_fused_act_bwd(old)
    return net._dmlb_arena.grads.clone(), float(loss), launches


@pytest.mark.parametrize("dtype", ["tf32", "bf16"])
def test_fused_dgrad_finish_matches_two_kernel_form(dtype):
    g0, l0, n0 = _grads(dtype, fused=False)
    g1, l1, n1 = _grads(dtype, fused=True)
    assert abs(l0 - l1) < 1e-5 * max(1.0, abs(l0))      # (split-K partial sums arrive in a different order from run to run)
    assert n1 == n0 - 4, (n0, n1)                  # four layer boundaries of AlexNet: (finish + act_bwd) -> one kernel each
    err = ((g0 - g1).norm() / g0.norm()).item()
    # identical products; only the bias gradients' summation differs (fp32 values before vs after the bf16 rounding of dy)
    assert err < 2e-2, err        # run-to-run split-K summation order already moves a few ReLU decisions
    assert torch.isfinite(g1).all() and g1.abs().sum() > 0


@pytest.mark.parametrize("model", ["lenet", "alexnet"])
@pytest.mark.parametrize("dtype", ["tf32", "bf16"])
def test_optimizer_kernel_refreshes_table_shadows(model, dtype):
    """The SGD kernel scatters every updated master value into the table-backed weight shadows (padded rows, permuted fc1, tap-padded
    conv1): after two steps they must equal what the stand-alone refresh kernel builds from the masters."""
    from distributed_ml_pytorch_b200.ops import kernels as K
    torch.manual_seed(0)
    net = build_model(model, dtype=dtype).cuda()
    opt = FusedSGD(net.parameters(), lr=0.05, model=net)
    x = torch.randn(16, 3, 32, 32, device="cuda")
    y = torch.randint(0, 10, (16,), device="cuda")
    n0 = None
    for _ in range(2):
        opt.zero_grad()
        OF.cross_entropy(net(x), y).backward()
        n0 = _native.launch_count()
        opt.step()
        assert _native.launch_count() - n0 == 1            # one kernel: update + accumulate + shadows (no refresh launch)
    torch.cuda.synchronize()
    ss = net._dmlb_shadows
    mods = [m for m in ss.modules if not m._flat]
    assert mods
    got = [m._shadow.clone() for m in mods]
    K.refresh_shadows(ss._table, ss._n_table, dt=mods[0].compute_dtype)
    torch.cuda.synchronize()
    for m, g in zip(mods, got):
        assert torch.equal(m._shadow, g), (type(m).__name__, tuple(m.weight.shape))


@pytest.mark.parametrize("dtype", ["tf32", "bf16"])
def test_cluster_splitk_matches_workspace_form(dtype):
    """Cluster split-K (DSMEM reduction + fused epilogue) against split-K through the fp32 workspace + finishing kernels: same gradients,
    eight launches fewer per AlexNet step (4 forward finishes, 4 backward finishes)."""
    from distributed_ml_pytorch_b200.ops import kernels as K

    def run(cluster):
        import os
        old = K.set_cluster_splitk(cluster)
        os.environ["DMLB_CLUSTER_MAXKB"] = "1000"           # force the cluster form on every split layer, whatever its loop length
        try:
            return _grads(dtype, fused=True)
        finally:
            K.set_cluster_splitk(old)
            os.environ.pop("DMLB_CLUSTER_MAXKB", None)

    g0, l0, n0 = run(False)
    g1, l1, n1 = run(True)
    assert abs(l0 - l1) < 1e-4 * max(1.0, abs(l0))
    assert n1 == n0 - 8, (n0, n1)
    err = ((g0 - g1).norm() / g0.norm()).item()
    assert err < 2e-2, err
    assert torch.isfinite(g1).all() and g1.abs().sum() > 0

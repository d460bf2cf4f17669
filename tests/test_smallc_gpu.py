"""Small-channel first-layer convolutions as implicit GEMMs (the reference's conv1 layers, example/models.py:8,29).

Cin <= 4: the input becomes one 16-byte granule per pixel, a k-block gathers eight filter taps by TMA im2col, no column matrix exists;
where the output width allows it the 2x2 max-pool runs in the GEMM epilogue.  Oracle: torch.nn.functional in fp32."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from distributed_ml_pytorch_b200 import _native
from distributed_ml_pytorch_b200 import nn as dnn
from distributed_ml_pytorch_b200.ops import functional as OF
from distributed_ml_pytorch_b200.parallel.fused_sgd import FusedSGD


def rel_err(a, b):
    return ((a.float() - b.float()).norm() / (b.float().norm() + 1e-12)).item()


def rnd(x, dtype):
    if dtype == "bf16":
        return x.to(torch.bfloat16).float()
    i = x.contiguous().view(torch.int32)
    return ((i + 0x1000) & ~0x1FFF).view(torch.float32)          # nearest tf32


CASES = [
    # cin, cout, k, s, p, hw, pool
    (3, 64, 11, 4, 5, 32, (2, 2, 0)),      # AlexNet conv1: 121 taps, pool fused in the epilogue (8x8 -> 4x4)
    (3, 6, 5, 1, 0, 32, (2, 2, 0)),        # LeNet conv1: 28x28 output, pool as its own kernel
    (1, 6, 5, 1, 0, 28, (2, 2, 0)),        # LeNet on the MNIST shape
    (3, 64, 7, 2, 3, 64, None),            # ResNet stem
    (3, 64, 3, 1, 1, 16, None),            # ResNet CIFAR stem: 16x16 outputs -> row windows
    (3, 64, 3, 1, 1, 32, (2, 2, 0)),       # 32x32 outputs: row windows, pool as its own kernel (width 32 > 16)
    (4, 32, 3, 2, 1, 17, None),            # full granule (fp32), odd extent
]


@pytest.mark.parametrize("dtype", ["tf32", "bf16"])
@pytest.mark.parametrize("cin,cout,k,s,p,hw,pool", CASES)
def test_smallc_conv_matches_fp32(dtype, cin, cout, k, s, p, hw, pool):
    torch.manual_seed(7)
    dev = "cuda"
    B = 16
    conv = dnn.set_compute_dtype(dnn.Conv2d(cin, cout, k, stride=s, padding=p, activation="relu", pool=pool), dtype).to(dev)
    assert conv._smallc()
    with torch.no_grad():
        conv.weight.copy_(rnd(conv.weight, dtype))
    x = rnd(torch.randn(B, cin, hw, hw, device=dev), dtype)
    y = conv(x)
    w = conv.weight.detach().clone().requires_grad_(True)
    b = conv.bias.detach().clone().requires_grad_(True)
    with torch.backends.cudnn.flags(allow_tf32=False):
        z = F.relu(F.conv2d(x, w, b, stride=s, padding=p))
        if dtype == "bf16":
            z = z + (rnd(z.detach(), dtype) - z.detach())         # the kernel stores the activation in bf16 (straight-through for the oracle)
        yr = z if pool is None else F.max_pool2d(z, *pool)
        assert y.shape == yr.shape
        assert rel_err(y, yr) < (1e-5 if dtype == "tf32" else 4e-3)
        g = rnd(torch.randn_like(yr), dtype)
        gin = OF.new_act4(*g.shape, g.device, y.dtype)
        gin.copy_(g)
        y.backward(gin)
        yr.backward(g)
    tol = 2e-3 if dtype == "tf32" else 2e-2
    assert rel_err(conv.weight.grad, w.grad) < tol          # PyTorch-layout gradient (no arena): pad channels dropped, (c,kh,kw) order
    assert rel_err(conv.bias.grad, b.grad) < tol


@pytest.mark.parametrize("dtype", ["tf32", "bf16"])
def test_alexnet_conv1_launches_and_arena_gradient(dtype):
    """Arena-backed (channels-last) weight gradient, and the launch budget of the forward pass (no im2col kernel, no column matrix)."""
    from distributed_ml_pytorch_b200.models import build_model
    torch.manual_seed(3)
    net = build_model("alexnet", dtype=dtype).cuda()
    FusedSGD(net.parameters(), lr=0.0, model=net)
    conv1 = net.features[0]
    x = torch.randn(64, 3, 32, 32, device="cuda")
    conv1(x)
    torch.cuda.synchronize()
    n0 = _native.launch_count()
    out = conv1(x)
    assert _native.launch_count() - n0 == 3          # layout kernel, split-K row-window GEMM, finish (+ pool); no column matrix
    ref = F.max_pool2d(F.relu(F.conv2d(x, conv1.weight, conv1.bias, stride=4, padding=5)), 2, 2)
    assert rel_err(out, ref) < (2e-3 if dtype == "tf32" else 1e-2)
    g = torch.randn_like(ref)
    gin = OF.new_act4(*g.shape, g.device, out.dtype)
    gin.copy_(g)
    conv1.weight.grad.zero_()
    out.backward(gin)
    w = conv1.weight.detach().clone().requires_grad_(True)
    F.max_pool2d(F.relu(F.conv2d(x, w, conv1.bias.detach(), stride=4, padding=5)), 2, 2).backward(g)
    assert rel_err(conv1.weight.grad, w.grad) < (4e-2 if dtype == "tf32" else 6e-2)       # unrounded operands: ReLU decisions may flip


def test_train_step_static_input_is_smallc_layout():
    from distributed_ml_pytorch_b200.engine import TrainStep
    from distributed_ml_pytorch_b200.models import build_model
    torch.manual_seed(0)
    net = build_model("alexnet").cuda()
    opt = FusedSGD(net.parameters(), lr=0.01, model=net)
    ts = TrainStep(net, opt, 16, (3, 32, 32))
    assert ts._smallc_in and getattr(ts.static_x, "_dmlb_rowwin", None) is not None      # 8x8 outputs: the row-window form
    assert ts.static_x._dmlb_rowwin.buf.shape == (16, 42, 44, 4)
    x = torch.randn(16, 3, 32, 32, device="cuda")
    y = torch.randint(0, 10, (16,), device="cuda")
    r = ts(x, y, sync=True)
    assert torch.equal(ts.static_x.contiguous(), x)           # the layout kernel is exact in fp32
    assert r.loss == r.loss and r.loss < 10

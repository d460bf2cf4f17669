"""Which convolution layers still go through an explicit column matrix (im2col / col2im launches) in a training step.

The implicit forms (TMA im2col, small-channel granule / row-window, 3-D weight map dgrad) cover every layer of AlexNet; what is left is
pinned here so that a regression (a layer silently falling back to the explicit path) fails a test:
  * LeNet conv2 (Cin = 6: more than one 16-byte granule of fp32, fewer channels than a k-block) — explicit wgrad column matrix and dgrad;
  * the stride-2 layers of the ResNets (three 3x3 s2 + three 1x1 s2 downsample convolutions): dgrad = GEMM + col2im.
"""
import pytest
import torch

from distributed_ml_pytorch_b200.models import build_model
from distributed_ml_pytorch_b200.ops import functional as OF
from distributed_ml_pytorch_b200.ops import kernels as K
from distributed_ml_pytorch_b200.parallel import FusedSGD

pytestmark = pytest.mark.gpu

_NAMES = ("im2col_nhwc", "im2col_nchw_f32", "col2im_nhwc")


def _explicit_launches(monkeypatch, model, hw, dtype, classes=10, batch=16):
    counts = {}
    for n in _NAMES:
        orig = getattr(K, n)

        def counted(*a, _n=n, _f=orig, **k):
            counts[_n] = counts.get(_n, 0) + 1
            return _f(*a, **k)
        monkeypatch.setattr(K, n, counted)
    torch.manual_seed(0)
    net = build_model(model, dtype=dtype, input_hw=hw, num_classes=classes).cuda().train()
    opt = FusedSGD(net.parameters(), lr=0.01, model=net)
    x = torch.randn(batch, 3, hw, hw, device="cuda")
    y = torch.randint(0, classes, (batch,), device="cuda")
    opt.zero_grad()
    loss = OF.cross_entropy(net(x), y)
    loss.backward()
    opt.step()
    torch.cuda.synchronize()
    assert torch.isfinite(loss)
    return counts


@pytest.mark.parametrize("dtype", ["tf32", "bf16"])
def test_alexnet_step_has_no_explicit_column_matrix(monkeypatch, dtype):
    assert _explicit_launches(monkeypatch, "alexnet", 32, dtype) == {}


@pytest.mark.parametrize("dtype", ["tf32", "bf16"])
def test_lenet_only_conv2_is_explicit(monkeypatch, dtype):
    c = _explicit_launches(monkeypatch, "lenet", 32, dtype)
    assert c.get("im2col_nchw_f32", 0) == 0                     # conv1 reads the NCHW batch through the small-channel form
    assert c.get("col2im_nhwc", 0) <= 1 and c.get("im2col_nhwc", 0) <= 2, c


@pytest.mark.parametrize("model,hw", [("resnet18", 32), ("resnet50", 64)])
def test_resnet_only_strided_dgrad_is_explicit(monkeypatch, model, hw):
    c = _explicit_launches(monkeypatch, model, hw, "bf16")
    assert c.get("im2col_nhwc", 0) == 0 and c.get("im2col_nchw_f32", 0) == 0, c
    assert c.get("col2im_nhwc", 0) == 6, c                      # 3 stages x (3x3 s2 + 1x1 s2 downsample)

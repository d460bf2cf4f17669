"""Implicit-GEMM convolution (TMA im2col gather, no column matrix) against torch fp32 conv on the same bf16-rounded data."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from distributed_ml_pytorch_b200.ops import functional as OF
from distributed_ml_pytorch_b200.ops import kernels as K


def bf(x):
    return x.to(torch.bfloat16).float()


def rel_err(a, b):
    return ((a.float() - b.float()).norm() / (b.float().norm() + 1e-12)).item()


GEOMS = [
    # B, Cin, Cout, H, W, k, stride, pad
    (8, 64, 192, 4, 4, 5, 1, 2),      # AlexNet conv2 (batch 8)
    (64, 192, 384, 2, 2, 3, 1, 1),    # AlexNet conv3: tensor < 128 KiB (descriptor workaround path)
    (4, 64, 64, 14, 14, 3, 1, 1),     # ResNet 3x3, tile crosses image rows and images
    (3, 128, 64, 9, 7, 3, 2, 1),      # strided, odd extents, ragged last tile
    (2, 64, 128, 8, 8, 1, 2, 0),      # 1x1 stride-2 downsample
]


@pytest.mark.parametrize("B,Cin,Cout,H,W,k,s,p", GEOMS)
def test_implicit_fprop_wgrad_dgrad(B, Cin, Cout, H, W, k, s, p):
    torch.manual_seed(0)
    dev = "cuda"
    x32 = bf(torch.randn(B, Cin, H, W, device=dev))
    w32 = bf(torch.randn(Cout, Cin, k, k, device=dev) * 0.05)
    bias = torch.randn(Cout, device=dev)
    x, ldx = OF.as_act4(x32.to(torch.bfloat16))
    wsh = w32.permute(0, 2, 3, 1).reshape(Cout, -1).to(torch.bfloat16).contiguous()      # [Cout, (kh,kw,cin)]
    Kd = k * k * Cin
    Ho, Wo = (H + 2 * p - k) // s + 1, (W + 2 * p - k) // s + 1
    M = B * Ho * Wo
    # ---- fprop (+bias +relu, bf16 out) ----
    y = OF.new_act4(B, Cout, Ho, Wo, dev)
    K.conv_implicit(0, x.data_ptr(), ldx, B, H, W, Cin, wsh.data_ptr(), Kd, k, k, s, s, p, p, Cout, y, K.pad8(Cout), bias=bias, relu=True)
    xr = x32.clone().requires_grad_(True)
    wr = w32.clone().requires_grad_(True)
    yr = F.relu(F.conv2d(xr, wr, bias, stride=s, padding=p))
    torch.cuda.synchronize()
    assert rel_err(y, yr) < 1e-2
    # ---- fprop split-K into fp32 (the path the small layers take) ----
    ws = torch.zeros(M, Cout, device=dev)
    K.conv_implicit(0, x.data_ptr(), ldx, B, H, W, Cin, wsh.data_ptr(), Kd, k, k, s, s, p, p, Cout, ws, Cout, split_k=3, atomic=True)
    assert rel_err(ws.view(B, Ho, Wo, Cout).permute(0, 3, 1, 2), F.conv2d(x32, w32, None, stride=s, padding=p)) < 1e-2
    # ---- backward reference ----
    g32 = bf(torch.randn_like(yr))
    F.conv2d(xr, wr, None, stride=s, padding=p).backward(g32)
    dy, lddy = OF.as_act4(g32.to(torch.bfloat16))
    # ---- wgrad: [Cout, (kh,kw,cin)] fp32, accumulated ----
    gw = torch.zeros(Cout, Kd, device=dev)
    K.conv_implicit(1, x.data_ptr(), ldx, B, H, W, Cin, dy.data_ptr(), lddy, k, k, s, s, p, p, Cout, gw, Kd, split_k=2, atomic=True)
    gw_ref = wr.grad.permute(0, 2, 3, 1).reshape(Cout, Kd)
    assert rel_err(gw, gw_ref) < 1e-2
    # ---- dgrad (stride 1, Cout % 64 == 0) ----
    if s == 1 and Cout % 64 == 0:
        dx = OF.new_act4(B, Cin, H, W, dev)
        K.conv_implicit(2, dy.data_ptr(), lddy, B, Ho, Wo, Cout, wsh.data_ptr(), Kd, k, k, 1, 1, p, p, Cin, dx, K.pad8(Cin))
        torch.cuda.synchronize()
        assert rel_err(dx, xr.grad) < 1e-2

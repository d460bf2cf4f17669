"""The tf32 compute dtype (fp32 tensors, tcgen05 kind::tf32) against plain PyTorch fp32 references of the same ops.

This is the reference's own precision — fp32 tensors end to end, cuDNN TF32 convolutions (reference example/main.py:41,53-54,70-73) —
so the tolerance is the one a TF32 tensor-core product allows (<= 2e-3 relative), far below the bf16 path's 1e-2.
"""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from distributed_ml_pytorch_b200 import nn as dnn
from distributed_ml_pytorch_b200.ops import functional as OF
from distributed_ml_pytorch_b200.ops import kernels as K

TOL = 2e-3


def rel_err(a, b):
    return ((a.float() - b.float()).norm() / (b.float().norm() + 1e-12)).item()


def tf32_round(x):
    """Round to the nearest tf32 value (10-bit mantissa), so a product of such values is exact in fp32."""
    i = x.contiguous().view(torch.int32)
    i = (i + 0x1000) & ~0x1FFF
    return i.view(torch.float32)


def act4(x):
    t = OF.new_act4(*x.shape, x.device, torch.float32)
    t.copy_(x)
    return t


def act2(x):
    t = OF.new_act2(x.shape[0], x.shape[1], x.device, torch.float32)
    t.copy_(x)
    return t


@pytest.mark.parametrize("M,N,Kd,amn,bmn", [(256, 384, 1728, 0, 0), (64, 10, 256, 0, 0), (100, 72, 40, 0, 0), (256, 1728, 384, 0, 1),
                                             (64, 256, 12, 0, 1), (384, 1728, 256, 1, 1), (12, 256, 64, 1, 1), (192, 1600, 1024, 1, 1)])
def test_gemm_tf32_all_majors(M, N, Kd, amn, bmn):
    torch.manual_seed(0)
    dev = "cuda"
    A = tf32_round(torch.randn(M, Kd, device=dev))
    B = tf32_round(torch.randn(N, Kd, device=dev))
    ref = (A.double() @ B.double().t()).float()
    a = (A.t() if amn else A).contiguous()
    b = (B.t() if bmn else B).contiguous()
    out = torch.zeros(M, N, device=dev, dtype=torch.float32)
    K.gemm(a, a.stride(0), bool(amn), b, b.stride(0), bool(bmn), M, N, Kd, out, N)
    torch.cuda.synchronize()
    assert rel_err(out, ref) < 1e-5          # tf32-exact operands: only the fp32 accumulation order differs
    # unrounded fp32 operands: the copy engine / tensor core reduce them to tf32
    A2, B2 = torch.randn(M, Kd, device=dev), torch.randn(N, Kd, device=dev)
    a2 = (A2.t() if amn else A2).contiguous()
    b2 = (B2.t() if bmn else B2).contiguous()
    out2 = torch.zeros(M, N, device=dev, dtype=torch.float32)
    K.gemm(a2, a2.stride(0), bool(amn), b2, b2.stride(0), bool(bmn), M, N, Kd, out2, N, split_k=2 if Kd >= 256 else 1)
    assert rel_err(out2, (A2.double() @ B2.double().t()).float()) < TOL


def test_gemm_tf32_bias_relu_epilogue():
    torch.manual_seed(1)
    dev = "cuda"
    M, N, Kd = 200, 84, 120
    A, B, bias = torch.randn(M, Kd, device=dev), torch.randn(N, Kd, device=dev), torch.randn(N, device=dev)
    out = torch.empty(M, N, device=dev)
    K.gemm(A, Kd, False, B, Kd, False, M, N, Kd, out, N, bias=bias, relu=True)
    assert rel_err(out, F.relu(A @ B.t() + bias)) < TOL


@pytest.mark.parametrize("cin,cout,k,s,p,hw,relu,pool", [
    (64, 192, 5, 1, 2, 4, True, (2, 2, 0)),       # AlexNet conv2: implicit fprop / wgrad / dgrad
    (192, 384, 3, 1, 1, 2, True, None),           # AlexNet conv3: split-K + finish
    (32, 64, 3, 1, 1, 8, True, None),             # smallest implicit channel count of the tf32 path
    (16, 32, 3, 2, 1, 9, False, None),            # explicit im2col / col2im, strided
    (6, 16, 5, 1, 0, 14, True, (2, 2, 0)),        # LeNet conv2
    (32, 64, 1, 1, 0, 7, False, None),            # 1x1: the activation is the GEMM operand
    (8, 16, 3, 1, 1, 8, True, (3, 2, 1)),
])
def test_conv_block_tf32_fwd_bwd(cin, cout, k, s, p, hw, relu, pool):
    """Operands are tf32-exact (inputs, weights, incoming gradient), so kernel and fp32 oracle compute the same products and differ only in
    accumulation order: ReLU masks and pool arg-maxes agree, and the gradients can be held to a tight tolerance.  (With unrounded operands
    the forward agrees to 1e-3 but ~0.1 % of the ReLU decisions flip, which alone moves a weight gradient by ~sqrt(1e-3) = 3 %.)"""
    # seed 1 gives the overlapping 3x3 pool a window whose two largest activations agree to the last bit in one accumulation order and
    # differ by one ulp in the other: the arg-max (a legitimate tie) then routes one gradient element to a different pixel
    torch.manual_seed(2 if pool == (3, 2, 1) else 1)
    dev = "cuda"
    B = 8
    conv = dnn.set_compute_dtype(dnn.Conv2d(cin, cout, k, stride=s, padding=p, activation="relu" if relu else None, pool=pool), "tf32").to(dev)
    with torch.no_grad():
        conv.weight.copy_(tf32_round(conv.weight))
    x32 = tf32_round(torch.randn(B, cin, hw, hw, device=dev))
    x = act4(x32).requires_grad_(True)
    y = conv(x)
    assert y.dtype == torch.float32
    xr = x32.clone().requires_grad_(True)
    w = conv.weight.detach().clone().requires_grad_(True)
    b = conv.bias.detach().clone().requires_grad_(True)
    with torch.backends.cudnn.flags(allow_tf32=False):
        z = F.conv2d(xr, w, b, stride=s, padding=p)
        if relu:
            z = F.relu(z)
        yr = z if pool is None else F.max_pool2d(z, pool[0], pool[1], pool[2])
        assert y.shape == yr.shape
        assert rel_err(y, yr) < 1e-5
        g = tf32_round(torch.randn_like(yr))
        y.backward(act4(g))
        yr.backward(g)
    torch.cuda.synchronize()
    # dY is tf32-exact only before the ReLU / pool routing, which keeps values as they are: still exact.  What remains: the rare
    # sign flip of a pre-activation that is zero to within accumulation order.
    assert rel_err(conv.weight.grad, w.grad) < TOL
    assert rel_err(conv.bias.grad, b.grad) < TOL
    assert rel_err(x.grad, xr.grad) < TOL


def test_conv_tf32_unrounded_operands_forward():
    """Plain fp32 tensors (what training feeds): the copy engine / tensor core reduce the operands to tf32, error ~2^-11 per product."""
    torch.manual_seed(11)
    dev = "cuda"
    conv = dnn.set_compute_dtype(dnn.Conv2d(64, 192, 5, padding=2, activation="relu", pool=(2, 2, 0)), "tf32").to(dev)
    x32 = torch.randn(16, 64, 4, 4, device=dev)
    y = conv(act4(x32))
    with torch.backends.cudnn.flags(allow_tf32=False):
        yr = F.max_pool2d(F.relu(F.conv2d(x32, conv.weight, conv.bias, padding=2)), 2, 2)
    assert rel_err(y, yr) < TOL


def test_conv_first_layer_tf32_from_fp32_nchw():
    torch.manual_seed(2)
    dev = "cuda"
    conv = dnn.set_compute_dtype(dnn.Conv2d(3, 64, 11, stride=4, padding=5, activation="relu", pool=(2, 2, 0)), "tf32").to(dev)
    with torch.no_grad():
        conv.weight.copy_(tf32_round(conv.weight))
    x = tf32_round(torch.randn(16, 3, 32, 32, device=dev))
    y = conv(x)
    w = conv.weight.detach().clone().requires_grad_(True)
    with torch.backends.cudnn.flags(allow_tf32=False):
        yr = F.max_pool2d(F.relu(F.conv2d(x, w, conv.bias.detach(), stride=4, padding=5)), 2, 2)
        assert rel_err(y, yr) < 1e-5
        g = tf32_round(torch.randn_like(yr))
        y.backward(act4(g))
        yr.backward(g)
    assert rel_err(conv.weight.grad, w.grad) < TOL


@pytest.mark.parametrize("i,o,relu,in_map", [(256, 10, False, None), (400, 120, True, (16, 5, 5)), (120, 84, True, None), (84, 10, False, None)])
def test_linear_tf32_fwd_bwd(i, o, relu, in_map):
    torch.manual_seed(3)
    dev = "cuda"
    B = 64
    lin = dnn.set_compute_dtype(dnn.Linear(i, o, activation="relu" if relu else None, in_map=in_map), "tf32").to(dev)
    x32 = torch.randn(B, i, device=dev)
    w = lin.weight.detach().clone().requires_grad_(True)
    b = lin.bias.detach().clone().requires_grad_(True)
    xr = x32.clone().requires_grad_(True)
    if in_map is None:
        xin = act2(x32)
    else:
        c, h, ww = in_map
        xin = act2(x32.view(B, c, h, ww).permute(0, 2, 3, 1).reshape(B, -1).contiguous())
    xin = xin.requires_grad_(True)
    y = lin(xin)
    yr = F.linear(xr, w, b)
    if relu:
        yr = F.relu(yr)
    assert rel_err(y, yr) < TOL
    g = torch.randn_like(yr)
    y.backward(act2(g))
    yr.backward(g)
    assert rel_err(lin.weight.grad, w.grad) < 2 * TOL
    assert rel_err(lin.bias.grad, b.grad) < 2 * TOL
    gx = xin.grad.float()
    if in_map is not None:
        c, h, ww = in_map
        gx = gx.view(B, h, ww, c).permute(0, 3, 1, 2).reshape(B, -1)
    assert rel_err(gx, xr.grad) < 2 * TOL


def test_pool_relu_dropout_fp32():
    torch.manual_seed(4)
    dev = "cuda"
    x32 = torch.randn(4, 12, 9, 9, device=dev)
    x = act4(x32).requires_grad_(True)
    xr = x32.clone().requires_grad_(True)
    y = OF.max_pool2d(OF.relu(x), 3, 2, 1)
    yr = F.max_pool2d(F.relu(xr), 3, 2, 1)
    assert y.dtype == torch.float32 and torch.equal(y.contiguous(), yr)
    g = torch.randn_like(yr)
    y.backward(act4(g))
    yr.backward(g)
    assert torch.allclose(x.grad.contiguous(), xr.grad)
    d = OF.dropout(act4(torch.ones(8, 16, 6, 6, device=dev)), 0.5, True, channel_mode=True)
    per_channel = d.float().reshape(8, 16, -1)
    assert ((per_channel.amax(2) == per_channel.amin(2))).all()          # whole channels are kept or dropped
    assert set(per_channel.unique().tolist()) <= {0.0, 2.0}


@pytest.mark.parametrize("model", ["alexnet", "lenet"])
def test_model_step_tf32_matches_fp32_torch(model):
    """A few SGD steps of the tf32 model against the same model run by ATen/cuDNN in strict fp32 on the CPU layout."""
    from distributed_ml_pytorch_b200.models import build_model
    from distributed_ml_pytorch_b200.parallel.fused_sgd import FusedSGD
    torch.manual_seed(5)
    dev = "cuda"
    ref = build_model(model).eval()                # eval: no dropout, so both sides are deterministic
    net = build_model(model, dtype="tf32")
    net.load_state_dict(ref.state_dict())
    net = net.to(dev).eval()
    opt = FusedSGD(net.parameters(), lr=0.05, model=net)
    ropt = torch.optim.SGD(ref.parameters(), lr=0.05)
    x = torch.randn(32, 3, 32, 32)
    y = torch.randint(0, 10, (32,))
    for step in range(4):
        opt.zero_grad()
        out = net(x.to(dev))
        assert out.dtype == torch.float32
        loss = OF.cross_entropy(out, y.to(dev))
        loss.backward()
        opt.step()
        ropt.zero_grad()
        rloss = F.cross_entropy(ref(x), y)
        rloss.backward()
        ropt.step()
        assert abs(float(loss) - float(rloss)) < 5e-3 * max(1.0, abs(float(rloss))), (step, float(loss), float(rloss))
    sd, rsd = net.state_dict(), ref.state_dict()
    for k in rsd:
        assert rel_err(sd[k].cpu(), rsd[k]) < 5e-3, k


def test_tf32_alexnet_weights_are_read_in_place():
    """Every arena-backed AlexNet weight but conv1 (K = 363) is a legal TMA operand as stored: its shadow aliases the fp32 master."""
    from distributed_ml_pytorch_b200.models import build_model
    from distributed_ml_pytorch_b200.parallel.fused_sgd import FusedSGD
    net = build_model("alexnet", dtype="tf32").cuda()
    FusedSGD(net.parameters(), lr=0.01, model=net)
    net(torch.randn(4, 3, 32, 32, device="cuda"))
    mods = [m for m in net.modules() if isinstance(m, (dnn.Conv2d, dnn.Linear))]
    aliased = [m._shadow.data_ptr() == m.weight.data_ptr() for m in mods]
    assert aliased == [False, True, True, True, True, True]
    assert net._dmlb_arena.shadow is None

"""Numerics of every CUDA op against a plain PyTorch fp32 reference of the same op."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from distributed_ml_pytorch_b200 import nn as dnn
from distributed_ml_pytorch_b200.ops import functional as OF
from distributed_ml_pytorch_b200.ops import kernels as K


def bf(x):
    return x.to(torch.bfloat16).float()


def rel_err(a, b):
    return ((a.float() - b.float()).norm() / (b.float().norm() + 1e-12)).item()


def to_act(x):
    t, _ = OF.as_act4(x.to(torch.bfloat16))
    return t


@pytest.mark.parametrize("M,N,Kd,amn,bmn", [(256, 384, 1728, 0, 0), (64, 10, 256, 0, 0), (256, 1728, 384, 0, 1), (384, 1728, 256, 1, 1)])
def test_gemm_matches_fp32(M, N, Kd, amn, bmn):
    torch.manual_seed(0)
    dev = "cuda"
    A = bf(torch.randn(M, Kd, device=dev))
    B = bf(torch.randn(N, Kd, device=dev))
    ref = A @ B.t()
    a = (A.t() if amn else A).contiguous().to(torch.bfloat16)
    b = (B.t() if bmn else B).contiguous().to(torch.bfloat16)
    out = torch.zeros(M, N, device=dev, dtype=torch.float32)
    K.gemm(a, a.stride(0), bool(amn), b, b.stride(0), bool(bmn), M, N, Kd, out, N, atomic=False)
    torch.cuda.synchronize()
    assert rel_err(out, ref) < 1e-5


@pytest.mark.parametrize("cin,cout,k,s,p,hw,relu,pool", [
    (64, 192, 5, 1, 2, 4, True, (2, 2, 0)),
    (192, 384, 3, 1, 1, 2, True, None),
    (16, 32, 3, 2, 1, 9, False, None),
    (6, 16, 5, 1, 0, 14, True, (2, 2, 0)),
    (32, 64, 1, 1, 0, 7, False, None),
    (8, 16, 3, 1, 1, 8, True, (3, 2, 1)),
])
def test_conv_block_fwd_bwd(cin, cout, k, s, p, hw, relu, pool):
    torch.manual_seed(1)
    dev = "cuda"
    B = 8
    conv = dnn.Conv2d(cin, cout, k, stride=s, padding=p, activation="relu" if relu else None, pool=pool).to(dev)
    x32 = bf(torch.randn(B, cin, hw, hw, device=dev))
    x = to_act(x32).requires_grad_(True)
    y = conv(x)
    # fp32 reference with bf16-rounded weights (the kernel's operand precision)
    xr = x32.clone().requires_grad_(True)
    w = bf(conv.weight.detach()).requires_grad_(True)
    b = conv.bias.detach().clone().requires_grad_(True)
    z = F.conv2d(xr, w, b, stride=s, padding=p)
    if relu:
        z = bf(F.relu(z))  # the kernel stores the activation in bf16 before pooling
    yr = z if pool is None else F.max_pool2d(z, pool[0], pool[1], pool[2])
    assert y.shape == yr.shape
    assert rel_err(y, yr) < 1e-2
    g = bf(torch.randn_like(yr))
    y.backward(to_act(g))
    yr.backward(g)
    torch.cuda.synchronize()
    assert rel_err(conv.weight.grad, w.grad) < 2e-2
    assert rel_err(conv.bias.grad, b.grad) < 2e-2
    assert rel_err(x.grad, xr.grad) < 2e-2


def test_conv_first_layer_from_fp32_nchw():
    torch.manual_seed(2)
    dev = "cuda"
    conv = dnn.Conv2d(3, 64, 11, stride=4, padding=5, activation="relu", pool=(2, 2, 0)).to(dev)
    x = torch.randn(16, 3, 32, 32, device=dev)
    y = conv(x)
    w = bf(conv.weight.detach()).requires_grad_(True)
    z = bf(F.relu(F.conv2d(bf(x), w, conv.bias.detach(), stride=4, padding=5)))
    yr = F.max_pool2d(z, 2, 2)
    assert rel_err(y, yr) < 1e-2
    g = bf(torch.randn_like(yr))
    y.backward(to_act(g))
    yr.backward(g)
    assert rel_err(conv.weight.grad, w.grad) < 2e-2


@pytest.mark.parametrize("i,o,relu,in_map", [(256, 10, False, None), (400, 120, True, (16, 5, 5)), (120, 84, True, None), (84, 10, False, None)])
def test_linear_fwd_bwd(i, o, relu, in_map):
    torch.manual_seed(3)
    dev = "cuda"
    B = 64
    lin = dnn.Linear(i, o, activation="relu" if relu else None, in_map=in_map).to(dev)
    x32 = bf(torch.randn(B, i, device=dev))
    w = bf(lin.weight.detach()).requires_grad_(True)
    b = lin.bias.detach().clone().requires_grad_(True)
    xr = x32.clone().requires_grad_(True)
    if in_map is None:
        xin, _ = OF.as_act2(x32.to(torch.bfloat16))
    else:
        # the kernel sees the channels-last flattening of the same feature map
        c, h, ww = in_map
        xin, _ = OF.as_act2(x32.view(B, c, h, ww).permute(0, 2, 3, 1).reshape(B, -1).contiguous().to(torch.bfloat16))
    xin = xin.requires_grad_(True)
    y = lin(xin)
    yr = F.linear(xr, w, b)
    if relu:
        yr = F.relu(yr)
    assert rel_err(y, yr) < 1e-2
    g = bf(torch.randn_like(yr))
    gin, _ = OF.as_act2(g.to(torch.bfloat16))
    y.backward(gin)
    yr.backward(g)
    assert rel_err(lin.weight.grad, w.grad) < 2e-2
    assert rel_err(lin.bias.grad, b.grad) < 2e-2
    gx = xin.grad.float()
    if in_map is not None:
        c, h, ww = in_map
        gx = gx.view(B, h, ww, c).permute(0, 3, 1, 2).reshape(B, -1)
    assert rel_err(gx, xr.grad) < 2e-2


@pytest.mark.parametrize("k,s,p,hw,c", [(2, 2, 0, 8, 64), (3, 2, 1, 9, 16), (2, 2, 0, 10, 6)])
def test_maxpool_fwd_bwd(k, s, p, hw, c):
    torch.manual_seed(4)
    dev = "cuda"
    x32 = bf(torch.randn(4, c, hw, hw, device=dev))
    x = to_act(x32).requires_grad_(True)
    xr = x32.clone().requires_grad_(True)
    y = OF.max_pool2d(x, k, s, p)
    yr = F.max_pool2d(xr, k, s, p)
    assert torch.equal(y.float(), yr)
    g = bf(torch.randn_like(yr))
    y.backward(to_act(g))
    yr.backward(g)
    assert rel_err(x.grad, xr.grad) < 1e-2


def test_relu_and_flatten():
    dev = "cuda"
    x32 = bf(torch.randn(4, 16, 5, 5, device=dev))
    x = to_act(x32).requires_grad_(True)
    y = OF.flatten(OF.relu(x))
    ref = F.relu(x32).permute(0, 2, 3, 1).reshape(4, -1)
    assert torch.equal(y.float(), ref)
    y.sum().backward()
    assert torch.equal(x.grad.float(), (x32 > 0).float())


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_softmax_xent(dtype):
    torch.manual_seed(5)
    dev = "cuda"
    B, C = 64, 10
    logits32 = bf(torch.randn(B, C, device=dev) * 3)
    labels = torch.randint(0, C, (B,), device=dev)
    metrics = torch.zeros(2, device=dev)
    if dtype == torch.bfloat16:
        lg, _ = OF.as_act2(logits32.to(torch.bfloat16))
    else:
        lg = logits32.clone()
    lg = lg.requires_grad_(True)
    loss = OF.cross_entropy(lg, labels, metrics)
    lr = logits32.clone().requires_grad_(True)
    ref = F.cross_entropy(lr, labels)
    assert abs(loss.item() - ref.item()) < 1e-4
    assert metrics[1].item() == (logits32.argmax(1) == labels).sum().item()
    (loss * 2.0).backward()
    (ref * 2.0).backward()
    assert rel_err(lg.grad, lr.grad) < 1e-2


def test_dropout_statistics_and_backward():
    dev = "cuda"
    x, _ = OF.as_act2(torch.ones(256, 128, device=dev, dtype=torch.bfloat16))
    x = x.requires_grad_(True)
    y = OF.dropout(x, 0.5, True)
    keep = (y.float() != 0).float().mean().item()
    assert 0.45 < keep < 0.55
    assert torch.all((y.float() == 0) | (y.float() == 2.0))
    y2 = OF.dropout(x, 0.5, True)
    assert not torch.equal(y.float(), y2.float())  # device-side step counter advances
    y.sum().backward()
    assert torch.equal(x.grad.float(), y.detach().float())
    # channel mode: whole (sample, channel) planes share one draw
    x4 = to_act(torch.ones(8, 16, 5, 5, device=dev))
    y4 = OF.dropout(x4, 0.5, True, channel_mode=True).float()
    per_plane = y4.reshape(8, 16, -1)
    assert torch.all(per_plane.min(-1).values == per_plane.max(-1).values)


def test_dropout_masks_are_independent_across_elements_and_calls():
    """The mask generator is a counter hash (splitmix of seed, call counter, element index), not Philox: check what training relies on.
    (a) keep-rate per call is binomial around 1-p (chi-square over 64 calls); (b) no correlation between neighbouring elements
    (lags 1, 2, 8, row pitch) inside a call; (c) no correlation between the masks of consecutive calls, element by element;
    (d) the 4-bit patterns of 4 consecutive elements are uniform (chi-square, 15 dof)."""
    dev = "cuda"
    rows, cols, p = 512, 256, 0.5
    n = rows * cols
    x, _ = OF.as_act2(torch.ones(rows, cols, device=dev, dtype=torch.bfloat16))
    calls = 64
    masks = torch.stack([(OF.dropout(x, p, True).float()[:, :cols] != 0).float() for _ in range(calls)])       # [calls, rows, cols]
    # (a) each call keeps Binomial(n, 1-p): z-scores of the per-call keep counts ~ N(0,1) -> sum of squares ~ chi2(calls)
    z = (masks.sum((1, 2)) - n * (1 - p)) / (n * p * (1 - p)) ** 0.5
    chi2 = float((z ** 2).sum())
    assert 30.0 < chi2 < 115.0, chi2                      # chi2(64): 0.01 % tails at ~29 / ~116
    assert float(z.abs().max()) < 4.5
    # (b) lag correlations inside a call: each estimate has sigma = 1/sqrt(n * calls) ~ 3.5e-4
    m = masks - masks.mean()
    flat = m.reshape(calls, -1)
    var = float((flat ** 2).mean())
    for lag in (1, 2, 8, cols):
        c = float((flat[:, :-lag] * flat[:, lag:]).mean()) / var
        assert abs(c) < 2.5e-3, (lag, c)
    # (c) consecutive calls, same element
    c = float((m[:-1] * m[1:]).mean()) / var
    assert abs(c) < 2.5e-3, c
    # (d) 4-element patterns
    bits = masks.reshape(calls, -1, 4)
    code = (bits * torch.tensor([1.0, 2.0, 4.0, 8.0], device=dev)).sum(-1).long().flatten()
    counts = torch.bincount(code, minlength=16).float()
    expect = code.numel() / 16.0
    chi2p = float(((counts - expect) ** 2 / expect).sum())
    assert chi2p < 50.0, chi2p                            # chi2(15): 1e-5 tail at ~48


def test_sgd_flat_and_shadow_refresh():
    torch.manual_seed(6)
    dev = "cuda"
    n = 4096
    p = torch.randn(n, device=dev)
    g = torch.randn(n, device=dev)
    acc = torch.zeros(n, device=dev)
    mom = torch.zeros(n, device=dev)
    lr = torch.tensor([0.1], device=dev)
    p0 = p.clone()
    K.sgd_flat(p, g, acc, mom, n, lr, mu=0.9, wd=0.0)
    K.sgd_flat(p, g, acc, mom, n, lr, mu=0.9, wd=0.0)
    v1 = g
    v2 = 0.9 * v1 + g
    torch.testing.assert_close(p, p0 - 0.1 * v1 - 0.1 * v2, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(acc, p - p0, rtol=1e-5, atol=1e-5)
    conv = dnn.Conv2d(6, 16, 5).to(dev)
    sh = conv.shadow()
    ref = conv.weight.detach().permute(0, 2, 3, 1).reshape(16, -1).to(torch.bfloat16)
    assert torch.equal(sh[:, :150], ref)
    with torch.no_grad():
        conv.weight.add_(1.0)
    assert torch.equal(conv.shadow()[:, :150], conv.weight.detach().permute(0, 2, 3, 1).reshape(16, -1).to(torch.bfloat16))


@pytest.mark.parametrize("name", ["alexnet", "lenet"])
def test_model_grads_match_fp32_reference(name):
    """Whole-network forward/backward against the same architecture in fp32 torch ops.

    bf16 activations flip a few ReLU gates, so deep gradients differ from fp32 by several percent no matter
    who computes them; the yardstick is therefore torch's own bf16 autocast run of the same network: our
    error against fp32 must be of the same size as autocast's.
    """
    torch.manual_seed(7)
    dev = "cuda"
    from distributed_ml_pytorch_b200.models import build_model
    net = build_model(name).to(dev)
    ref = build_model(name)  # CPU instance -> runs the torch.nn.functional path in fp32
    ref.load_state_dict(net.state_dict())
    net.eval(); ref.eval()   # no dropout noise in the comparison
    x = torch.randn(64, 3, 32, 32)
    labels = torch.randint(0, 10, (64,))
    out = net(x.to(dev))
    loss = OF.cross_entropy(out, labels.to(dev))
    loss.backward()
    out_r = ref(x)
    loss_r = F.cross_entropy(out_r, labels)
    loss_r.backward()
    g32 = [p.grad.clone() for p in ref.parameters()]
    ref.zero_grad()
    with torch.autocast("cpu", dtype=torch.bfloat16):
        loss_a = F.cross_entropy(ref(x), labels)
    loss_a.backward()
    assert abs(loss.item() - loss_r.item()) < 3e-2
    assert rel_err(out.cpu(), out_r) < 3e-2
    for (n1, p1), p2, g in zip(net.named_parameters(), ref.parameters(), g32):
        assert p1.grad is not None, n1
        ours = rel_err(p1.grad.cpu(), g)
        autocast = rel_err(p2.grad, g)
        cos = F.cosine_similarity(p1.grad.cpu().flatten(), g.flatten(), dim=0).item()
        assert cos > 0.97, (n1, cos)
        assert ours < max(2.5 * autocast, 3e-2), (n1, ours, autocast)


@pytest.mark.parametrize("M,N,Kd,split", [(256, 384, 1728, 13), (64, 10, 4096, 8), (100, 72, 1000, 4)])
def test_gemm_bias_act_splitk_c_abi(M, N, Kd, split):
    """Split-K forward entry point (fp32 workspace scratch + fused finish) against fp32 matmul."""
    from distributed_ml_pytorch_b200 import _native
    torch.manual_seed(8)
    dev = "cuda"
    A = bf(torch.randn(M, Kd, device=dev))
    Wt = bf(torch.randn(N, Kd, device=dev) * 0.05)
    bias = torch.randn(N, device=dev)
    ref = F.relu(A @ Wt.t() + bias)
    a, w = A.to(torch.bfloat16).contiguous(), Wt.to(torch.bfloat16).contiguous()
    ldy = K.pad8(N)
    y = torch.zeros(M, ldy, device=dev, dtype=torch.bfloat16)
    ws = torch.zeros(M * ((N + 3) // 4 * 4), device=dev, dtype=torch.float32)
    for _ in range(2):   # the second call proves the workspace was left clean
        _native.call("dmlb_gemm_bias_act_splitk", a.data_ptr(), Kd, w.data_ptr(), Kd, M, N, Kd, ws.data_ptr(), bias.data_ptr(), 1,
                     y.data_ptr(), ldy, split, K.stream_ptr())
        torch.cuda.synchronize()
        assert rel_err(y[:, :N], ref) < 1e-2
        assert ws.abs().max().item() == 0.0


@pytest.mark.gpu
def test_resident_loader_kernel_matches_totensor_normalize():
    from distributed_ml_pytorch_b200.utils.data import GpuResidentLoader
    g = torch.Generator().manual_seed(0)
    imgs = torch.randint(0, 256, (300, 32, 32, 3), generator=g, dtype=torch.uint8)
    labels = torch.randint(0, 10, (300,), generator=g)
    ld = GpuResidentLoader(imgs, labels, batch_size=64, device="cuda", shuffle=True, seed=5)
    idx = torch.randperm(300, device="cuda")[:64]
    x, y = ld.gather(idx)
    ref = (imgs[idx.cpu()].permute(0, 3, 1, 2).float() / 255 - 0.5) / 0.5
    assert torch.allclose(x.cpu(), ref, atol=1e-6) and torch.equal(y.cpu(), labels[idx.cpu()])
    n = sum(yy.numel() for _, yy in ld)
    assert n == 4 * 64                                  # drop_last keeps the shape fixed for the captured step


@pytest.mark.gpu
def test_train_step_orders_its_copy_behind_the_producer_of_a_device_batch():
    """GpuResidentLoader batches are produced on the caller's stream; TrainStep copies them into its staging slot on another stream.
    Put a long-running kernel in front of the gather: without the stream ordering the copy would read the batch before it is written."""
    from distributed_ml_pytorch_b200.engine import TrainStep
    from distributed_ml_pytorch_b200.models import build_model
    from distributed_ml_pytorch_b200.parallel import FusedSGD
    from distributed_ml_pytorch_b200.utils.data import GpuResidentLoader
    g = torch.Generator().manual_seed(1)
    imgs = torch.randint(0, 256, (256, 32, 32, 3), generator=g, dtype=torch.uint8)
    labels = torch.randint(0, 10, (256,), generator=g)
    ld = GpuResidentLoader(imgs, labels, batch_size=64, device="cuda", shuffle=False)
    net = build_model("lenet").cuda().train()
    ts = TrainStep(net, FusedSGD(net.parameters(), lr=0.01, model=net), 64, (3, 32, 32), device=torch.device("cuda", 0))
    big = torch.randn(8192, 8192, device="cuda")
    for step in range(4):
        idx = torch.arange(step * 64, step * 64 + 64, device="cuda")
        for _ in range(3):
            big = torch.tanh(big @ big * 1e-4)                       # ~ms of work queued on the current stream, ahead of the gather
        x, y = ld.gather(idx)
        slot = ts.step_count & 1
        ts(x, y)
        del x, y                                                    # the allocator may hand the blocks out again right away
        junk = torch.full((64, 3, 32, 32), 7.0, device="cuda")      # ... to this
        ts.flush()
        sx, sy = ts._staging[slot]
        ref = (imgs[step * 64:step * 64 + 64].permute(0, 3, 1, 2).float() / 255 - 0.5) / 0.5
        assert torch.allclose(sx.float().cpu(), ref, atol=1e-6), step
        assert torch.equal(sy.cpu(), labels[step * 64:step * 64 + 64]), step
        del junk


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["alexnet", "lenet"])
def test_eval_forward_at_reference_test_batch_size(name):
    """The reference evaluates with one 10000-image batch (main.py:143 --test-batch-size 10000)."""
    from distributed_ml_pytorch_b200 import models
    torch.manual_seed(0)
    net = (models.AlexNet() if name == "alexnet" else models.LeNet()).cuda().eval()
    ref = (models.AlexNet() if name == "alexnet" else models.LeNet()).eval()
    ref.load_state_dict({k: v.cpu() for k, v in net.state_dict().items()})
    x = torch.randn(10000, 3, 32, 32)
    with torch.no_grad():
        y = net(x.cuda()).float().cpu()
        y_ref = ref(x)
    assert y.shape == (10000, 10)
    assert ((y - y_ref).norm() / y_ref.norm()).item() < 3e-2


@pytest.mark.gpu
def test_gemm_on_every_visible_device_in_one_process():
    """The > 48 KB shared-memory opt-in is a per-device function attribute; a process that drives several GPUs must get it on each."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    for d in range(min(2, torch.cuda.device_count())):
        with torch.cuda.device(d):
            dev = torch.device("cuda", d)
            torch.manual_seed(d)
            A = bf(torch.randn(256, 512, device=dev))
            B = bf(torch.randn(192, 512, device=dev))
            out = torch.zeros(256, 192, device=dev)
            K.gemm(A.to(torch.bfloat16), 512, False, B.to(torch.bfloat16), 512, False, 256, 192, 512, out, 192)
            torch.cuda.synchronize(dev)
            assert rel_err(out, A @ B.t()) < 1e-5


@pytest.mark.gpu
@pytest.mark.parametrize("M,N,Kd,bias,relu,out_f32", [
    (512, 512, 512, False, False, True),       # 2 x 2 CTA pairs, one tile per cluster
    (300, 700, 200, True, True, False),        # ragged everything, fused bias + ReLU, bf16 out
    (2600, 3600, 128, True, False, True),      # 165 tiles: persistent pairs, double-buffered TMEM, banded tile order
    (4096, 2048, 1024, False, False, False),   # persistent, several ring revolutions per tile
])
def test_gemm_cta_pair_kernel_matches_fp32(M, N, Kd, bias, relu, out_f32):
    """256-wide tiles of a K-major GEMM run on CTA pairs (tcgen05 cta_group::2): csrc/gemm_tcgen05.cu pair kernels."""
    torch.manual_seed(0)
    dev = "cuda"
    A = bf(torch.randn(M, Kd, device=dev))
    B = bf(torch.randn(N, Kd, device=dev) * 0.1)
    b = torch.randn(N, device=dev) if bias else None
    ref = A @ B.t() + (b if bias else 0)
    ref = F.relu(ref) if relu else ref
    ldo = K.pad8(N)
    out = torch.zeros(M, ldo, device=dev, dtype=torch.float32 if out_f32 else torch.bfloat16)
    K.gemm(A.to(torch.bfloat16), Kd, False, B.to(torch.bfloat16), Kd, False, M, N, Kd, out, ldo, bias=b, relu=relu, tile_n=256)
    torch.cuda.synchronize()
    assert rel_err(out[:, :N], ref) < (1e-5 if out_f32 else 1e-2)

"""Batch norm (+ residual + ReLU), global average pooling and the ResNets against plain PyTorch fp32 references."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from distributed_ml_pytorch_b200 import nn as dnn
from distributed_ml_pytorch_b200.ops import functional as OF


def bf(x):
    return x.to(torch.bfloat16).float()


def rel_err(a, b):
    return ((a.float() - b.float()).norm() / (b.float().norm() + 1e-12)).item()


def to_act(x):
    return OF.as_act4(x.to(torch.bfloat16))[0]


@pytest.mark.parametrize("C,hw,res,relu", [(64, 8, True, True), (24, 5, False, True), (256, 4, True, False), (2048, 2, False, False)])
def test_batch_norm_fwd_bwd(C, hw, res, relu):
    torch.manual_seed(0)
    dev = "cuda"
    B = 8
    bn = dnn.BatchNorm2d(C, activation="relu" if relu else None).to(dev)
    with torch.no_grad():
        bn.weight.uniform_(0.5, 1.5)
        bn.bias.uniform_(-0.5, 0.5)
    x32 = bf(torch.randn(B, C, hw, hw, device=dev) * 2 + 0.5)
    r32 = bf(torch.randn(B, C, hw, hw, device=dev)) if res else None
    x = to_act(x32).requires_grad_(True)
    r = to_act(r32).requires_grad_(True) if res else None
    y = bn(x, r)
    xr = x32.clone().requires_grad_(True)
    rr = r32.clone().requires_grad_(True) if res else None
    w = bn.weight.detach().clone().requires_grad_(True)
    b = bn.bias.detach().clone().requires_grad_(True)
    rm, rv = torch.zeros(C, device=dev), torch.ones(C, device=dev)
    yr = F.batch_norm(xr, rm, rv, w, b, True, 0.1, 1e-5)
    if res:
        yr = yr + rr
    if relu:
        yr = F.relu(yr)
    assert rel_err(y, yr) < 1e-2
    torch.testing.assert_close(bn.running_mean, rm, rtol=1e-3, atol=1e-3)
    torch.testing.assert_close(bn.running_var, rv, rtol=1e-3, atol=1e-3)
    g = bf(torch.randn_like(yr))
    y.backward(to_act(g))
    yr.backward(g)
    assert rel_err(x.grad, xr.grad) < 3e-2
    assert rel_err(bn.weight.grad, w.grad) < 2e-2
    assert rel_err(bn.bias.grad, b.grad) < 2e-2
    if res:
        assert rel_err(r.grad, rr.grad) < 1e-2
    # inference mode uses the running statistics
    bn.eval()
    ye = bn(to_act(x32))
    yer = F.batch_norm(x32, bn.running_mean, bn.running_var, bn.weight, bn.bias, False, 0.1, 1e-5)
    if relu:
        yer = F.relu(yer)
    assert rel_err(ye, yer) < 1e-2


def test_global_avg_pool():
    dev = "cuda"
    x32 = bf(torch.randn(4, 64, 7, 7, device=dev))
    x = to_act(x32).requires_grad_(True)
    y = OF.global_avg_pool(x)
    assert rel_err(y, x32.mean(dim=(2, 3))) < 5e-3
    g = bf(torch.randn(4, 64, device=dev))
    y.backward(OF.as_act2(g.to(torch.bfloat16))[0])
    assert rel_err(x.grad, (g / 49)[:, :, None, None].expand(4, 64, 7, 7)) < 5e-3


@pytest.mark.parametrize("name,hw,classes,batch", [("resnet18", 32, 10, 16), ("resnet50", 64, 100, 8)])
def test_resnet_grads_match_fp32_reference(name, hw, classes, batch):
    torch.manual_seed(1)
    dev = "cuda"
    from distributed_ml_pytorch_b200.models import resnet
    net = getattr(resnet, name)(num_classes=classes, small_input=(hw <= 32)).to(dev).train()
    ref = getattr(resnet, name)(num_classes=classes, small_input=(hw <= 32)).train()
    ref.load_state_dict(net.state_dict())
    x = torch.randn(batch, 3, hw, hw)
    labels = torch.randint(0, classes, (batch,))
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
    assert abs(loss.item() - loss_r.item()) < 0.1
    # yardstick: torch's own bf16 autocast run of the same network against fp32 (tiny batch-norm batches make deep
    # bf16 gradients noisy for everybody, so the comparison is relative, on the whole gradient vector)
    ours = torch.cat([p.grad.float().cpu().flatten() for p in net.parameters()])
    auto = torch.cat([p.grad.flatten() for p in ref.parameters()])
    full = torch.cat([g.flatten() for g in g32])
    cos_ours = F.cosine_similarity(ours, full, dim=0).item()
    cos_auto = F.cosine_similarity(auto, full, dim=0).item()
    print(name, "cosine to fp32 gradient: ours", cos_ours, "torch autocast", cos_auto)
    # as good as torch autocast; an untrained 50-layer net at batch 8 is chaotic for both (ours: 0.115-0.139 over runs, split-K atomics
    # reorder the sums; autocast: 0.138), so the whole-vector bound is loose there and the last layer, which sees only forward error, is
    # checked tightly instead
    assert cos_ours > cos_auto - 0.15
    last_ours = list(net.parameters())[-2].grad.float().cpu().flatten()
    last_full = g32[-2].flatten()
    cos_last = F.cosine_similarity(last_ours, last_full, dim=0).item()
    print(name, "cosine of the classifier weight gradient to fp32:", cos_last)
    assert cos_last > 0.9
    if name == "resnet18":
        assert cos_ours > 0.9
    for (n1, p1) in net.named_parameters():
        assert p1.grad is not None and torch.isfinite(p1.grad).all(), n1


def test_resnet18_train_step_graph_learns():
    from distributed_ml_pytorch_b200.engine import TrainStep
    from distributed_ml_pytorch_b200.models import build_model
    from distributed_ml_pytorch_b200.parallel import FusedSGD
    torch.manual_seed(2)
    dev = torch.device("cuda")
    net = build_model("resnet18").to(dev).train()
    opt = FusedSGD(net.parameters(), lr=0.05, momentum=0.9, model=net)
    ts = TrainStep(net, opt, 32, (3, 32, 32))
    assert ts.graph is not None
    x = torch.randn(32, 3, 32, 32, device=dev)
    y = torch.randint(0, 10, (32,), device=dev)
    losses = [ts(x, y, sync=True).loss for _ in range(30)]
    assert all(l == l for l in losses)
    assert losses[-1] < 0.5 * losses[0]       # memorises one batch

import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from distributed_ml_pytorch_b200 import nn as dnn
from distributed_ml_pytorch_b200.models import build_model
from distributed_ml_pytorch_b200.ops import functional as OF, kernels as K
from distributed_ml_pytorch_b200.parallel.fused_sgd import FusedSGD
for model in ("lenet", "alexnet"):
  for dtype in ("tf32", "bf16"):
    torch.manual_seed(0)
    net = build_model(model, dtype=dtype).cuda()
    opt = FusedSGD(net.parameters(), lr=0.05, model=net)
    x = torch.randn(16, 3, 32, 32, device="cuda"); y = torch.randint(0, 10, (16,), device="cuda")
    for _ in range(2):
        opt.zero_grad(); OF.cross_entropy(net(x), y).backward(); opt.step()
    torch.cuda.synchronize()
    ss = net._dmlb_shadows
    mods = [m for m in ss.modules if not m._flat]
    got = [m._shadow.clone() for m in mods]
    K.refresh_shadows(ss._table, ss._n_table, dt=mods[0].compute_dtype)
    torch.cuda.synchronize()
    for m, g in zip(mods, got):
        d = (m._shadow.float() - g.float()).abs().max().item()
        print(model, dtype, type(m).__name__, tuple(m.weight.shape), "shadow", tuple(m._shadow.shape), "max |fused - refresh| =", d)

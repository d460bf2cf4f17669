import sys, torch
sys.path.insert(0, ".")
from distributed_ml_pytorch_b200.models import build_model
from distributed_ml_pytorch_b200.ops import functional as OF, kernels as K
from distributed_ml_pytorch_b200.parallel import FusedSGD
names = ["im2col_nhwc", "im2col_nchw_f32", "col2im_nhwc"]
counts = {}
for n in names:
    def mk(n, f):
        def w(*a, **k):
            counts[n] = counts.get(n, 0) + 1
            return f(*a, **k)
        return w
    setattr(K, n, mk(n, getattr(K, n)))
for model, hw, dt in [("alexnet", 32, "tf32"), ("alexnet", 32, "bf16"), ("lenet", 32, "tf32"), ("lenet", 32, "bf16"), ("resnet18", 32, "bf16"), ("resnet50", 64, "bf16")]:
    counts.clear()
    torch.manual_seed(0)
    net = build_model(model, dtype=dt).cuda().train()
    opt = FusedSGD(net.parameters(), lr=0.01, model=net)
    x = torch.randn(16, 3, hw, hw, device="cuda"); y = torch.randint(0, 10, (16,), device="cuda")
    opt.zero_grad(); OF.cross_entropy(net(x), y).backward(); opt.step()
    torch.cuda.synchronize()
    print(model, dt, dict(counts))

import torch, torch.nn.functional as F
from distributed_ml_pytorch_b200 import nn as dnn
from distributed_ml_pytorch_b200.ops import functional as OF
import sys
sys.path.insert(0, 'tests')
from test_tf32_gpu import tf32_round, act4, rel_err
dev='cuda'
for rounded in (True, False):
  for seed in (1,2,3):
    torch.manual_seed(seed)
    cin,cout,k,s,p,hw,pool = 8,16,3,1,1,8,(3,2,1)
    conv = dnn.set_compute_dtype(dnn.Conv2d(cin, cout, k, stride=s, padding=p, activation="relu", pool=pool), "tf32").to(dev)
    R = tf32_round if rounded else (lambda t: t)
    with torch.no_grad(): conv.weight.copy_(R(conv.weight))
    x32 = R(torch.randn(8, cin, hw, hw, device=dev))
    x = act4(x32).requires_grad_(True)
    y = conv(x)
    xr = x32.clone().requires_grad_(True)
    w = conv.weight.detach().clone().requires_grad_(True)
    b = conv.bias.detach().clone().requires_grad_(True)
    with torch.backends.cudnn.flags(allow_tf32=False):
        z = F.relu(F.conv2d(xr, w, b, stride=s, padding=p))
        z.retain_grad()
        yr = F.max_pool2d(z, *pool)
        g = R(torch.randn_like(yr))
        y.backward(act4(g)); yr.backward(g)
    print(rounded, seed, 'fwd', rel_err(y,yr), 'w', rel_err(conv.weight.grad,w.grad), 'b', rel_err(conv.bias.grad,b.grad), 'x', rel_err(x.grad,xr.grad))

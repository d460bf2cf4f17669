import torch, torch.nn.functional as F
from distributed_ml_pytorch_b200 import nn as dnn
from distributed_ml_pytorch_b200.ops import functional as OF
import sys
sys.path.insert(0, 'tests')
from test_tf32_gpu import tf32_round, act4, rel_err
dev='cuda'
torch.manual_seed(1)
cin,cout,k,s,p,hw,pool = 8,16,3,1,1,8,(3,2,1)
conv = dnn.set_compute_dtype(dnn.Conv2d(cin, cout, k, stride=s, padding=p, activation="relu", pool=pool), "tf32").to(dev)
with torch.no_grad(): conv.weight.copy_(tf32_round(conv.weight))
x32 = tf32_round(torch.randn(8, cin, hw, hw, device=dev))
# kernel-side pieces, step by step
x = act4(x32)
yk = dnn.Conv2d.forward  # noqa
convnp = dnn.set_compute_dtype(dnn.Conv2d(cin, cout, k, stride=s, padding=p, activation="relu", pool=None), "tf32").to(dev)
with torch.no_grad():
    convnp.weight.copy_(conv.weight); convnp.bias.copy_(conv.bias)
zk = convnp(x).contiguous()                      # kernel's post-relu activation
with torch.backends.cudnn.flags(allow_tf32=False):
    zr = F.relu(F.conv2d(x32, conv.weight, conv.bias, stride=s, padding=p))
g = tf32_round(torch.randn(8, cout, 4, 4, device=dev))
zr1 = zr.detach().clone().requires_grad_(True); F.max_pool2d(zr1, *pool).backward(g)
zk1 = zk.detach().clone().requires_grad_(True); F.max_pool2d(zk1, *pool).backward(g)
print('max |zk-zr|', (zk-zr).abs().max().item())
d = (zr1.grad != zk1.grad)
print('ATen pool-bwd routing differs between kernel activation and reference activation at', int(d.sum()), 'elements')
# our kernel's pool backward on the kernel activation
zin = act4(zk.detach()).requires_grad_(True)
yo = OF.max_pool2d(zin, 3, 2, 1); yo.backward(act4(g))
d2 = (zin.grad.contiguous() != zk1.grad)
print('our pool-bwd vs ATen pool-bwd on the SAME activation differ at', int(d2.sum()), 'elements')
idx = d.nonzero()[:4]
for (b,c,h,w) in idx.tolist():
    h0,w0 = max(0,h-2), max(0,w-2)
    print('pos',b,c,h,w,'ref patch', zr[b,c,h0:h+3,w0:w+3].flatten().tolist())
    print('              ker patch', zk[b,c,h0:h+3,w0:w+3].flatten().tolist())

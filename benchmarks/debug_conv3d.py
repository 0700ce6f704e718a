import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch, torch.nn as nn, torch.nn.functional as F
from maavss_b200 import layers as L
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
def rel(a, b): return float((a - b).abs().max()) / max(float(b.abs().max()), 1e-12)
torch.manual_seed(0)
for cin, cout, hw, pad, pk in ((64, 64, 32, (1, 2, 2), 3), (32, 64, 64, (1, 2, 2), 2), (16, 32, 128, (1, 2, 2), 2), (64, 16, 10, (1, 3, 3), 3)):
    ref = nn.Conv3d(cin, cout, (3, 5, 5), 1, pad, bias=False).cuda()
    ours = L.Conv3d(cin, cout, (3, 5, 5), 1, pad, bias=False).cuda()
    ours.load_state_dict(ref.state_dict())
    x = torch.randn(2, cin, 8, hw, hw).cuda()
    xo, xr = x.clone().requires_grad_(), x.clone().requires_grad_()
    yo, yr = ours(xo), ref(xr)
    po, pr = L.MaxPool3d((1, pk, pk))(yo), F.max_pool3d(yr, (1, pk, pk))
    zo, zr = L.LeakyReLU()(po), F.leaky_relu(pr)
    dy = torch.randn_like(zr)
    zo.backward(dy); zr.backward(dy)
    print(cin, cout, hw, "fwd", rel(yo, yr), "pool", rel(zo, zr), "dx", rel(xo.grad, xr.grad), "dw", rel(ours.weight.grad, ref.weight.grad), flush=True)

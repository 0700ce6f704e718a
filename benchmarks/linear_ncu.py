"""The head Linear of the fc fusion model with the four accumulation windows stacked (M = 64, N = 32768, K = 512):
forward, input gradient, weight gradient, for `ncu --set full -k regex:"lin_"`; prints CUDA-event times without ncu."""
import json
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from maavss_b200 import layers as L  # noqa: E402

M, N, K = 64, 32768, 512
lin = L.Linear(K, N).cuda()
x = torch.randn(M, K, device="cuda", requires_grad=True)
dy = torch.randn(M, N, device="cuda")
for _ in range(3):
    y = lin(x)
    y.backward(dy)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    y = lin(x)
    y.backward(dy)
e1.record()
torch.cuda.synchronize()
print(json.dumps({"shape": [M, N, K], "fwd_bwd_us": round(e0.elapsed_time(e1) * 100, 1)}))

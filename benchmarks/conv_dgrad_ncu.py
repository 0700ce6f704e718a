"""Forward + backward of the headline model's heaviest strided 2-D layers (batched windows: N = 64) for
`ncu --set full -k regex:dgrad_kernel` / `regex:conv_igemm`."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from maavss_b200 import nn_ops as ops  # noqa: E402

for Cin, Cout, k, s, p, H, W in [(64, 64, (1, 1, 9), (1, 1, 2), (0, 0, 4), 8, 64), (32, 64, (1, 5, 5), (1, 2, 2), (0, 2, 2), 16, 32)]:
    x = torch.randn(64, Cin, 1, H, W, device="cuda", requires_grad=True)
    w = (torch.randn(Cout, Cin, *k, device="cuda") * 0.02).requires_grad_()
    for _ in range(2):
        y = ops.conv(x, w, None, s, p)
        y.backward(torch.randn_like(y))
    torch.cuda.synchronize()
print("ok")

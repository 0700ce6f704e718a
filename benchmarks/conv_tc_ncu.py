"""One forward + backward of each tensor-core Conv3d layer geometry (layers 2-4 of the Frames visual encoder) for
`ncu --set full -k regex:tc_conv`."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from maavss_b200 import nn_ops as ops  # noqa: E402

ops.TENSOR_CORE_CONV = True
B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
for Cin, Cout, HW in [(16, 32, 128), (32, 64, 64), (64, 64, 32)]:
    x = torch.randn(B, Cin, 8, HW, HW, device="cuda", requires_grad=True)
    w = (torch.randn(Cout, Cin, 3, 5, 5, device="cuda") * 0.02).requires_grad_()
    for _ in range(2):
        y = ops.conv(x, w, None, (1, 1, 1), (1, 2, 2))
        y.backward(torch.randn_like(y))
    torch.cuda.synchronize()
print("ok")

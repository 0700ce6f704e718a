import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
from maavss_b200 import nn_ops as ops
M, N, K = 64, 32768, 512
W = torch.randn(N, K, device="cuda"); x = torch.randn(M, K, device="cuda"); y = torch.empty(M, N, device="cuda")
for _ in range(3):
    ops.linear_fwd_into(x, W, None, y)
torch.cuda.synchronize()

"""tc_gemm_bf16 on the ViT-S/8 shapes of one 64-frame clip (M = 65600 tokens): time and TFLOP/s per shape / epilogue."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from maavss_b200 import vit_ops  # noqa: E402

M = 65600
dev = "cuda"
cases = [("qkv", 1152, 384, True, False, torch.bfloat16, 0), ("proj+res", 384, 384, True, True, torch.float32, 0),
         ("fc1+gelu", 1536, 384, True, False, torch.bfloat16, vit_ops.ACT_GELU), ("fc1 no act", 1536, 384, True, False, torch.bfloat16, 0),
         ("fc1 no bias/act", 1536, 384, False, False, torch.bfloat16, 0), ("fc2+res", 384, 1536, True, True, torch.float32, 0)]
for name, N, K, has_bias, has_res, odt, act in cases:
    a = (torch.randn(M, K, device=dev) * 0.5).bfloat16()
    b = (torch.randn(N, K, device=dev) * 0.05).bfloat16()
    bias = torch.randn(N, device=dev) if has_bias else None
    res = torch.randn(M, N, device=dev) if has_res else None
    out = torch.empty(M, N, device=dev, dtype=odt)
    for _ in range(3):
        vit_ops.gemm_bf16(a, b, bias=bias, residual=res, out=out, out_dtype=odt, act=act)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        vit_ops.gemm_bf16(a, b, bias=bias, residual=res, out=out, out_dtype=odt, act=act)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / 20 * 1e3
    print(json.dumps({"case": name, "M": M, "N": N, "K": K, "us": round(us, 1), "tflops": round(2.0 * M * N * K / us / 1e6, 1)}))

#!/usr/bin/env python
"""Micro-benchmark of the tcgen05 Conv3d kernel on the AV_Fusion_Model_Frames visual-encoder geometries
(avse_model_final.py:33-59, layers 2-4: k(3,5,5), pad (1,2,2), B clips x 8 frames): TFLOP/s of the forward and
data-gradient calls (prepass + main kernel), against the fp32 SIMT implicit GEMM and the measured bf16 peak."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from maavss_b200 import nn_ops as ops  # noqa: E402


def time_ms(fn, iters=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main(B=8, simt=True):
    try:
        pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"]
    except Exception:
        pk = 1590.0
    for name, Cin, Cout, HW in [("L2", 16, 32, 128), ("L3", 32, 64, 64), ("L4", 64, 64, 32)]:
        x = torch.randn(B, Cin, 8, HW, HW, device="cuda")
        w = torch.randn(Cout, Cin, 3, 5, 5, device="cuda") * 0.02
        dy = torch.randn(B, Cout, 8, HW, HW, device="cuda")
        gflop = 2.0 * B * 8 * HW * HW * Cout * Cin * 75 / 1e9
        row = {"layer": name, "B": B, "Cin": Cin, "Cout": Cout, "HW": HW, "gflop": round(gflop, 1)}
        for tc in ([True, False] if simt else [True]):
            ops.TENSOR_CORE_CONV = tc
            xo = x.clone().requires_grad_()
            y = ops.conv(xo, w, None, (1, 1, 1), (1, 2, 2))
            ms_f = time_ms(lambda: ops.conv(x, w, None, (1, 1, 1), (1, 2, 2)))
            ms_b = time_ms(lambda: torch.autograd.grad(y, xo, dy, retain_graph=True))
            tag = "tc" if tc else "simt"
            row[f"{tag}_fwd_ms"] = round(ms_f, 3)
            row[f"{tag}_fwd_tflops"] = round(gflop / ms_f, 1)
            row[f"{tag}_dgrad_ms"] = round(ms_b, 3)
            row[f"{tag}_dgrad_tflops"] = round(gflop / ms_b, 1)
            if tc:
                row["tc_fwd_frac_of_bf16_peak"] = round(gflop / ms_f / pk, 3)
        ops.TENSOR_CORE_CONV = False
        print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main(B=int(sys.argv[1]) if len(sys.argv) > 1 else 8, simt="--no-simt" not in sys.argv)

#!/usr/bin/env python
"""3xTF32 tensor-core GEMM (csrc/tc_gemm3.cu) on the Linear shapes of avse_model.AV_Fusion_Model with the accumulation windows
stacked (M = 64) and the Bi-LSTM projections, beside the fp32 SIMT kernels they replace: device time (CUDA events, L2 flushed
by rotating over weight copies > 126 MB), algorithmic GB/s (every operand once) against the measured HBM peak."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from maavss_b200 import nn_ops as ops

PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0


def time_ms(fn, n_sets, iters=20):
    """Device time per call with the host out of the way: n_sets calls (rotating weight copies) captured into one CUDA graph."""
    for i in range(min(3, n_sets)):
        fn(i % n_sets)
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for i in range(n_sets):
                fn(i)
    reps = max(1, iters // n_sets)
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (reps * n_sets)


def main():
    rows = []
    only = sys.argv[1].split(",") if len(sys.argv) > 1 else None
    only_ops = sys.argv[2].split(",") if len(sys.argv) > 2 else None
    iters = int(sys.argv[3]) if len(sys.argv) > 3 else 20
    for name, M, N, K in (("v_fc1", 64, 32768, 512), ("a_fc1", 64, 16384, 512), ("fc1", 64, 2048, 4096), ("fc2", 64, 512, 2048), ("lstm_ih", 512, 1024, 512)):
        if only and name not in only:
            continue
        n_sets = max(2, int(200e6 // (4 * N * K)) + 1)
        W = [torch.randn(N, K, device="cuda") * K ** -0.5 for _ in range(n_sets)]
        dW = [torch.zeros(N, K, device="cuda") for _ in range(n_sets)]
        x = torch.randn(M, K, device="cuda")
        dz = torch.randn(M, N, device="cuda")
        b = torch.randn(N, device="cuda")
        y, dx = torch.empty(M, N, device="cuda"), torch.empty(M, K, device="cuda")
        rec = {"layer": name, "M": M, "N": N, "K": K}
        for tc in (True, False):
            ops.TENSOR_CORE_LINEAR = tc
            tag = "tc" if tc else "simt"
            for op, fn, nbytes in (("fwd", lambda i: ops.linear_fwd_into(x, W[i], b, y, ops.ACT_LEAKY, 0.3), 4 * (M * K + N * K + M * N + N)),
                                   ("dgrad", lambda i: ops.linear_dgrad_into(dz, W[i], dx), 4 * (M * N + N * K + M * K)),
                                   ("wgrad", lambda i: ops.linear_wgrad_into(dz, x, dW[i], None, 1), 4 * (M * N + M * K + 2 * N * K))):
                if only_ops and (op not in only_ops or not tc):
                    continue
                ms = time_ms(fn, n_sets, iters)
                rec[f"{op}_{tag}_us"] = round(ms * 1e3, 1)
                rec[f"{op}_{tag}_frac"] = round(nbytes / ms / 1e6 / PEAK, 3)
        ops.TENSOR_CORE_LINEAR = True
        rows.append(rec)
        print(json.dumps(rec), flush=True)


if __name__ == "__main__":
    main()

import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
from maavss_b200 import nn_ops as ops
torch.manual_seed(0)
def run(P, Q, K, a_mn, b_mn, tr, small=False):
    A = torch.randn((K, P) if a_mn else (P, K)).cuda()
    B = torch.randn((K, Q) if b_mn else (Q, K)).cuda()
    shape = (Q, P) if tr else (P, Q)
    out = torch.full(shape, float("nan")).cuda()
    ok = ops.gemm_tf32x3(A, a_mn, A.shape[1], B, b_mn, B.shape[1], P, Q, K, out, shape[1], tr)
    torch.cuda.synchronize()
    Ad = (A.t() if a_mn else A).double().cpu(); Bd = (B.t() if b_mn else B).double().cpu()
    ref = Ad @ Bd.t()
    if tr: ref = ref.t()
    o = out.double().cpu()
    err = (o - ref).abs()
    nanc = int(torch.isnan(o).sum())
    scale = float(ref.abs().max())
    L = ops.lib()
    print(f"P{P} Q{Q} K{K} a_mn{a_mn} b_mn{b_mn} tr{tr}: ok={ok} splits_ws={L.maavss_gemm_tf32x3_workspace_bytes(P,Q,K)} nan={nanc} rel err {float(err[~torch.isnan(err)].max())/scale:.3e}", flush=True)
    if small:
        bad = (err > 1e-4 * scale) | torch.isnan(err)
        rows = bad.any(1).nonzero().flatten().tolist(); cols = bad.any(0).nonzero().flatten().tolist()
        print("   bad rows", rows[:40], "... n", len(rows), " bad cols", cols[:40], "... n", len(cols))
        r, c = (rows[0], cols[0]) if rows else (0, 0)
        print("   sample out/ref", o[r, c].item(), ref[r, c].item(), "ratio", (o[r, c] / ref[r, c]).item())
for cfg in [(128, 64, 32, 0, 0, 0), (128, 64, 32, 0, 0, 1), (128, 64, 8, 0, 0, 0), (128, 64, 64, 0, 0, 0), (128, 64, 128, 0, 0, 0), (256, 64, 32, 0, 0, 0),
            (128, 128, 32, 0, 0, 0), (128, 256, 32, 0, 0, 0), (128, 64, 32, 1, 0, 0), (128, 64, 32, 0, 1, 0), (128, 64, 32, 1, 1, 0),
            (1024, 512, 512, 0, 0, 1), (1, 64, 512, 0, 0, 1), (512, 64, 2048, 0, 0, 1), (32768, 64, 512, 0, 0, 1)]:
    run(*cfg, small=True)

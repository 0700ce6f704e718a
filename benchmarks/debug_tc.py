"""Bring-up probe for the tcgen05 kernels: prints per-case error against fp32 torch (not a test, not a benchmark)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden"))
import torch
from maavss_b200 import vit_ops


def rel(a, b):
    return float((a.float() - b.float()).abs().max()) / max(float(b.float().abs().max()), 1e-12)


torch.manual_seed(0)
for (M, N, K) in [(128, 128, 64), (128, 128, 128), (256, 256, 384), (300, 384, 384), (128, 64, 64)]:
    a = (torch.randn(M, K) * 0.5).cuda().bfloat16()
    b = (torch.randn(N, K) * K ** -0.5).cuda().bfloat16()
    try:
        out = vit_ops.gemm_bf16(a, b, out_dtype=torch.float32)
        torch.cuda.synchronize()
        ref = a.float() @ b.float().t()
        print("gemm", M, N, K, "rel", rel(out, ref), flush=True)
        if rel(out, ref) > 1e-3:
            d = (out - ref).abs()
            print("  bad rows", (d.max(1).values > 1e-3).nonzero().flatten()[:16].tolist(), "bad cols", (d.max(0).values > 1e-3).nonzero().flatten()[:16].tolist())
    except Exception as e:
        print("gemm", M, N, K, "FAILED", e, flush=True)
        sys.exit(1)
for (T, N, H) in [(1, 128, 1), (1, 256, 1), (1, 300, 2), (2, 1025, 6)]:
    qkv = torch.randn(T * N, 3 * H * 64).cuda().bfloat16()
    try:
        out = vit_ops.attention(qkv, T, N, H).float().view(T, N, H, 64)
        torch.cuda.synchronize()
        q, k, v = qkv.float().view(T, N, 3, H, 64).permute(2, 0, 3, 1, 4)
        p = ((q @ k.transpose(-2, -1)) * 0.125).softmax(-1)
        ref = (p @ v).permute(0, 2, 1, 3)
        print("attn", T, N, H, "rel", rel(out, ref), flush=True)
    except Exception as e:
        print("attn", T, N, H, "FAILED", e, flush=True)
        sys.exit(1)

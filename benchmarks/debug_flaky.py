"""Stress the AV_Fusion_Model visual / audio auto-encoder sub-forwards for run-to-run differences (a race or a read of
uninitialised memory shows up as a deviation far above fp32 summation-order noise) and name the first layer that deviates."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
from maavss_b200 import avse_model  # noqa: E402

torch.manual_seed(0)
m = avse_model.AV_Fusion_Model([2, 2, 64, 128], [2, 1, 8, 4096], 8)
m.train()
x_v = (torch.randn(2, 1, 8, 4096) * 0.3).cuda()
x_a = (torch.randn(2, 2, 64, 128) * 0.3).cuda()
outs = {}


def hook(name):
    def f(mod, inp, out):
        outs.setdefault(name, []).append(out.detach().clone())
    return f


for name, mod in m.named_modules():
    if name.count(".") == 1 and (name.startswith("phasegram_") or name.startswith("stft_")) and "autoencoder" not in name:
        mod.register_forward_hook(hook(name))
N = int(sys.argv[1]) if len(sys.argv) > 1 else 100
junk = []
for it in range(N):
    if it % 10 == 0:          # churn the caching allocator so recycled blocks hold garbage
        junk = [torch.randn(1 << 20, device="cuda") * 1e6 for _ in range(8)]
        junk = []
    m.visual_ae_forward(x_v)
    m.audio_ae_forward(x_a)
torch.cuda.synchronize()
bad = 0
for name, lst in outs.items():
    ref = lst[0]
    worst = max(float((t - ref).abs().max()) for t in lst)
    scale = float(ref.abs().max())
    flag = "  <-- DEVIATES" if worst > 1e-4 * max(scale, 1e-6) else ""
    bad += bool(flag)
    print(f"{name:28s} max|diff| over {len(lst)} runs = {worst:.3e}  (scale {scale:.3e}){flag}")
print("deviating layers:", bad)

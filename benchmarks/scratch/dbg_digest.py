"""Scratch: which digest rows differ between batched and sequential accumulation (and run-to-run)?"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import numpy as np, torch, synth
from oracle import models as om
from maavss_b200 import avse_model, optim, train_step
from test_gpu_models import param_digest

ref = om.AVFusionOracle([2, 2, 64, 128], [2, 1, 8, 4096], 8)
sd = synth.fill_state_dict(ref.state_dict(), seed=31)
xs, ys = (z.cuda() for z in synth.stft_batch(2, 96, 128, seed=37))
attn = synth.attn_clip(2, 12, 256, 256, seed=17).cuda()
res = []
for batched in (False, True, False, True):
    m = avse_model.AV_Fusion_Model([2, 2, 64, 128], [2, 1, 8, 4096], 8)
    m.load_state_dict(sd); m.train()
    opt = optim.Adam(m.parameters(), lr=1e-5)
    st = train_step.FusionTrainStep(m, opt, graph=False, batch_windows=batched)
    losses = [[float(v) for v in st(xs, ys, attn)] for _ in range(2)]
    # keep the gradient of a third accumulation (no optimizer step) for inspection
    opt.zero_grad(); st.micro_steps(xs, ys, attn)
    grads = {k: p.grad.detach().clone() for k, p in m.named_parameters() if p.grad is not None}
    res.append((losses, param_digest(m), m, grads))
names = [k for k, _ in res[0][2].named_parameters()]
names = [k for i, k in enumerate(names) if k not in names[:i]]
def cmp(a, b, tag):
    pa, pb = res[a][1], res[b][1]
    bad = np.argwhere(~np.isclose(pa, pb, rtol=1e-5, atol=1e-7))
    print(tag, "mismatches", len(bad))
    for r, c in bad:
        k = names[r]
        ga, gb = res[a][3].get(k), res[b][3].get(k)
        print("  ", k, "col", c, pa[r, c], pb[r, c], "numel", dict(res[a][2].named_parameters())[k].numel(),
              "grad absmax", float(ga.abs().max()) if ga is not None else None, "grad diff max", float((ga - gb).abs().max()) if ga is not None else None)
cmp(0, 1, "seq vs batched")
cmp(0, 2, "seq vs seq (run to run)")
cmp(1, 3, "batched vs batched (run to run)")

"""Which parameter gradients of AV_Fusion_Model_Frames deviate from the goldens (bring-up probe)."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "golden")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch, synth
from oracle import models as om
from maavss_b200 import avse_model_final
from test_gpu_models import grad_digest, micro_step_loss
g = np.load(os.path.join(ROOT, "tests", "golden", "frames.npz"))
xf, yf = synth.stft_batch(2, 64, 129, seed=43)
fr = synth.attn_clip(2, 8, 256, 256, seed=47)
ref = om.AVFusionFramesOracle([2, 2, 64, 129], [2, 1, 8, 256, 256], 8)
sd = synth.fill_state_dict(ref.state_dict(), seed=53)
m = avse_model_final.AV_Fusion_Model_Frames([2, 2, 64, 129], [2, 1, 8, 256, 256], 8)
m.load_state_dict(sd); m.train()
loss, a, v = micro_step_loss(m, xf.cuda(), fr.cuda(), yf[:, :, 32:40].cuda(), fr[:, :, 4].cuda())
loss.backward()
names, norms, samples = grad_digest(m)
for n, no, gn, s, gs in zip(names, norms, g["grad_norms"], samples, g["grad_samples"]):
    err = np.abs(s - gs).max() / max(np.abs(gs).max(), 1e-30)
    print(f"{n:40s} norm {no:.6e} ref {gn:.6e} sample relerr {err:.2e} maxref {np.abs(gs).max():.2e}")

"""Three-way gradient comparison for AV_Fusion_Model_Frames: ours vs oracle-on-CUDA (torch) vs CPU goldens."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "golden")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch, synth
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
from oracle import models as om
from test_gpu_models import grad_digest
g = np.load(os.path.join(ROOT, "tests", "golden", "frames.npz"))
xf, yf = synth.stft_batch(2, 64, 129, seed=43)
fr = synth.attn_clip(2, 8, 256, 256, seed=47)
ref = om.AVFusionFramesOracle([2, 2, 64, 129], [2, 1, 8, 256, 256], 8)
ref.load_state_dict(synth.fill_state_dict(ref.state_dict(), seed=53))
ref = ref.cuda(); ref.train()
loss, a, v = om.micro_step_loss(ref, xf.cuda(), fr.cuda(), yf[:, :, 32:40].cuda(), fr[:, :, 4].cuda())
loss.backward()
names, norms, samples = grad_digest(ref)
for n, no, gn, s, gs in list(zip(names, norms, g["grad_norms"], samples, g["grad_samples"]))[:16]:
    err = np.abs(s - gs).max() / max(np.abs(gs).max(), 1e-30)
    print(f"torch-cuda vs golden {n:32s} norm {no:.6e} ref {gn:.6e} sample relerr {err:.2e}")

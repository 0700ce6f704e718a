"""Small STFT/iSTFT launches for compute-sanitizer memcheck (all n_fft, a few hops, ragged lengths)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from maavss_b200 import audio_ops  # noqa: E402

cases = [(4096, 512, 160000, 3), (4096, 2048, 30001, 2), (2048, 512, 20000, 3), (1024, 256, 9000, 2), (512, 128, 5000, 3),
         (256, 66, 6336, 5), (128, 32, 700, 2), (64, 16, 500, 2)]
if len(sys.argv) > 1:
    cases = cases[:int(sys.argv[1])]
for n_fft, hop, n, B in cases:
    audio = torch.randn(B, n, device="cuda")
    w = torch.hamming_window(n_fft)
    for trim in (False, True):
        p = audio_ops.stft_planes(audio, w, n_fft, hop, trim)
        torch.cuda.synchronize()
        r = audio_ops.istft_audio(p, w, n_fft, hop, trim)
        torch.cuda.synchronize()
    print(n_fft, hop, n, "ok", float(p.abs().max()), float(r.abs().max()), flush=True)

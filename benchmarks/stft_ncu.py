"""A few STFT / iSTFT launches for `ncu --set full` (one launch per kernel variant, inputs > L2)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from maavss_b200 import audio_ops  # noqa: E402

clips, n = 512, 160000
audio = torch.randn(clips, n, device="cuda")
for n_fft, hop in [(512, 128), (2048, 512), (256, 66)]:
    w = torch.hamming_window(n_fft)
    T, F = audio_ops.stft_out_dims(n, n_fft, hop, False)
    planes = torch.empty(clips, 2, T, F, device="cuda")
    out = torch.empty(clips, hop * (T - 1), device="cuda")
    for _ in range(2):
        audio_ops.stft_planes(audio, w, n_fft, hop, False, out=planes)
        audio_ops.istft_audio(planes, w, n_fft, hop, False, out=out)
    torch.cuda.synchronize()

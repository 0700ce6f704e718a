import os, sys, torch
sys.path.insert(0, os.getcwd())
from maavss_b200 import audio_ops
n=160000
for clips in (8, 128, 1024):
    audio = torch.randn(clips, n, device="cuda")
    for n_fft, hop in [(4096,512),(4096,1024),(4096,2048)]:
        w = torch.hamming_window(n_fft)
        T, F = audio_ops.stft_out_dims(n, n_fft, hop, False)
        planes = torch.empty(clips, 2, T, F, device="cuda")
        audio_ops.stft_planes(audio, w, n_fft, hop, False, out=planes); torch.cuda.synchronize()
        print(clips, n_fft, hop, "fwd ok", flush=True)
        out = audio_ops.istft_audio(planes, w, n_fft, hop, False); torch.cuda.synchronize()
        print(clips, n_fft, hop, "inv ok", float((out[:, 4096:-4096]-audio[:, 4096:out.shape[1]-4096]*1.5863).abs().max()), flush=True)
        del planes, out

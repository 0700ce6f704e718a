"""Config 5: STFT / iSTFT bandwidth sweep (n_fft 512..4096 x hop n/8,n/4,n/2 + 256/66) on 10 s clips.
Algorithmic bytes: STFT 4*N + 8*F*T per clip, iSTFT 8*F*T + 4*hop*(T-1).  CUDA events on
the launching stream; inputs+outputs > L2 so HBM is measured."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from maavss_b200 import audio_ops  # noqa: E402

PEAK = 6463.3
try:
    PEAK = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass


def time_ms(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main(clips=1024, n=160000, torch_ref=True):
    audio = torch.randn(clips, n, device="cuda")
    rows = []
    cfgs = [(256, 66)] + [(nf, nf // d) for nf in (512, 1024, 2048, 4096) for d in (8, 4, 2)]
    for n_fft, hop in cfgs:
        w = torch.hamming_window(n_fft)
        T, F = audio_ops.stft_out_dims(n, n_fft, hop, False)
        B = clips
        while B * (8 * F * T) > 24e9:
            B //= 2
        a = audio[:B]
        planes = torch.empty(B, 2, T, F, device="cuda")
        out_a = torch.empty(B, hop * (T - 1), device="cuda")
        ms_f = time_ms(lambda: audio_ops.stft_planes(a, w, n_fft, hop, False, out=planes))
        ms_i = time_ms(lambda: audio_ops.istft_audio(planes, w, n_fft, hop, False, out=out_a))
        by_f = B * (4 * n + 8 * F * T)
        by_i = B * (8 * F * T + 4 * hop * (T - 1))
        row = dict(n_fft=n_fft, hop=hop, clips=B, stft_ms=round(ms_f, 3), stft_gbs=round(by_f / ms_f / 1e6, 1),
                   stft_frac=round(by_f / ms_f / 1e6 / PEAK, 3), istft_ms=round(ms_i, 3),
                   istft_gbs=round(by_i / ms_i / 1e6, 1), istft_frac=round(by_i / ms_i / 1e6 / PEAK, 3))
        if torch_ref:
            wd = w.cuda()
            ms_t = time_ms(lambda: torch.stft(a, n_fft, hop, n_fft, wd, center=True, pad_mode="reflect",
                                              return_complex=True), iters=3, warm=1)
            row["torch_stft_ms"] = round(ms_t, 3)
        rows.append(row)
        print(json.dumps(row), flush=True)
        del planes, out_a
    return rows


if __name__ == "__main__":
    main(clips=int(sys.argv[1]) if len(sys.argv) > 1 else 1024)

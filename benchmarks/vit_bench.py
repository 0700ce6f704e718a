#!/usr/bin/env python
"""BASELINE.json configs[2]: motion-attention front end, frame-diff + DINO ViT-S/8 attention maps over 64-frame clips.
Prints one JSON line: frames/s, achieved TFLOP/s against the measured bf16 peak, per-entry-point time shares."""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import torch
from maavss_b200 import _lib, video_attention

GFLOP_PER_FRAME = 58.1     # 11 full blocks + patch embed + last-block K projection (CLS-row shortcut)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=64)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--torch-ref", action="store_true", help="also time the oracle on CUDA through torch (library bar)")
    a = ap.parse_args()
    dev = torch.device("cuda")
    va = video_attention.VideoAttention(state_dict=video_attention.random_state_dict(seed=0))
    clips = [torch.randn(a.frames, 3, 256, 256, generator=torch.Generator().manual_seed(i)).to(dev) for i in range(3)]
    for i in range(3):
        va.clip_attention(clips[i % 3], attn_diff=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(a.iters):
        out = va.clip_attention(clips[i % 3], attn_diff=True)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.iters
    try:
        pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pk = {"bf16_tflops_sustained": 1400.0}
    tf = GFLOP_PER_FRAME * a.frames / ms        # GFLOP / ms == TFLOP/s
    _lib.start_profile()
    va.clip_attention(clips[0], attn_diff=True)
    agg = {}
    for name, args, t in _lib.stop_profile():
        agg[name] = agg.get(name, 0.0) + t
    tot = sum(agg.values())
    line = {"metric": "ViT-S/8 attention-map frames/s", "value": round(a.frames / ms * 1e3, 1), "unit": "frames/s", "frames": a.frames,
            "ms_per_clip": round(ms, 3), "achieved_tflops": round(tf, 1), "peak_tflops": pk["bf16_tflops_sustained"],
            "frac_of_bf16_peak": round(tf / pk["bf16_tflops_sustained"], 4), "dtype": "bf16 operands / fp32 accumulate",
            "shares": {k: round(v / tot, 4) for k, v in sorted(agg.items(), key=lambda kv: -kv[1])}, "profile_ms": round(tot, 3)}
    if a.torch_ref:
        from oracle import vit as ov
        sd = {k: v.to(dev) for k, v in video_attention.random_state_dict(seed=0).items()}
        with torch.no_grad():
            ov_last = lambda x: ov.last_selfattention(sd, x)
            for _ in range(2):
                ov_last(clips[0][:16])
            torch.cuda.synchronize()
            e0.record()
            for s in range(0, a.frames, 16):
                ov_last(clips[0][s:s + 16])
            e1.record()
            torch.cuda.synchronize()
        line["torch_eager_fp32_frames_per_s"] = round(a.frames / e0.elapsed_time(e1) * 1e3, 1)
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()

"""HBM roofline of the attention-frame cache kernels: pack fp32 -> uint8 / fp16 and the fused
dequantise + temporal diff + clip normalise, on a training batch of cached clips (256 x 12 x 256 x 256, config 4's
batch).  Algorithmic bytes: pack 4 + s per pixel, load s + 4 per pixel (s = 1 or 2 stored bytes; the max pass's second
read of the stored frames is overhead, not algorithmic).  Working set (201-403 MB stored + 805 MB fp32) exceeds the
126 MB L2.  Timed with CUDA events on the current stream; torch eager of the same arithmetic is the library bar.
One JSON line per (format, op)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from maavss_b200 import attn_cache  # noqa: E402


def timed(fn, reps=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def eager_load(stored, diff):
    x = stored.float().div(255) if stored.dtype == torch.uint8 else stored.float()
    if diff:
        x = torch.cat((torch.zeros_like(x[:, :1]), torch.diff(x, dim=1)), 1)
    return x * (1 / x.amax(dim=(1, 2, 3), keepdim=True))


def main():
    peaks = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    peak = json.load(open(peaks)).get("hbm_gbs", 6450.6) if os.path.exists(peaks) else 6450.6
    clips, T, H, W = 256, 12, 256, 256
    maps = torch.rand(clips, T, H, W, device="cuda")
    px = maps.numel()
    for fmt, s in (("uint8", 1), ("float16", 2)):
        ms = timed(lambda: attn_cache.pack(maps, fmt))
        stored = attn_cache.pack(maps, fmt)
        gbs = px * (4 + s) / ms / 1e6
        print(json.dumps({"op": "attn_cache_pack", "format": fmt, "pixels": px, "ms": round(ms, 4), "gbs": round(gbs, 1),
                          "frac": round(gbs / peak, 3)}), flush=True)
        for diff in (False, True):
            ms = timed(lambda: attn_cache.load(stored, diff))
            ems = timed(lambda: eager_load(stored, diff), reps=5, warm=2)
            gbs = px * (4 + s) / ms / 1e6
            print(json.dumps({"op": "attn_cache_load", "format": fmt, "diff": diff, "pixels": px, "ms": round(ms, 4),
                              "gbs": round(gbs, 1), "frac": round(gbs / peak, 3), "torch_eager_ms": round(ems, 4)}), flush=True)


if __name__ == "__main__":
    main()

"""video_phasegram at the headline shape (16 clips x 8 frames of 256x256 -> 64x64): CUDA-event time of the whole call
(4 rotating inputs, 134 MB > L2) and, under `ncu --set full -k regex:phasegram_frame`, one capture of the frame kernel."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from maavss_b200 import utilities  # noqa: E402

B, T, H, W = 16, 8, 256, 256
xs = [torch.rand(B, 1, T, H, W, device="cuda") for _ in range(4)]
for x in xs:
    utilities.video_phasegram(x, resize=(64, 64))
torch.cuda.synchronize()
if "--time" in sys.argv:
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for x in xs:
            utilities.video_phasegram(x, resize=(64, 64))
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / 200 * 1e3
    print(json.dumps({"call": "video_phasegram", "shape": [B, T, H, W], "us_per_call_in_graph": round(us, 2),
                      "algorithmic_GBps": round(4 * B * T * (H * W + 4096) / us / 1e3, 1)}))

"""N eager headline train steps (no CUDA graph) for ncu launch lists: python benchmarks/one_step_eager.py [steps]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402
from maavss_b200 import av_dataset, avse_model, optim, train_step  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
B = bench.B_PER_GPU
torch.manual_seed(0)
model = avse_model.AV_Fusion_Model([B, 2, 64, 128], [B, 1, 8, 4096], 8)
model.train()
opt = optim.Adam(model.parameters(), lr=bench.LR)
ds = av_dataset.AV_Dataset(num_frames=bench.CLIP_FRAMES, fft_len=bench.N_FFT, hops_per_frame=bench.HOPS_PER_FRAME,
                           noise_std=bench.NOISE_STD, framerate=bench.FPS, samplerate=bench.SR)
step = train_step.FusionTrainStep(model, opt, num_seq=bench.NUM_SEQ, num_frames=bench.NUM_FRAMES, hops_per_frame=bench.HOPS_PER_FRAME,
                                  p_size=64, loss_coeff=bench.LOSS_COEFF, graph=False)
audio, noise, attn = (t.cuda() for t in bench.synth_batch(torch, B, 1234))
for i in range(steps):
    x, y = ds.gen_stft_batch(audio, noise)
    out = step(x, y, attn)
torch.cuda.synchronize()
print("loss", float(out[0]))

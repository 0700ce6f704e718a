#!/usr/bin/env python
"""BASELINE.json configs[3]: avse_model_final.AV_Fusion_Model_Frames full AV training step (train_avse_frames.py:150-181),
data-parallel over N GPUs (torchrun) with the NCCL gradient all-reduce; B clips per GPU (256 global at B=32, N=8).

One step = 4 accumulation windows (8 raw 256x256 attention frames + 64 STFT frames each) forward + backward, gradient
all-reduce, Adam.  Prints one JSON line: clips/s, ms/step, achieved TFLOP/s of the model arithmetic (78.97 GFLOP per
clip-window forward + backward) and the per-entry-point shares of one eager step.

    python benchmarks/frames_bench.py --clips 8 --steps 3
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 benchmarks/frames_bench.py --clips 8
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

GFLOP_PER_CLIP_WINDOW = 78.97


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--clips", type=int, default=8, help="clips per GPU")
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--breakdown", default=None)
    ap.add_argument("--tc", action="store_true", help="stride-1 Conv3d forward / data gradient on the tcgen05 kernel (bf16 operands)")
    a = ap.parse_args()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    import torch
    import torch.distributed as dist
    from maavss_b200 import _lib, avse_model_final, nn_ops, optim, parallel, train_step
    nn_ops.TENSOR_CORE_CONV = bool(a.tc)

    rank, world = parallel.init()
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    B = a.clips
    torch.manual_seed(0)
    model = avse_model_final.AV_Fusion_Model_Frames([B, 2, 64, 129], [B, 1, 8, 256, 256], 8, latent_channels=16, fc_size=4096)
    parallel.broadcast_module(model)
    model.train()
    opt = optim.Adam(model.parameters(), lr=1e-5)
    step = train_step.FramesTrainStep(model, opt, num_seq=4, num_frames=8, hops_per_frame=8, loss_coeff=0.001, graph=False, tensor_core=bool(a.tc))
    g = torch.Generator().manual_seed(1234 + rank)
    y = torch.randn(B, 2, 96, 129, generator=g).to(dev)
    x = y + 0.1 * torch.randn(B, 2, 96, 129, generator=g).to(dev)
    attn = torch.rand(B, 1, 12, 256, 256, generator=g).to(dev)
    for _ in range(a.warmup):
        out = step(x, y, attn)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        out = step(x, y, attn)
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / a.steps], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms)
    _lib.start_profile()
    step(x, y, attn)
    agg = {}
    for name, args, t in _lib.stop_profile():
        e = agg.setdefault(name, [0.0, 0])
        e[0] += t
        e[1] += 1
    tot = sum(v[0] for v in agg.values())
    try:
        pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pk = {"bf16_tflops_sustained": 1400.0}
    tf = GFLOP_PER_CLIP_WINDOW * 4 * B / ms      # GFLOP / ms == TFLOP/s, per GPU
    line = {"metric": "AV_Fusion_Model_Frames train clips/sec", "value": round(world * B / (ms / 1e3), 2), "unit": "clips/s",
            "n_gpus": world, "clips_per_gpu": B, "global_batch": world * B, "ms_per_step": round(ms, 2), "steps": a.steps,
            "dtype": "bf16 operands / f32 accumulate on tcgen05 for the Conv3d layers 2-4 (fwd, dgrad, wgrad), f32 elsewhere" if a.tc else "f32 (SIMT implicit-GEMM convolutions)", "achieved_tflops_per_gpu": round(tf, 2),
            "frac_of_bf16_tensor_peak": round(tf / pk["bf16_tflops_sustained"], 4), "execution": "eager", "sync_bn": world > 1,
            "final_loss": float(out[0]),
            "shares": {k: round(v[0] / tot, 4) for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:10]}}
    if a.breakdown and rank == 0:
        with open(a.breakdown, "w") as f:
            for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0]):
                f.write(f"{100 * v[0] / tot:6.2f}%  calls={v[1]:4d}  total_ms={v[0]:9.3f}  {k}\n")
    if rank == 0:
        os.write(real_stdout, (json.dumps(line) + "\n").encode())
    if world > 1:
        torch.cuda.synchronize()
        dist.barrier()
        os._exit(0)


if __name__ == "__main__":
    main()

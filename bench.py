#!/usr/bin/env python
"""Benchmark of the MAAVSS training hot path on B200 -- ONE JSON line carrying the whole BASELINE.json metric
("train clips/sec at 1/2/4/8 B200; STFT GB/s and attn TC util vs peak").

Headline (`value`, `e2e`, `roofline`) = BASELINE.json configs[1]: "avse_model train step, batch 16 synthetic audio+frames, 1 B200".
One step = the hot path over one batch of 16 synthetic clips per GPU:
    audio front end   STFT (n_fft 256, hop 66) -> real/imag planes -> 1/max|.| -> + noise      (av_dataset.py:335-342)
    4 micro-steps     video_phasegram of 8 attention frames -> AV_Fusion_Model forward -> MSE(a) + 0.001 MSE(v)
                      -> backward, gradients accumulated                                         (train.py:136-159)
    optimizer         [N > 1: NCCL all-reduce of the flat gradient arena] -> Adam(lr 1e-5)       (train.py:161-162)
"frames" are the per-clip motion-attention frames (12 x 256 x 256) the dataset hands to the loop.

Sub-records on the same line, each with its own clock sample and >= 20 timed iterations:
    "frames"  configs[3]: avse_model_final.AV_Fusion_Model_Frames train step (tcgen05 Conv3d), 32 clips / GPU, CUDA graph,
              data parallel at N > 1 (global batch 256 at N = 8)                                 (train_avse_frames.py:150-181)
    "vit"     configs[2]: frame-diff + DINO ViT-S/8 attention maps over a 64-frame clip           (video_attention.py:38-103)   [N = 1]
    "stft"    configs[4]: STFT / iSTFT bandwidth sweep, n_fft 512-4096 x hop n/8,n/4,n/2 (+256/66), 1024 x 10 s clips           [N = 1]
    "torch_eager_b200"    the same two train steps through PyTorch eager (cuDNN / cuBLAS, TF32 off) on the same GPU            [N = 1]
    "pipeline"            the headline step fed from RAW inputs: audio memmap + uint8 frames -> H2D -> ViT -> STFT -> train step      [N = 1]
    "cpu_baseline"        the oracle port of the headline step on the box's host cores                                          [N = 1]

    python bench.py --gpus N --steps K --warmup W            our arm  (torchrun for N > 1, weak scaling: 16 clips / GPU)
    python bench.py --impl reference ...                     the reference's CPU path (oracle port) on the host cores
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))

B_PER_GPU = 16
N_FFT, HOPS_PER_FRAME, NUM_FRAMES, NUM_SEQ, FPS, SR, FRAME = 256, 8, 8, 4, 30, 16000, 256
CLIP_FRAMES = NUM_FRAMES + NUM_SEQ          # 12 (train.py:33)
NOISE_STD, LOSS_COEFF, LR = 0.1, 0.001, 1e-5
HOP = int((SR / FPS) / HOPS_PER_FRAME)
WORKLOAD = "avse_model.AV_Fusion_Model train step, 16 synthetic clips/GPU (audio 6336 samples + 12x256x256 attention frames)"
FRAMES_B_PER_GPU = 32                       # configs[3]: global batch 256 over 8 GPUs
GFLOP_PER_FRAMES_CLIP = 4 * 78.97           # 78.97 GFLOP forward + backward per clip-window, 4 windows per clip
GFLOP_PER_VIT_FRAME = 58.1                  # 11 full blocks + patch embed + last-block K projection


def config_record(world):
    """Identical in both arms (the driver compares the two `config` objects)."""
    return {"workload": WORKLOAD, "clips_per_gpu": B_PER_GPU, "num_seq": NUM_SEQ, "num_frames": NUM_FRAMES, "n_fft": N_FFT, "hop": HOP,
            "frame": f"{CLIP_FRAMES}x{FRAME}x{FRAME}", "optimizer": f"Adam lr {LR}", "l2": "inputs rotate over 4 batches (208 MB) + 147 MB "
            "weights + 3 x 147 MB optimizer arenas per step > 126 MB L2"}


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


def synth_batch(torch, B, seed):
    g = torch.Generator().manual_seed(seed)
    n = HOPS_PER_FRAME * HOP * CLIP_FRAMES
    audio = torch.randn(B, n, generator=g) * 0.1
    noise = torch.randn(B, 2, n // HOP, N_FFT // 2, generator=g)
    attn = torch.rand(B, 1, CLIP_FRAMES, FRAME, FRAME, generator=g)
    return audio, noise, attn


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING a timed region (B200_PROFILING.md)."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self, min_s=0.0):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.thread.join(timeout=2)
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].startswith("Active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def hold_for_clocks(torch, fn, seconds):
    """Keep `fn` running for about `seconds` so that the 100 ms nvidia-smi sampler sees the kernel under load."""
    t0 = time.perf_counter()
    while time.perf_counter() - t0 < seconds:
        fn()
        torch.cuda.synchronize()


# ----------------------------------------------------------------------------------------------- reference arm
def oracle_step_factory(torch, B, device="cpu"):
    """The reference's path for this step, restated in oracle/ (the reference is pure Python and not on this box)."""
    from oracle import audio as oa
    from oracle import models as om
    from oracle import phasegram as op
    torch.manual_seed(0)
    model = om.AVFusionOracle([B, 2, 64, 128], [B, 1, 8, 4096], 8).to(device)
    model.train()
    opt = torch.optim.Adam(model.parameters(), lr=LR)
    hop = oa.calc_hop_size(NUM_FRAMES, HOPS_PER_FRAME, FPS, SR)[0]
    window = oa.hamming_window(N_FFT).to(device)

    def step(audio, noise, attn):
        y = torch.stack([oa.stft_planes(audio[b], window, N_FFT, hop, True, True, normalize_output=True) for b in range(B)])
        x = oa.add_noise(y, noise, NOISE_STD)
        loss = None
        for j in range(NUM_SEQ):
            pg = op.video_phasegram(attn[:, :, j:j + NUM_FRAMES], resize=(64, 64))
            s0, s1 = HOPS_PER_FRAME * j, HOPS_PER_FRAME * (j + NUM_FRAMES)
            loss, _, _ = om.micro_step_loss(model, x[:, :, s0:s1], pg, y[:, :, s0:s1], pg, LOSS_COEFF, NUM_SEQ)
            loss.backward()
        opt.step()
        opt.zero_grad()
        return loss.detach()
    return step


def oracle_frames_step_factory(torch, B, device):
    from oracle import models as om
    torch.manual_seed(0)
    model = om.AVFusionFramesOracle([B, 2, 64, 129], [B, 1, 8, 256, 256], 8).to(device)
    model.train()
    opt = torch.optim.Adam(model.parameters(), lr=LR)
    mid = NUM_FRAMES // 2

    def step(x_stft, y_stft, attn):
        loss = None
        for j in range(NUM_SEQ):
            s0 = HOPS_PER_FRAME * j
            m0 = s0 + HOPS_PER_FRAME * mid
            loss, _, _ = om.micro_step_loss(model, x_stft[:, :, s0:s0 + HOPS_PER_FRAME * NUM_FRAMES], attn[:, :, j:j + NUM_FRAMES],
                                            y_stft[:, :, m0:m0 + HOPS_PER_FRAME], attn[:, :, j + mid], LOSS_COEFF, NUM_SEQ)
            loss.backward()
        opt.step()
        opt.zero_grad()
        return loss.detach()
    return step


def time_cpu(torch, B, steps, warmup, budget_s=None):
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    step = oracle_step_factory(torch, B)
    batch = synth_batch(torch, B, 1234)
    for _ in range(warmup):
        step(*batch)
    t0, done = time.perf_counter(), 0
    while done < steps:
        step(*batch)
        done += 1
        if budget_s is not None and time.perf_counter() - t0 > budget_s:
            break
    dt = time.perf_counter() - t0
    return B * done / dt, dt / done * 1e3, threads, done


def run_reference(args):
    import torch
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cps, ms, threads, done = time_cpu(torch, B_PER_GPU, args.steps, args.warmup, budget_s=240)
    line = {"impl": "reference", "metric": "train clips/sec", "value": round(cps, 3), "unit": "clips/s", "n_gpus": args.gpus,
            "steps": done, "warmup": args.warmup, "ms_per_step": round(ms, 3), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_record(args.gpus), "device": "host CPU (rank 0 only)",
            "cpu_baseline": {"value": round(cps, 3), "unit": "clips/s", "cores": threads, "kind": "port",
                             "sample": f"{done} full steps of {B_PER_GPU} clips (same step as the GPU arm), torch CPU, all host threads"},
            "e2e": {"value": round(cps, 3), "unit": "clips/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


# ----------------------------------------------------------------------------------------------------- our arm
def algorithmic_bytes(name, a):
    """Algorithmic HBM bytes of one C-ABI call (DESIGN.md 'Kernels'): every operand read or written once, fp32."""
    iv = lambda x: int(getattr(x, "value", x))
    if name == "maavss_linear_fwd":
        M, N, K = iv(a[4]), iv(a[5]), iv(a[6])
        return 4 * (M * K + N * K + M * N + N), f"M{M} N{N} K{K}"
    if name == "maavss_linear_dgrad":
        M, N, K = iv(a[3]), iv(a[4]), iv(a[5])
        return 4 * (M * N + N * K + M * K), f"M{M} N{N} K{K}"
    if name == "maavss_linear_wgrad":
        M, N, K, acc = iv(a[4]), iv(a[5]), iv(a[6]), iv(a[7])
        return 4 * (M * N + M * K + (2 if acc else 1) * N * K), f"M{M} N{N} K{K}"
    if name in ("maavss_conv_fwd", "maavss_conv_dgrad", "maavss_conv_wgrad"):
        d = a[0]._obj
        x = d.N * d.Cin * d.D * d.H * d.W
        y = d.N * d.Cout * d.OD * d.OH * d.OW
        w = d.Cout * d.Cin * d.KD * d.KH * d.KW
        tag = f"N{d.N} C{d.Cin}->{d.Cout} k{d.KD}x{d.KH}x{d.KW} s{d.SH}x{d.SW} in{d.D}x{d.H}x{d.W}"
        return 4 * (x + y + (2 * w if name.endswith("wgrad") else w)), tag
    if name in ("maavss_conv_tc3_fwd", "maavss_conv_tc3_dgrad", "maavss_conv_tc3_wgrad"):
        # the layer's algorithmic traffic (x, y, w once each), NOT the im2col matrix the path moves through L2
        d = a[0]._obj
        x = d.N * d.Cin * d.D * d.H * d.W
        y = d.N * d.Cout * d.OD * d.OH * d.OW
        w = d.Cout * d.Cin * d.KD * d.KH * d.KW
        tag = f"N{d.N} C{d.Cin}->{d.Cout} k{d.KH}x{d.KW} s{d.SH}x{d.SW} in{d.H}x{d.W}"
        return 4 * (x + y + (2 * w if name.endswith("wgrad") else w)), tag
    if name == "maavss_adam_step":
        return 28 * iv(a[4]), f"n{iv(a[4])}"
    if name == "maavss_adam_apply":
        return (32 if iv(a[11]) else 28) * iv(a[4]), f"n{iv(a[4])}"
    if name in ("maavss_bn_apply_act", "maavss_bn_bwd_apply", "maavss_bn_reduce", "maavss_bn_bwd_reduce"):
        k = {"maavss_bn_apply_act": (4, 2), "maavss_bn_bwd_apply": (6, 4), "maavss_bn_reduce": (1, 1), "maavss_bn_bwd_reduce": (4, 3)}[name]
        N, C, inner = iv(a[k[0]]), iv(a[k[0] + 1]), iv(a[k[0] + 2])
        return 4 * k[1] * N * C * inner, f"N{N} C{C} inner{inner}"
    if name == "maavss_channel_sum":
        N, C, inner = iv(a[1]), iv(a[2]), iv(a[3])
        return 4 * (N * C * inner + C), f"N{N} C{C} inner{inner}"
    if name in ("maavss_bn_reduce_grouped", "maavss_bn_apply_grouped", "maavss_bn_bwd_reduce_grouped", "maavss_bn_bwd_apply_grouped"):
        # tensors of N*C*inner floats touched once each: x | x, y | dy, y, x | dy, y, x, dx
        k = {"maavss_bn_reduce_grouped": (1, 1), "maavss_bn_apply_grouped": (2, 2), "maavss_bn_bwd_reduce_grouped": (4, 3),
             "maavss_bn_bwd_apply_grouped": (7, 4)}[name]
        N, C, inner = iv(a[k[0]]), iv(a[k[0] + 1]), iv(a[k[0] + 2])
        return 4 * k[1] * N * C * inner, f"N{N} C{C} inner{inner}"
    if name in ("maavss_lstm_step_fwd", "maavss_lstm_step_bwd"):
        B, T, H = (iv(a[i]) for i in ((6, 7, 8) if name.endswith("fwd") else (7, 8, 9)))
        # per direction: W_hh (4H x H), one time slice of gin / gates (B x 4H each), cells, h (B x H each)
        return 4 * 2 * (4 * H * H + 2 * B * 4 * H + 3 * B * H), f"B{B} T{T} H{H}"
    if name == "maavss_gemm_tf32x3":
        P, Q, K = iv(a[6]), iv(a[7]), iv(a[8])
        acc = iv(a[16])
        return 4 * (P * K + Q * K + (2 if acc else 1) * P * Q), f"P{P} Q{Q} K{K} a_mn{iv(a[1])} b_mn{iv(a[4])}"
    if name == "maavss_phasegram_frames":
        B, T, H, W, R = (iv(a[i]) for i in range(3, 8))
        return 4 * B * T * (H * W + R * R), f"B{B} T{T} {H}x{W}->{R}"
    return None, ""


def headline_algorithmic_bytes(n_arena, n_used, B):
    """Floor of the HBM traffic of one headline step (DESIGN.md 'Measurement'): every trained weight read once forward and
    once for the data gradient, its gradient written once (the 4 windows are batched, so once per step, not per window),
    Adam's 28 B/param over the arena + the zero_grad fill, the attention frames and STFT planes read once."""
    weights = 4 * n_used * 3                      # read fwd, read dgrad, write grad
    optimizer = 28 * n_arena + 4 * n_arena         # Adam + zero_grad
    inputs = 4 * B * (CLIP_FRAMES * FRAME * FRAME + HOPS_PER_FRAME * HOP * CLIP_FRAMES + 3 * 2 * 96 * 128)
    return weights + optimizer + inputs, {"weights_and_grads": weights, "adam_and_zero_grad": optimizer, "inputs": inputs}


_REAL_STDOUT = None


def emit(line):
    """The ONE JSON line goes to the process's original stdout; everything else (NCCL's version banner, library
    chatter) was redirected to stderr by quiet_stdout()."""
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def quiet_stdout():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)


class Ctx:
    pass


def timed_steps(ctx, fn, steps):
    """EXACTLY `steps` calls bracketed by barrier + synchronize on both sides, CUDA events, MAX over ranks -> total ms."""
    torch, dist = ctx.torch, ctx.dist
    def barrier():
        if ctx.world > 1:
            dist.barrier()
        torch.cuda.synchronize()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        fn(i)
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device=ctx.dev)
    if ctx.world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms)


def leg_frames(ctx, steps):
    """configs[3]: avse_model_final full AV training (train_avse_frames.py:150-181), 32 clips per GPU, the Conv3d layers on
    tcgen05, one CUDA graph per optimizer step (4 windows fwd + bwd, gradient all-reduce, Adam); data parallel at N > 1."""
    torch = ctx.torch
    from maavss_b200 import avse_model_final, optim, parallel, train_step
    B = FRAMES_B_PER_GPU
    torch.manual_seed(0)
    model = avse_model_final.AV_Fusion_Model_Frames([B, 2, 64, 129], [B, 1, 8, 256, 256], 8, latent_channels=16, fc_size=4096)
    parallel.broadcast_module(model)
    model.train()
    opt = optim.Adam(model.parameters(), lr=LR)
    step = train_step.FramesTrainStep(model, opt, num_seq=NUM_SEQ, num_frames=NUM_FRAMES, hops_per_frame=HOPS_PER_FRAME,
                                      loss_coeff=LOSS_COEFF, graph=not ctx.args.eager, tensor_core=True)
    sets = []
    for i in range(2):                              # 2 x (32 x 12 x 256 x 256 x 4 B = 101 MB of frames) + 288 MB weights > L2
        g = torch.Generator().manual_seed(4321 + 97 * ctx.rank + i)
        y = torch.randn(B, 2, 96, 129, generator=g)
        x = y + 0.1 * torch.randn(B, 2, 96, 129, generator=g)
        attn = torch.rand(B, 1, CLIP_FRAMES, FRAME, FRAME, generator=g)
        sets.append(tuple(t.to(ctx.dev) for t in (x, y, attn)))
    out = None
    for i in range(3):
        out = step(*sets[i % 2])
    torch.cuda.synchronize()
    clocks = ClockSampler(ctx.local).start()
    ms = timed_steps(ctx, lambda i: step(*sets[i % 2]), steps)
    clk = clocks.stop()
    pk, _ = peaks()
    per_step = ms / steps
    tf = GFLOP_PER_FRAMES_CLIP * B / per_step                     # GFLOP / ms == TFLOP/s, per GPU
    rec = {"metric": "AV_Fusion_Model_Frames train clips/sec", "value": round(ctx.world * B / (per_step / 1e3), 2), "unit": "clips/s",
           "n_gpus": ctx.world, "clips_per_gpu": B, "global_batch": ctx.world * B, "steps": steps, "ms_per_step": round(per_step, 3),
           "scaling": "weak", "execution": "eager" if ctx.args.eager else "cuda-graph",
           "dtype": "bf16 operands / f32 accumulate on tcgen05 for Conv3d layers 2-4 (fwd, dgrad, wgrad); f32 elsewhere",
           "gflop_per_clip": GFLOP_PER_FRAMES_CLIP, "achieved_tflops_per_gpu": round(tf, 2), "peak_tflops": pk["bf16_tflops_sustained"],
           "frac_of_bf16_sustained": round(tf / pk["bf16_tflops_sustained"], 4), "sync_bn": ctx.world > 1,
           "final_loss": float(out[0]), "clocks": clk,
           "l2": "2 rotating input sets (2 x 101 MB frames) + 288 MB weights + optimizer arenas > 126 MB L2"}
    step._graph = None
    del step, model, opt, sets
    torch.cuda.empty_cache()
    return rec


def leg_vit(ctx, iters):
    """configs[2]: frame-diff + DINO ViT-S/8 attention maps over 64-frame clips (video_attention.py:38-103, av_dataset.py:323-328)."""
    torch = ctx.torch
    from maavss_b200 import _lib, video_attention
    frames = 64
    va = video_attention.VideoAttention(state_dict=video_attention.random_state_dict(seed=0))
    clips = [torch.randn(frames, 3, 256, 256, generator=torch.Generator().manual_seed(i)).to(ctx.dev) for i in range(3)]   # 3 x 50 MB
    fn = lambda i: va.clip_attention(clips[i % 3], attn_diff=True)
    for i in range(3):
        fn(i)
    torch.cuda.synchronize()
    L = _lib.lib()
    c0 = L.maavss_launch_count()
    fn(0)
    launches = L.maavss_launch_count() - c0
    clocks = ClockSampler(ctx.local).start()
    hold_for_clocks(torch, lambda: fn(0), 0.4)
    ms = timed_steps(ctx, fn, iters) / iters
    clk = clocks.stop()
    pk, _ = peaks()
    tf = GFLOP_PER_VIT_FRAME * frames / ms
    _lib.start_profile()
    fn(0)
    agg = {}
    for name, _, t in _lib.stop_profile():
        agg[name] = agg.get(name, 0.0) + t
    tot = sum(agg.values())
    rec = {"metric": "ViT-S/8 attention-map frames/sec (frame diff + clip normalise fused on device)", "value": round(frames / ms * 1e3, 1),
           "unit": "frames/s", "frames_per_clip": frames, "iters": iters, "ms_per_clip": round(ms, 3), "gflop_per_frame": GFLOP_PER_VIT_FRAME,
           "achieved_tflops": round(tf, 1), "peak_tflops": pk["bf16_tflops_sustained"], "frac_of_bf16_sustained": round(tf / pk["bf16_tflops_sustained"], 4),
           "dtype": "bf16 operands / f32 accumulate (tcgen05), f32 residual stream", "gpu_launches_per_clip": int(launches), "clocks": clk,
           "l2": "3 rotating 64-frame clips (3 x 50 MB) + 43 MB bf16 weights + activations (qkv 151 MB) > 126 MB L2",
           "shares": {k: round(v / tot, 4) for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:6]}}
    del va, clips
    torch.cuda.empty_cache()
    return rec


def leg_stft(ctx, iters, clips=1024, n=160000):
    """configs[4]: STFT / iSTFT bandwidth sweep on 10-second clips.  Algorithmic bytes per clip:
    STFT 4 N + 8 F T, iSTFT 8 F T + 4 hop (T - 1); inputs + outputs of every point are > 2 GB, far beyond L2."""
    torch = ctx.torch
    from maavss_b200 import audio_ops
    pk, _ = peaks()
    peak = pk["hbm_gbs"]
    audio = torch.randn(clips, n, device=ctx.dev)
    cfgs = [(256, 66)] + [(nf, nf // d) for nf in (512, 1024, 2048, 4096) for d in (8, 4, 2)]
    rows = []
    clocks = ClockSampler(ctx.local).start()
    for n_fft, hop in cfgs:
        w = torch.hamming_window(n_fft)
        T, F = audio_ops.stft_out_dims(n, n_fft, hop, False)
        B = clips
        while B * (8 * F * T) > 24e9:
            B //= 2
        a = audio[:B]
        planes = torch.empty(B, 2, T, F, device=ctx.dev)
        out_a = torch.empty(B, hop * (T - 1), device=ctx.dev)
        f_fwd = lambda i: audio_ops.stft_planes(a, w, n_fft, hop, False, out=planes)
        f_inv = lambda i: audio_ops.istft_audio(planes, w, n_fft, hop, False, out=out_a)
        for f in (f_fwd, f_inv):
            for i in range(3):
                f(i)
        ms_f = timed_steps(ctx, f_fwd, iters) / iters
        ms_i = timed_steps(ctx, f_inv, iters) / iters
        by_f = B * (4 * n + 8 * F * T)
        by_i = B * (8 * F * T + 4 * hop * (T - 1))
        rows.append({"n_fft": n_fft, "hop": hop, "clips": B, "stft_ms": round(ms_f, 4), "stft_gbs": round(by_f / ms_f / 1e6, 1),
                     "stft_frac": round(by_f / ms_f / 1e6 / peak, 4), "istft_ms": round(ms_i, 4), "istft_gbs": round(by_i / ms_i / 1e6, 1),
                     "istft_frac": round(by_i / ms_i / 1e6 / peak, 4)})
        del planes, out_a
    clk = clocks.stop()
    torch.cuda.empty_cache()
    summ = lambda key: {"min": min(r[key] for r in rows), "median": round(statistics.median(r[key] for r in rows), 4),
                        "max": max(r[key] for r in rows)}
    ref = next(r for r in rows if r["n_fft"] == 256)
    return {"metric": "STFT / iSTFT achieved HBM GB/s (algorithmic bytes / CUDA-event time)", "unit": "GB/s", "clips": clips, "samples_per_clip": n,
            "iters": iters, "peak": peak, "bound": "hbm", "stft_frac": summ("stft_frac"), "istft_frac": summ("istft_frac"),
            "stft_gbs": summ("stft_gbs"), "istft_gbs": summ("istft_gbs"),
            "reference_geometry_256_66": {"stft_frac": ref["stft_frac"], "istft_frac": ref["istft_frac"]}, "sweep": rows, "clocks": clk}


def leg_torch_eager(ctx):
    """The library bar: the same two train steps through PyTorch eager on this B200 -- cuFFT / cuDNN / cuBLAS,
    TF32 off, fp32 -- running the oracle modules on CUDA.  Baseline only; nothing of it is on the product path."""
    torch = ctx.torch
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.benchmark = True
    out = {}
    B = B_PER_GPU
    step = oracle_step_factory(torch, B, ctx.dev)
    sets = [tuple(t.to(ctx.dev) for t in synth_batch(torch, B, 1234 + i)) for i in range(4)]
    for i in range(3):
        step(*sets[i % 4])
    ms = timed_steps(ctx, lambda i: step(*sets[i % 4]), 20) / 20
    out["fusion_step"] = {"value": round(B / ms * 1e3, 1), "unit": "clips/s", "ms_per_step": round(ms, 3), "clips": B, "steps": 20}
    del step, sets
    torch.cuda.empty_cache()
    B = FRAMES_B_PER_GPU
    fstep = oracle_frames_step_factory(torch, B, ctx.dev)
    g = torch.Generator().manual_seed(4321)
    y = torch.randn(B, 2, 96, 129, generator=g).to(ctx.dev)
    x = y + 0.1 * torch.randn(B, 2, 96, 129, generator=g).to(ctx.dev)
    attn = torch.rand(B, 1, CLIP_FRAMES, FRAME, FRAME, generator=g).to(ctx.dev)
    for _ in range(2):
        fstep(x, y, attn)
    ms = timed_steps(ctx, lambda i: fstep(x, y, attn), 5) / 5
    out["frames_step"] = {"value": round(B / ms * 1e3, 1), "unit": "clips/s", "ms_per_step": round(ms, 2), "clips": B, "steps": 5}
    out["how"] = "oracle.models on CUDA through torch eager, fp32, allow_tf32 False, cudnn.benchmark True; same synthetic shapes"
    del fstep, x, y, attn
    torch.cuda.empty_cache()
    return out


def leg_pipeline(ctx, steps):
    """The train.py loop END TO END from raw inputs (reference: Dataset.__getitem__ with num_workers=0 inside
    the training process, av_dataset.py:302-342,365-377): per step, 16 clips are sliced out of the float32 audio memmap and their
    12 decoded uint8 frames gathered on the host, copied to the device, normalised, pushed through the DINO ViT-S/8 extractor
    (192 frames, frame difference + clip normalise fused), STFT'd, and trained on (4 windows + Adam) -- the batch of step i+1
    staged on side streams while step i trains.  Host gather, H2D, front end and the D2H of the losses are all inside the timed
    region."""
    import tempfile
    torch = ctx.torch
    import numpy as np
    from maavss_b200 import av_dataset, avse_model, input_pipeline, optim, train_step, video_attention
    B = B_PER_GPU
    n = HOPS_PER_FRAME * HOP * CLIP_FRAMES
    tmp = tempfile.mkdtemp(prefix="maavss_bench_")
    g = torch.Generator().manual_seed(99)
    input_pipeline.save_audio_memmap(tmp, [(torch.randn(40 * n, generator=g) * 0.1).numpy() for _ in range(4)])
    amap = input_pipeline.AudioMemmap(tmp)
    videos = [torch.randint(0, 256, (CLIP_FRAMES, FRAME, FRAME, 3), generator=g, dtype=torch.uint8).numpy() for _ in range(2 * B)]
    va = video_attention.VideoAttention(state_dict=video_attention.random_state_dict(seed=0))
    ds = av_dataset.AV_Dataset(num_frames=CLIP_FRAMES, frame_hop=2, fft_len=N_FFT, hops_per_frame=HOPS_PER_FRAME, noise_std=NOISE_STD,
                               framerate=FPS, samplerate=SR, attention_extractor=va)
    torch.manual_seed(0)
    model = avse_model.AV_Fusion_Model([B, 2, 64, 128], [B, 1, 8, 4096], 8)
    model.train()
    opt = optim.Adam(model.parameters(), lr=LR)
    step = train_step.FusionTrainStep(model, opt, num_seq=NUM_SEQ, num_frames=NUM_FRAMES, hops_per_frame=HOPS_PER_FRAME, p_size=64,
                                      loss_coeff=LOSS_COEFF, graph=not ctx.args.eager)
    pipe = input_pipeline.ClipBatchPipeline(ds, amap, lambda i: videos[i], va, device=ctx.dev)
    locs = lambda i: [((i + b) % 4, (3 * i + b) % 30) for b in range(B)]
    idxs = lambda i: [(i * B + b) % (2 * B) for b in range(B)]
    losses = []

    def one(i):
        pipe.submit(locs(i + 1), idxs(i + 1))
        x, y, attn, _ = pipe.get()
        out = step(x, y, attn)
        losses.append(torch.stack(out).cpu())

    pipe.submit(locs(0), idxs(0))
    for i in range(3):
        one(i)
    clocks = ClockSampler(ctx.local).start()
    ms = timed_steps(ctx, lambda i: one(i + 3), steps) / steps
    clk = clocks.stop()
    rec = {"metric": "train clips/sec from raw inputs (audio memmap + decoded uint8 frames -> ViT -> train step)", "value": round(B / ms * 1e3, 1),
           "unit": "clips/s", "ms_per_step": round(ms, 3), "steps": steps, "clips_per_step": B, "frames_through_vit_per_step": B * CLIP_FRAMES,
           "h2d_bytes_per_step": B * (4 * n + CLIP_FRAMES * FRAME * FRAME * 3), "d2h_bytes_per_step": 12, "final_loss": float(losses[-1][0]),
           "vit": "ViT-S/8 depth 12, random dino init, bf16 tcgen05", "clocks": clk,
           "input_pipeline": "input_pipeline.ClipBatchPipeline: host gather + H2D + front end of step i+1 on side streams while step i trains"}
    step._graph = None
    del step, model, opt, pipe, va
    torch.cuda.empty_cache()
    return rec


def leg_flow(ctx, iters):
    """flow.py's loop body (Shi-Tomasi corners + 3-level pyramidal LK, flow.py:11-20,28,37) on 256 x 256 synthetic frame pairs,
    beside OpenCV -- the arithmetic the reference itself calls -- on one host core."""
    torch = ctx.torch
    import synth
    from maavss_b200 import flow
    pairs = [synth.flow_frames(100 + i) for i in range(8)]
    dev_pairs = [(torch.from_numpy(a).to(ctx.dev), torch.from_numpy(b).to(ctx.dev)) for a, b in pairs]
    fn = lambda i: flow.track(*dev_pairs[i % 8])
    for i in range(3):
        fn(i)
    ms = timed_steps(ctx, fn, iters) / iters
    rec = {"metric": "sparse optical flow frame pairs/sec (goodFeaturesToTrack + calcOpticalFlowPyrLK, flow.py parameters)",
           "value": round(1e3 / ms, 1), "unit": "pairs/s", "ms_per_pair": round(ms, 4), "frame": "256x256 uint8", "iters": iters,
           "note": "latency bound: ~10 small launches and one 4-byte D2H (corner count) per pair"}
    try:
        import cv2
        cv2.setNumThreads(1)
        t0 = time.perf_counter()
        for i in range(iters):
            a, b = pairs[i % 8]
            p0 = cv2.goodFeaturesToTrack(a, mask=None, maxCorners=100, qualityLevel=0.3, minDistance=7, blockSize=7)
            cv2.calcOpticalFlowPyrLK(a, b, p0, None, winSize=(15, 15), maxLevel=2, criteria=(3, 10, 0.03))
        rec["opencv_cpu_1_thread_pairs_per_s"] = round(iters / (time.perf_counter() - t0), 1)
    except ImportError:
        pass
    return rec


def safe(ctx, name, fn, *a):
    try:
        if ctx.args.verbose:
            print(f"[bench rank {ctx.rank}] leg {name}", file=sys.stderr, flush=True)
        return fn(ctx, *a)
    except Exception as e:                 # a secondary leg must not take the headline line down with it
        import traceback
        traceback.print_exc(file=sys.stderr)
        return {"error": f"{type(e).__name__}: {e}"[:400]}


def run_ours(args):
    quiet_stdout()
    import torch
    import torch.distributed as dist
    from maavss_b200 import _lib, av_dataset, avse_model, optim, parallel, train_step

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (maavss_b200 has no CPU path); use --impl reference for the CPU arm")
    rank, world = parallel.init()
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun --nproc-per-node {args.gpus}")
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    ctx = Ctx()
    ctx.torch, ctx.dist, ctx.rank, ctx.world, ctx.local, ctx.dev, ctx.args = torch, dist, rank, world, local, dev, args
    B = B_PER_GPU
    L = _lib.lib()

    def note(msg):
        if args.verbose:
            print(f"[bench rank {rank}] {msg}", file=sys.stderr, flush=True)

    torch.manual_seed(0)
    model = avse_model.AV_Fusion_Model([B, 2, 64, 128], [B, 1, 8, 4096], 8)
    parallel.broadcast_module(model)
    model.train()
    opt = optim.Adam(model.parameters(), lr=LR)
    ds = av_dataset.AV_Dataset(num_frames=CLIP_FRAMES, fft_len=N_FFT, hops_per_frame=HOPS_PER_FRAME, noise_std=NOISE_STD,
                               framerate=FPS, samplerate=SR)
    step = train_step.FusionTrainStep(model, opt, num_seq=NUM_SEQ, num_frames=NUM_FRAMES, hops_per_frame=HOPS_PER_FRAME,
                                      p_size=64, loss_coeff=LOSS_COEFF, graph=not args.eager)

    # 4 rotating input batches per rank: 4 x 50.7 MB of frames + 147 MB of weights + 3 x 147 MB of optimizer arenas
    # stream through HBM every step, far beyond the 126 MB L2.  (Data parallel, a rank's attention frames are those of
    # parallel.rolled_clip_indices -- synthetic and independent per clip, so any seeded draw stands for them.)
    n_sets = 4
    host = [tuple(t.pin_memory() for t in synth_batch(torch, B, 1234 + 97 * rank + i)) for i in range(n_sets)]
    resident = [tuple(t.to(dev) for t in h) for h in host]

    def full_step(audio, noise, attn):
        x, y = ds.gen_stft_batch(audio, noise)            # STFT + normalise + noise on the device
        return step(x, y, attn)

    note("model built, capturing")
    full_step(*resident[0])                                # builds plans, captures the graph (warm-up steps inside, state restored)
    torch.cuda.synchronize()
    c1 = L.maavss_launch_count()
    full_step(*resident[1])
    torch.cuda.synchronize()
    front_launches = L.maavss_launch_count() - c1          # launches issued outside the graph per step
    for i in range(args.warmup):
        full_step(*resident[i % n_sets])
    torch.cuda.synchronize()
    note("warm-up done")
    clocks = ClockSampler(local).start()
    hold_for_clocks(torch, lambda: full_step(*resident[0]), 0.3)
    ms_total = timed_steps(ctx, lambda i: full_step(*resident[i % n_sets]), args.steps)
    clk = clocks.stop()

    # end to end: pinned host buffers -> H2D inside the timed region -> step -> D2H of the losses.  The input pipeline is
    # the package's double-buffered prefetcher: the copy of step i+1 is issued before step i is launched and overlaps
    # it; every timed step still issues one full H2D copy (K copies in K steps) and reads its losses back.
    feed = train_step.InputPrefetcher(host[0], dev)
    losses = []

    def e2e_step(i):
        feed.put(host[(i + 1) % n_sets])                   # next step's inputs, on the copy stream
        out = full_step(*feed.get())
        feed.release()
        losses.append(torch.stack(out).cpu())              # synchronising D2H read of (loss, a_loss, v_loss)

    feed.put(host[0])                                      # pipeline fill (untimed), then one untimed step
    e2e_step(0)
    ms_e2e_f32 = timed_steps(ctx, lambda i: e2e_step(i + 1), args.steps)
    h2d_f32 = sum(t.numel() * 4 for t in host[0])

    ms_e2e, h2d = ms_e2e_f32, h2d_f32

    # Informational, one GPU only: the same with the attention frames in the format they have on the host in the reference's
    # data path, the frame cache (av_dataset.py:251-278 keeps them as 8-bit images; maavss_b200.attn_cache.AttnFrameCache as
    # raw uint8).  The step's H2D copy is then 14.6 MB instead of 52 MB per 16 clips and maavss_attn_cache_load dequantises +
    # clip-normalises on the device inside the timed region (av_dataset.py:261-269,323-328).
    u8_record = None
    if world == 1:
        from maavss_b200 import attn_cache
        host_u8 = [(a, nz, attn_cache.pack(at.to(dev).clamp(0, 1), "uint8").cpu().pin_memory()) for a, nz, at in host]
        feed8 = train_step.InputPrefetcher(host_u8[0], dev)

        def e2e_step_u8(i):
            feed8.put(host_u8[(i + 1) % n_sets])
            a, nz, q = feed8.get()
            attn = attn_cache.load(q.view(B, CLIP_FRAMES, FRAME, FRAME), attn_diff=False).view(B, 1, CLIP_FRAMES, FRAME, FRAME)
            out = full_step(a, nz, attn)
            feed8.release()
            losses.append(torch.stack(out).cpu())

        feed8.put(host_u8[0])
        e2e_step_u8(0)
        ms_u8 = timed_steps(ctx, lambda i: e2e_step_u8(i + 1), args.steps)
        u8_record = {"value": round(world * B * args.steps / (ms_u8 / 1e3), 2), "ms_per_step": round(ms_u8 / args.steps, 4),
                     "h2d_bytes_per_step": sum(t.numel() * t.element_size() for t in host_u8[0]),
                     "how": "attention frames travel as uint8 (frame-cache format), attn_cache.load dequantises + clip-normalises on the device"}

    graph_launches = 0
    if not args.eager:
        # launches captured into the graph = library launches during one eager replica of the captured body
        c2 = L.maavss_launch_count()
        step._eager(*step.static_inputs())
        torch.cuda.synchronize()
        graph_launches = L.maavss_launch_count() - c2
    launches_per_step = front_launches + graph_launches
    per_step = ms_total / args.steps
    pk, pk_kind = peaks()

    line = {"metric": "train clips/sec", "value": round(world * B * args.steps / (ms_total / 1e3), 2), "unit": "clips/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(per_step, 4),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_record(world),
            "run": {"global_batch": world * B, "parallelism": f"dp{world}", "execution": "eager" if args.eager else "cuda-graph",
                    "sync_bn": world > 1, "global_phasegram_semantics": world > 1,
                    "bn_exchange": ("nvlink peer memory kernel" if parallel.peer_exchange() is not None else "nccl") if world > 1 else None,
                    "grad_allreduce": ("overlapped with the encoder backward" if step.early is not None else "after backward") if world > 1 else None},
            "clocks": clk,
            "e2e": {"value": round(world * B * args.steps / (ms_e2e / 1e3), 2), "unit": "clips/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": 12, "ms_per_step": round(ms_e2e / args.steps, 4),
                    "input_pipeline": "train_step.InputPrefetcher: H2D of step i+1 on a copy stream overlaps step i (2 device slots)",
                    "uint8_frames": u8_record},
            "gpu_launches": launches_per_step * args.steps, "gpu_launches_per_step": launches_per_step,
            "final_loss": float(losses[-1][0])}
    note("timed regions done")

    # ---- roofline.  Whole step: algorithmic bytes of one step / device time of one step (the graph replay measured
    # above).  Per entry point: CUDA events around each C-ABI call of eager steps in the SAME configuration as the graph
    # body (two-stream encoders, overlapped all-reduce); every rank runs them (they contain the collectives).
    flat = opt.flat
    prof_step = train_step.FusionTrainStep(model, opt, num_seq=NUM_SEQ, num_frames=NUM_FRAMES, hops_per_frame=HOPS_PER_FRAME,
                                           p_size=64, loss_coeff=LOSS_COEFF, graph=False)
    agg = {}
    reps = 3
    for rep in range(reps):
        _lib.start_profile()
        x, y = ds.gen_stft_batch(resident[rep % n_sets][0], resident[rep % n_sets][1])
        prof_step(x, y, resident[rep % n_sets][2])
        for name, a, ms in _lib.stop_profile():
            nbytes, tag = algorithmic_bytes(name, a)
            e = agg.setdefault((name, tag), [0.0, 0, nbytes])
            e[0] += ms
            e[1] += 1
    tot = sum(e[0] for e in agg.values())
    top = sorted(agg.items(), key=lambda kv: -kv[1][0])
    # trained parameters that receive gradient in this forward (the unused decoders do not): non-zero entries of the
    # gradient arena after the accumulation windows, before Adam / zero_grad
    opt.zero_grad()
    x, y = ds.gen_stft_batch(resident[0][0], resident[0][1])
    prof_step.micro_steps(x, y, resident[0][2])
    n_used = int((flat.grad != 0).sum())
    opt.zero_grad()
    step_bytes, parts = headline_algorithmic_bytes(flat.numel, n_used, B)
    achieved = step_bytes / (per_step * 1e-3) / 1e9
    try:
        traffic_tbl = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
    except Exception:
        traffic_tbl = {}
    tops = []
    for (n, t), e in top[:3]:
        avg = e[0] / e[1]
        gbs = e[2] / (avg * 1e-3) / 1e9 if e[2] else None
        tops.append({"kernel": f"{n}[{t}]", "share_of_step": round(e[0] / tot, 4), "calls_per_step": e[1] // reps, "avg_ms": round(avg, 5),
                     "algorithmic_bytes": e[2], "achieved": round(gbs, 1) if gbs else None, "frac": round(gbs / pk["hbm_gbs"], 4) if gbs else None,
                     "traffic": traffic_tbl.get(f"{n}[{t}]")})
    line["roofline"] = {"bound": "hbm", "scope": "whole optimizer step (CUDA-graph replay timed above)", "achieved": round(achieved, 1),
                        "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": round(achieved / pk["hbm_gbs"], 4),
                        "algorithmic_bytes": step_bytes, "algorithmic_bytes_parts": parts, "trained_params_with_grad": n_used,
                        "traffic": traffic_tbl.get("whole_step"),
                        "peak_source": (pk_kind + " (MEASURED_PEAKS.json hbm_gbs)") if pk_kind == "measured" else "fallback",
                        "top_kernels": tops,
                        "how": "step: algorithmic bytes / graph-replay ms; kernels: CUDA events around each C-ABI call of eager steps in the "
                               "graph-equivalent configuration (two-stream encoders, overlapped all-reduce), 3 repetitions"}
    line["kernel_breakdown"] = [{"call": f"{n}[{t}]", "share": round(e[0] / tot, 4), "avg_ms": round(e[0] / e[1], 5),
                                 "calls_per_step": e[1] // reps,
                                 "gbs": round(e[2] / (e[0] / e[1] * 1e-3) / 1e9, 1) if e[2] else None} for (n, t), e in top[:12]]
    if args.breakdown and rank == 0:
        with open(args.breakdown, "w") as f:
            for (n, t), e in top:
                f.write(f"{100 * e[0] / tot:6.2f}%  calls/step={e[1] // reps:3d}  avg_ms={e[0] / e[1]:8.5f}  "
                        f"GB/s={(e[2] / (e[0] / e[1] * 1e-3) / 1e9) if e[2] else float('nan'):8.1f}  {n}[{t}]\n")
            f.write(f"# eager step total {tot / reps:.3f} ms (sum of per-call CUDA-event times)\n")
    note("profile steps done")

    # ---- release the headline step before the other legs
    step._graph = None
    del step, prof_step, feed
    torch.cuda.empty_cache()

    if not args.no_frames:
        line["frames"] = safe(ctx, "frames", leg_frames, max(20, min(args.steps, 20)))
    if world == 1:
        if not args.no_vit:
            line["vit"] = safe(ctx, "vit", leg_vit, 20)
        if not args.no_stft:
            line["stft"] = safe(ctx, "stft", leg_stft, 20)
        if not args.no_pipeline:
            line["pipeline"] = safe(ctx, "pipeline", leg_pipeline, 20)
            line["flow"] = safe(ctx, "flow", leg_flow, 40)
        if not args.no_torch:
            line["torch_eager_b200"] = safe(ctx, "torch_eager", leg_torch_eager)
        if not args.no_cpu:
            cps, ms_cpu, threads, done = time_cpu(torch, B, 1000, 1, budget_s=args.cpu_seconds)
            line["cpu_baseline"] = {"value": round(cps, 3), "unit": "clips/s", "cores": threads, "kind": "port",
                                    "sample": f"{done} full steps of {B} clips (same step) in {ms_cpu * done / 1e3:.1f} s, oracle on torch CPU"}
    if rank == 0:
        emit(line)
    if world > 1:
        # Tearing the NCCL communicator down while captured graphs still hold its kernels hangs (seen at N=2 on B200):
        # drain the device, meet at a barrier and leave without destroy_process_group.
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--eager", action="store_true", help="run the steps without CUDA-graph capture")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-frames", action="store_true", help="skip the configs[3] leg")
    ap.add_argument("--no-vit", action="store_true", help="skip the configs[2] leg")
    ap.add_argument("--no-stft", action="store_true", help="skip the configs[4] leg")
    ap.add_argument("--no-torch", action="store_true", help="skip the PyTorch-eager-on-B200 leg")
    ap.add_argument("--no-pipeline", action="store_true", help="skip the raw-inputs-to-step pipeline leg")
    ap.add_argument("--only-headline", action="store_true", help="headline + roofline only (profiling runs)")
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    ap.add_argument("--verbose", action="store_true", help="progress notes on stderr")
    ap.add_argument("--breakdown", default=None, help="write the full per-call breakdown of the eager profiling steps to this file")
    args = ap.parse_args()
    if args.only_headline:
        args.no_cpu = args.no_frames = args.no_vit = args.no_stft = args.no_torch = args.no_pipeline = True
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Headline benchmark: train clips/s of the avse_model train step (BASELINE.json configs[1]:
"avse_model train step, batch 16 synthetic audio+frames, 1 B200").

One step = the hot path over one batch of 16 synthetic clips per GPU:
    audio front end   STFT (n_fft 256, hop 66) -> real/imag planes -> 1/max|.| -> + noise      (av_dataset.py:335-342)
    4 micro-steps     video_phasegram of 8 attention frames -> AV_Fusion_Model forward -> MSE(a) + 0.001 MSE(v)
                      -> backward, gradients accumulated                                         (train.py:136-159)
    optimizer         [N > 1: NCCL all-reduce of the flat gradient arena] -> Adam(lr 1e-5)       (train.py:161-162)
"frames" are the per-clip motion-attention frames (12 x 256 x 256) the dataset hands to the loop; the ViT that
produces them is configs[2] and is not part of this step.

    python bench.py --gpus N --steps K --warmup W            our arm  (torchrun for N > 1, weak scaling: 16 clips / GPU)
    python bench.py --impl reference ...                     the reference's CPU path (oracle port) on the host cores

Prints ONE JSON line (see DESIGN.md "Measurement").
"""
import argparse
import json
import os
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
WORKLOAD = "avse_model.AV_Fusion_Model train step, 16 synthetic clips/GPU (audio 6336 samples + 12x256x256 attention frames)"


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


def synth_batch(torch, B, seed):
    g = torch.Generator().manual_seed(seed)
    hop = int((SR / FPS) / HOPS_PER_FRAME)
    n = HOPS_PER_FRAME * hop * CLIP_FRAMES
    audio = torch.randn(B, n, generator=g) * 0.1
    noise = torch.randn(B, 2, n // hop, N_FFT // 2, generator=g)
    attn = torch.rand(B, 1, CLIP_FRAMES, FRAME, FRAME, generator=g)
    return audio, noise, attn


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
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

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
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


# ----------------------------------------------------------------------------------------------- reference arm
def oracle_step_factory(torch, B):
    """The reference's CPU path for this step, restated in oracle/ (the reference is pure Python and not on this box)."""
    from oracle import audio as oa
    from oracle import models as om
    from oracle import phasegram as op
    torch.manual_seed(0)
    model = om.AVFusionOracle([B, 2, 64, 128], [B, 1, 8, 4096], 8)
    model.train()
    opt = torch.optim.Adam(model.parameters(), lr=LR)
    hop = oa.calc_hop_size(NUM_FRAMES, HOPS_PER_FRAME, FPS, SR)[0]
    window = oa.hamming_window(N_FFT)

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
        return float(loss.detach())
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
    cps, ms, threads, done = time_cpu(torch, B_PER_GPU, args.steps, min(args.warmup, 2), budget_s=150)
    line = {"impl": "reference", "metric": "train clips/sec", "value": round(cps, 3), "unit": "clips/s", "n_gpus": args.gpus,
            "steps": done, "warmup": min(args.warmup, 2), "ms_per_step": round(ms, 3), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "batch_per_step": B_PER_GPU, "device": "host CPU"},
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
    if name == "maavss_adam_step":
        return 28 * iv(a[4]), f"n{iv(a[4])}"
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
    if name == "maavss_video_phasegram":
        B, T, H, W, R = (iv(a[i]) for i in range(4, 9))
        return 4 * B * T * (H * W + R * R), f"B{B} T{T} {H}x{W}->{R}"
    return None, ""


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
    B = B_PER_GPU
    L = _lib.lib()

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
    # stream through HBM every step, far beyond the 126 MB L2.
    n_sets = 4
    host = [tuple(t.pin_memory() for t in synth_batch(torch, B, 1234 + 97 * rank + i)) for i in range(n_sets)]
    resident = [tuple(t.to(dev) for t in h) for h in host]

    def full_step(audio, noise, attn):
        x, y = ds.gen_stft_batch(audio, noise)            # STFT + normalise + noise on the device
        return step(x, y, attn)

    c0 = L.maavss_launch_count()
    if args.verbose:
        print(f"[bench rank {rank}] model built, capturing", file=sys.stderr, flush=True)
    full_step(*resident[0])                                # builds plans, captures the graph (2 warm-up steps inside)
    torch.cuda.synchronize()
    c1 = L.maavss_launch_count()
    full_step(*resident[1])
    torch.cuda.synchronize()
    front_launches = L.maavss_launch_count() - c1          # launches issued outside the graph per step
    for i in range(args.warmup):
        full_step(*resident[i % n_sets])
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(args.steps):
            fn(i)
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms)

    if args.verbose:
        print(f"[bench rank {rank}] warm-up done", file=sys.stderr, flush=True)
    clocks = ClockSampler(local)
    clocks.start()
    ms_total = timed(lambda i: full_step(*resident[i % n_sets]))
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
    ms_e2e = timed(lambda i: e2e_step(i + 1))
    h2d = sum(t.numel() * 4 for t in host[0])

    graph_launches = None
    if not args.eager:
        # launches captured into the graph = library launches during one eager replica of the captured body
        c2 = L.maavss_launch_count()
        step._eager(*step.static_inputs())
        torch.cuda.synchronize()
        graph_launches = L.maavss_launch_count() - c2
    launches_per_step = front_launches + (graph_launches or 0)

    line = {"metric": "train clips/sec", "value": round(world * B * args.steps / (ms_total / 1e3), 2), "unit": "clips/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms_total / args.steps, 4),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "clips_per_gpu": B, "global_batch": world * B, "parallelism": f"dp{world}",
                       "num_seq": NUM_SEQ, "n_fft": N_FFT, "hop": ds.hop, "execution": "eager" if args.eager else "cuda-graph",
                       "l2": f"inputs rotate over {n_sets} batches ({n_sets * h2d / 1e6:.0f} MB) + 147 MB weights + optimizer arenas > 126 MB L2",
                       "sync_bn": world > 1, "grad_allreduce": ("overlapped with the encoder backward" if step.early is not None else "after backward") if world > 1 else None},
            "clocks": clk,
            "e2e": {"value": round(world * B * args.steps / (ms_e2e / 1e3), 2), "unit": "clips/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": 12, "ms_per_step": round(ms_e2e / args.steps, 4),
                    "input_pipeline": "train_step.InputPrefetcher: H2D of step i+1 on a copy stream overlaps step i (2 device slots)"},
            "gpu_launches": launches_per_step * args.steps,
            "final_loss": float(losses[-1][0])}

    def note(msg):
        if args.verbose:
            print(f"[bench rank {rank}] {msg}", file=sys.stderr, flush=True)

    note("timed regions done")
    # ---- roofline of the dominant kernel: per-call CUDA-event timing of one eager step (same kernels, same shapes).
    # Every rank runs it (the step contains the sync-BN and gradient collectives); rank 0 reports.
    if True:
        pk, pk_kind = peaks()
        prof_step = train_step.FusionTrainStep(model, opt, num_seq=NUM_SEQ, num_frames=NUM_FRAMES, hops_per_frame=HOPS_PER_FRAME,
                                               p_size=64, loss_coeff=LOSS_COEFF, graph=False, overlap_allreduce=False)
        agg = {}
        for rep in range(3):
            _lib.start_profile()
            x, y = ds.gen_stft_batch(resident[rep % n_sets][0], resident[rep % n_sets][1])
            prof_step(x, y, resident[rep % n_sets][2])
            for name, a, ms in _lib.stop_profile():
                nbytes, tag = algorithmic_bytes(name, a)
                k = (name, tag)
                e = agg.setdefault(k, [0.0, 0, nbytes])
                e[0] += ms
                e[1] += 1
        tot = sum(e[0] for e in agg.values())
        top = sorted(agg.items(), key=lambda kv: -kv[1][0])
        (name, tag), (ms_sum, calls, nbytes) = top[0]
        achieved = (nbytes / (ms_sum / calls * 1e-3) / 1e9) if nbytes else None
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json"))).get(f"{name}[{tag}]")
        except Exception:
            traffic = None
        line["roofline"] = {"bound": "hbm", "kernel": f"{name}[{tag}]", "achieved": round(achieved, 1) if achieved else None,
                            "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": round(achieved / pk["hbm_gbs"], 4) if achieved else None,
                            "traffic": traffic, "traffic_source": "profiles/ncu_traffic.json (ncu --set full, dram read + write per launch)" if traffic else None,
                            "algorithmic_bytes": nbytes, "peak_source": pk_kind + " (MEASURED_PEAKS.json hbm_gbs)" if pk_kind == "measured" else "fallback",
                            "share_of_step": round(ms_sum / tot, 4), "avg_ms": round(ms_sum / calls, 5), "calls_per_step": calls // 3,
                            "how": "CUDA events around each C-ABI call of one eager (non-graph) step, 3 repetitions"}
        line["kernel_breakdown"] = [{"call": f"{n}[{t}]", "share": round(e[0] / tot, 4), "avg_ms": round(e[0] / e[1], 5),
                                     "calls_per_step": e[1] // 3,
                                     "gbs": round(e[2] / (e[0] / e[1] * 1e-3) / 1e9, 1) if e[2] else None} for (n, t), e in top[:12]]
        if args.breakdown and rank == 0:
            with open(args.breakdown, "w") as f:
                for (n, t), e in top:
                    f.write(f"{100 * e[0] / tot:6.2f}%  calls/step={e[1] // 3:3d}  avg_ms={e[0] / e[1]:8.5f}  "
                            f"GB/s={(e[2] / (e[0] / e[1] * 1e-3) / 1e9) if e[2] else float('nan'):8.1f}  {n}[{t}]\n")
                f.write(f"# eager step total {tot / 3:.3f} ms (sum of per-call CUDA-event times)\n")
        note("profile steps done")
        if rank == 0 and world == 1 and not args.no_cpu:
            cps, ms_cpu, threads, done = time_cpu(torch, B, 1000, 1, budget_s=args.cpu_seconds)
            line["cpu_baseline"] = {"value": round(cps, 3), "unit": "clips/s", "cores": threads, "kind": "port",
                                    "sample": f"{done} full steps of {B} clips (same step) in {ms_cpu * done / 1e3:.1f} s, oracle on torch CPU"}
        if rank == 0:
            emit(line)
    if world > 1:
        # Tearing the NCCL communicator down while captured graphs still hold its kernels hangs (seen at N=2 on B200):
        # drop the graphs, drain the device, meet at a barrier and leave without destroy_process_group.
        step._graph = None
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
    ap.add_argument("--eager", action="store_true", help="run the step without CUDA-graph capture")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    ap.add_argument("--verbose", action="store_true", help="progress notes on stderr")
    ap.add_argument("--breakdown", default=None, help="write the full per-call breakdown of the eager profiling steps to this file")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()

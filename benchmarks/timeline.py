"""Kernel timeline of the captured headline train step (CUPTI activity records through torch.profiler): start, duration and
stream of every kernel of a few graph replays, so that the critical path and the overlap of the side streams can be read.
usage: python benchmarks/timeline.py [out.csv]      (GPU box; writes gpurun_out/timeline.csv by default)
       python benchmarks/timeline.py --view timeline.csv [step_index]      (anywhere)"""
import csv
import os
import re
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def capture(out):
    import torch
    import bench
    from maavss_b200 import av_dataset, avse_model, optim, train_step
    from maavss_b200 import parallel
    rank, world = parallel.init()              # torchrun: data-parallel step (sync-BN, gradient exchange); rank 0 writes the CSV
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    B = bench.B_PER_GPU
    torch.manual_seed(0)
    model = avse_model.AV_Fusion_Model([B, 2, 64, 128], [B, 1, 8, 4096], 8)
    parallel.broadcast_module(model)
    model.train()
    opt = optim.Adam(model.parameters(), lr=bench.LR)
    ds = av_dataset.AV_Dataset(num_frames=bench.CLIP_FRAMES, fft_len=bench.N_FFT, hops_per_frame=bench.HOPS_PER_FRAME,
                               noise_std=bench.NOISE_STD, framerate=bench.FPS, samplerate=bench.SR)
    step = train_step.FusionTrainStep(model, opt, num_seq=bench.NUM_SEQ, num_frames=bench.NUM_FRAMES,
                                      hops_per_frame=bench.HOPS_PER_FRAME, p_size=64, loss_coeff=bench.LOSS_COEFF, graph=True)
    sets = [tuple(t.to(dev) for t in bench.synth_batch(torch, B, 1234 + i)) for i in range(4)]

    def full_step(audio, noise, attn):
        x, y = ds.gen_stft_batch(audio, noise)
        return step(x, y, attn)

    for i in range(8):
        full_step(*sets[i % 4])
    torch.cuda.synchronize()
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for i in range(3):
            full_step(*sets[i % 4])
        torch.cuda.synchronize()
    rows = []
    for e in prof.events():
        if e.device_type is not None and "cuda" in str(e.device_type).lower():
            rows.append((e.time_range.start, e.time_range.end - e.time_range.start, getattr(e, "device_resource_id", -1), e.name))
    rows.sort()
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        if rank != 0:
            os._exit(0)
    t0 = rows[0][0] if rows else 0
    os.makedirs(os.path.dirname(out) or ".", exist_ok=True)
    with open(out, "w") as f:
        f.write("start_us,dur_us,stream,name\n")
        for s, d, st, n in rows:
            n = n.replace('"', "'")
            f.write(f"{s - t0:.3f},{d:.3f},{st},\"{n[:120]}\"\n")
    print(f"{len(rows)} device records -> {out}", flush=True)
    if world > 1:
        os._exit(0)


def view(path, which=-1):
    rows = [(float(r["start_us"]), float(r["dur_us"]), r["stream"], r["name"]) for r in csv.DictReader(open(path))]
    starts = [i for i, r in enumerate(rows) if "stft_fwd_kernel" in r[3]]
    bounds = starts + [len(rows)]
    k = which % len(starts)
    step = rows[bounds[k]:bounds[k + 1]]
    t0 = step[0][0]
    busy_end, idle = t0, 0.0
    for s, d, st, n in step:
        n = re.sub(r"\(.*$", "", n)
        n = re.sub(r"\(anonymous namespace\)::|void ", "", n)[:46]
        gap = s - busy_end
        if gap > 0:
            idle += gap
        print(f"{s - t0:9.1f} {d:7.1f} s{st:>3} {'gap %5.1f' % gap if gap > 0.3 else '         '} {n}")
        busy_end = max(busy_end, s + d)
    print(f"# step span {busy_end - t0:.1f} us, GPU fully idle {idle:.1f} us, kernels {len(step)}, sum of durations {sum(r[1] for r in step):.1f} us")


if __name__ == "__main__":
    if "--view" in sys.argv:
        i = sys.argv.index("--view")
        view(sys.argv[i + 1], int(sys.argv[i + 2]) if len(sys.argv) > i + 2 else -1)
    else:
        capture(next((a for a in sys.argv[1:] if not a.startswith("--")), "gpurun_out/timeline.csv"))

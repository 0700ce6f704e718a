"""Data-parallel parity on real GPUs: N ranks over NCCL, each on its contiguous shard of the clip batch,
must reproduce ONE process on the concatenated batch -- losses, parameters after two optimizer steps and BatchNorm
running statistics -- for both train steps:

  * FusionTrainStep (train.py:136-162): sync-BN partial sums, the overlapped gradient all-reduce inside the captured
    CUDA graph, and the whole-batch semantics of utilities.video_phasegram (max over the global batch, utilities.py:226-227;
    batch-axis roll of the all-dims fftshift, utilities.py:210 -> parallel.rolled_clip_indices);
  * FramesTrainStep (train_avse_frames.py:150-181): sync-BN over BatchNorm3d/2d + gradient all-reduce, eager.

Needs >= 2 GPUs (skipped otherwise): run with `gpurun --gpus 2 -- python -m pytest tests/test_gpu_multi.py -m gpu`.
"""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

import synth
from oracle import models as om

pytestmark = pytest.mark.gpu

B_GLOBAL, WORLD, STEPS = 4, 2, 2


def _free_port():
    """A port nothing listens on, outside the ephemeral range the kernel hands to outgoing connections (a port taken from
    bind(0) was lost to such a connection once between this probe and the rendezvous: EADDRINUSE, 22 GPU-minutes)."""
    import random
    for _ in range(64):
        p = random.randint(20000, 30000)
        s = socket.socket()
        try:
            s.bind(("127.0.0.1", p))
            return p
        except OSError:
            continue
        finally:
            s.close()
    raise RuntimeError("no free rendezvous port")


def _digest(model):
    out, seen = [], set()
    for k, p in model.named_parameters():
        if p.data_ptr() in seen:
            continue
        seen.add(p.data_ptr())
        f = p.detach().flatten().cpu()
        out.append(np.concatenate(([float(f.double().sum())], f[synth.sample_idx(f.numel())].double().numpy())))
    bn = np.concatenate([v.detach().flatten().double().cpu().numpy() for k, v in model.state_dict().items()
                         if k.endswith("running_mean") or k.endswith("running_var") or k.endswith("num_batches_tracked")])
    return np.stack(out), bn


def _inputs(kind):
    if kind == "fusion":
        xs, ys = synth.stft_batch(B_GLOBAL, 96, 128, seed=37)
    else:
        xs, ys = synth.stft_batch(B_GLOBAL, 96, 129, seed=39)
    attn = synth.attn_clip(B_GLOBAL, 12, 256, 256, seed=17)
    return xs, ys, attn


def _build(kind, B):
    from maavss_b200 import avse_model, avse_model_final, optim, train_step
    if kind == "fusion":
        ref = om.AVFusionOracle([2, 2, 64, 128], [2, 1, 8, 4096], 8)
        sd = synth.fill_state_dict(ref.state_dict(), seed=31)
        m = avse_model.AV_Fusion_Model([B, 2, 64, 128], [B, 1, 8, 4096], 8)
    else:
        ref = om.AVFusionFramesOracle([2, 2, 64, 129], [2, 1, 8, 256, 256], 8)
        sd = synth.fill_state_dict(ref.state_dict(), seed=53)
        m = avse_model_final.AV_Fusion_Model_Frames([B, 2, 64, 129], [B, 1, 8, 256, 256], 8)
    m.load_state_dict(sd)
    m.train()
    opt = optim.Adam(m.parameters(), lr=1e-5)
    if kind == "fusion":
        step = train_step.FusionTrainStep(m, opt, graph=True)
    else:
        step = train_step.FramesTrainStep(m, opt, graph=False)
    return m, step


def _run(kind, xs, ys, attn):
    m, step = _build(kind, xs.shape[0])
    xs, ys, attn = xs.cuda().contiguous(), ys.cuda().contiguous(), attn.cuda().contiguous()
    losses = [[float(v) for v in step(xs, ys, attn)] for _ in range(STEPS)]
    torch.cuda.synchronize()
    return (np.array(losses),) + _digest(m)


def _worker(rank, world, port, kind, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    torch.cuda.set_device(rank)
    import torch.distributed as dist
    from maavss_b200 import parallel
    parallel.init(backend="nccl", sync_bn=True)
    xs, ys, attn = _inputs(kind)
    lo, hi = parallel.shard_bounds(B_GLOBAL, rank, world, require_even=True)
    # the train.py path pairs the audio of global clip i with the phasegram of clip (i - B/2) mod B (utilities.py:210)
    a_idx = parallel.rolled_clip_indices(B_GLOBAL, rank, world) if kind == "fusion" else list(range(lo, hi))
    res = _run(kind, xs[lo:hi], ys[lo:hi], attn[a_idx])
    assert parallel.peer_exchange() is not None, "NVLink peer exchange did not come up on this box"
    assert parallel.peer_status() == 0, "a rank gave up waiting for a peer in the BatchNorm exchange"
    q.put((rank,) + res)
    dist.barrier()
    torch.cuda.synchronize()
    sys.stdout.flush()
    os._exit(0)          # no destroy_process_group while captured graphs hold NCCL kernels (DESIGN.md "Measurement")


@pytest.mark.timeout(600)
@pytest.mark.parametrize("kind", ["fusion", "frames"])
def test_two_ranks_equal_one_process(kind):
    if torch.cuda.device_count() < WORLD:
        pytest.skip("needs 2 GPUs: gpurun --gpus 2")
    single = _run(kind, *_inputs(kind))
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, WORLD, port, kind, q)) for r in range(WORLD)]
    for p in procs:
        p.start()
    # fail fast when a rank dies (a worker that cannot bind the rendezvous port must not cost the whole timeout)
    import queue
    import time
    res, deadline = [], time.time() + 240
    while len(res) < WORLD:
        try:
            res.append(q.get(timeout=2))
        except queue.Empty:
            dead = [p.exitcode for p in procs if p.exitcode not in (None, 0)]
            if dead or time.time() > deadline:
                for p in procs:
                    if p.is_alive():
                        p.kill()
                pytest.fail(f"data-parallel workers failed (exit codes {dead}, {len(res)} of {WORLD} results)")
    res.sort(key=lambda r: r[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    l1, p1, bn1 = single
    # reported losses are means over the rank's shard: equal shards -> their mean is the global mean
    lN = np.mean([r[1] for r in res], axis=0)
    np.testing.assert_allclose(lN, l1, rtol=2e-5)
    noisy, visual, numel = _row_classes(kind)
    for r in res:
        # every rank ends with the same parameters and running statistics as the single process
        if kind == "fusion":
            # Sampled entries: equal to 1e-5 -- except the few whose gradient is rounding noise: the rank-local batch (2 clips)
            # and the whole batch (4 clips) take different kernels for some layers (the tensor-core convolutions need >= 128
            # output positions) and Adam turns a flipped sign into a full +-lr move per step.  So >= 99 % tight, nothing beyond
            # 2 steps x lr, and the per-tensor sums (column 0) within 1e-4 of the largest possible drift 2 x lr x numel.
            a, b = r[2][~noisy], p1[~noisy]
            close = np.isclose(a[:, 1:], b[:, 1:], rtol=1e-5, atol=2e-7)
            assert close.mean() > 0.99, close.mean()
            np.testing.assert_allclose(a[:, 1:], b[:, 1:], rtol=1e-5, atol=2.5e-5)
            assert (np.abs(a[:, 0] - b[:, 0]) <= 1e-5 * np.abs(b[:, 0]) + 1e-4 * 2 * 1e-5 * numel[~noisy] + 2e-7).all()
        else:
            # AV_Fusion_Model_Frames: the sync-BN partial sums are added in another order (1e-7 noise), which flips the arg-max of
            # a few MaxPool3d windows and with it the SIGN of some near-zero visual-encoder gradient entries; Adam turns a sign
            # into a full +-lr move.  Sampled entries: exact to 1e-5 outside the visual encoder, within 2 steps x 2 lr inside it;
            # the per-tensor sums (column 0) add such moves up and are not compared for this model.
            plain = ~noisy & ~visual
            np.testing.assert_allclose(r[2][plain][:, 1:], p1[plain][:, 1:], rtol=1e-5, atol=2e-7)
            np.testing.assert_allclose(r[2][visual][:, 1:], p1[visual][:, 1:], rtol=0, atol=4.5e-5)
            assert np.isclose(r[2][visual][:, 1:], p1[visual][:, 1:], rtol=1e-5, atol=2e-7).mean() > 0.9
        np.testing.assert_allclose(r[2][noisy][:, 1:], p1[noisy][:, 1:], rtol=0, atol=2.5e-5)
        # running statistics: 1e-3; near-zero entries of the Frames model's deep layers carry the pool arg-max noise (seen: 1e-5)
        np.testing.assert_allclose(r[3], bn1, rtol=1e-3, atol=1e-6 if kind == "fusion" else 2e-5)
    np.testing.assert_array_equal(res[0][2], res[1][2])


def _row_classes(kind):
    """-> (noisy, visual) masks over the digest rows and the element count of every row's tensor.  noisy: conv biases in front of a train-mode BatchNorm (exactly-zero
    gradient, Adam acts on rounding noise; see test_gpu_models.test_fusion_train_step_golden).  visual: the Conv3d /
    BatchNorm3d parameters below the MaxPool3d layers of AV_Fusion_Model_Frames."""
    m, _ = _build(kind, 2)
    names = [k for k, _ in m.named_parameters()]
    names = [k for i, k in enumerate(names) if k not in names[:i]]
    mods = dict(m.named_modules())
    noisy = np.array([k.endswith(".bias") and type(mods[k.rsplit(".", 1)[0]]).__name__.startswith("Conv") for k in names])
    visual = np.array([k.startswith("visual_encoder") for k in names])
    params = dict(m.named_parameters())
    numel = np.array([params[k].numel() for k in names], dtype=np.float64)
    return noisy, visual, numel

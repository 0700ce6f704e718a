"""GPU parity of utilities.video_phasegram (fused resize + fft2 + fftshift + angle + cumsum + diff + max-normalise)
against the reference golden output and the oracle.  The angle of a (near-)zero bin is ill-conditioned (a rounding
difference can move it by 2*pi, which the cumsum then carries), so the comparison is robust: median error plus the
fraction of elements that differ visibly -- the same criterion that pins the oracle itself."""
import numpy as np
import pytest
import torch

import synth
from oracle import phasegram as op

pytestmark = pytest.mark.gpu


def t(a):
    return torch.from_numpy(np.asarray(a))


def test_phasegram_golden(golden):
    from maavss_b200 import utilities
    g = golden("phasegram")
    attn = synth.attn_clip(4, 8, 256, 256, seed=17)
    pg = utilities.video_phasegram(attn.cuda(), resize=(64, 64)).cpu()
    ref = t(g["pgram"])
    assert pg.shape == ref.shape == (4, 1, 8, 4096)
    assert float((pg - ref).abs().median()) < 2e-6
    assert float(((pg - ref).abs() > 2e-3).float().mean()) < 2e-3
    assert float(pg.abs().max()) == pytest.approx(1.0, abs=1e-6)
    assert float(pg[:, :, 0].abs().max()) == 0.0                    # zero-padded first frame of the temporal diff
    pn = utilities.video_phasegram(attn[:2].cuda(), resize=(64, 64), diff=False, cumulative=False).cpu()
    refn = t(g["pgram_nodiff"])
    assert float(((pn - refn).abs() > 1e-4).float().mean()) < 1e-3


def test_phasegram_fftshift_rolls_batch_and_time():
    """fftshift over all dims rolls the batch axis by B/2 and time by T/2; reproduced."""
    from maavss_b200 import utilities
    attn = synth.attn_clip(4, 6, 128, 128, seed=3)
    pg = utilities.video_phasegram(attn.cuda(), resize=(64, 64), diff=False, normalize=False).cpu()
    ref = op.video_phasegram(attn, resize=(64, 64), diff=False, normalize=False)
    bad = ((pg - ref).abs() > 1e-3).float().mean(dim=(1, 2, 3))
    assert float(bad.max()) < 5e-3
    # un-rolled comparison must NOT match (guards against silently 'fixing' the quirk)
    unrolled = torch.roll(ref, shifts=(2, 3), dims=(0, 2))
    assert float(((pg - unrolled).abs() > 1e-3).float().mean()) > 0.5


def test_phasegram_no_resize_matches_oracle():
    from maavss_b200 import utilities
    fr = torch.rand(2, 1, 4, 64, 64, generator=synth.gen(8))
    pg = utilities.video_phasegram(fr.cuda()).cpu()
    ref = op.video_phasegram(fr)
    assert float((pg - ref).abs().median()) < 2e-6
    assert float(((pg - ref).abs() > 2e-3).float().mean()) < 2e-3


@pytest.mark.parametrize("shape", [(2, 5, 100, 90), (2, 4, 130, 256), (1, 3, 300, 384), (2, 4, 66, 1000)])
def test_phasegram_resize_paths(shape):
    """Chunked shared-memory resize: scalar staging (W % 4 != 0), a last partial chunk (H % 32 != 0), wide rows."""
    from maavss_b200 import utilities
    B, T, H, W = shape
    attn = synth.attn_clip(B, T, H, W, seed=5)
    pg = utilities.video_phasegram(attn.cuda(), resize=(64, 64), diff=False, cumulative=False, normalize=False).cpu()
    ref = op.video_phasegram(attn, resize=(64, 64), diff=False, cumulative=False, normalize=False)
    assert float((pg - ref).abs().median()) < 2e-6
    assert float(((pg - ref).abs() > 1e-4).float().mean()) < 2e-3


def test_phasegram_time_slice_view():
    """train.py:139 passes a time slice of the clip's attention frames: strided batch, no copy."""
    from maavss_b200 import utilities
    attn = synth.attn_clip(3, 12, 256, 256, seed=9)
    view = attn.cuda()[:, :, 2:10]
    pg = utilities.video_phasegram(view, resize=(64, 64)).cpu()
    ref = utilities.video_phasegram(view.contiguous(), resize=(64, 64)).cpu()
    assert torch.equal(pg, ref)
    orc = op.video_phasegram(attn[:, :, 2:10], resize=(64, 64))
    assert float((pg - orc).abs().median()) < 2e-6


def test_phasegram_windows_equal_per_window_calls():
    """train.py:136-139 calls video_phasegram on 4 overlapping 8-frame windows of a 12-frame clip; the one-pass version
    transforms each of the 12 frames once and must reproduce the 4 calls bit for bit (same per-frame arithmetic)."""
    from maavss_b200 import phasegram_ops
    attn = synth.attn_clip(4, 12, 256, 256, seed=21).cuda()
    stacked = phasegram_ops.video_phasegram_windows(attn, 4, 8, resize=(64, 64))
    assert stacked.shape == (16, 1, 8, 4096)
    for j in range(4):
        one = phasegram_ops.video_phasegram(attn[:, :, j:j + 8], resize=(64, 64))
        assert torch.equal(stacked[4 * j:4 * j + 4], one), j
        orc = op.video_phasegram(attn.cpu()[:, :, j:j + 8], resize=(64, 64))
        assert float((one.cpu() - orc).abs().median()) < 2e-6
        assert float(((one.cpu() - orc).abs() > 2e-3).float().mean()) < 2e-3
    # odd batch / odd window length: rolls are B // 2 and Tw // 2 (torch.fft.fftshift)
    attn = synth.attn_clip(3, 7, 128, 128, seed=22)
    st = phasegram_ops.video_phasegram_windows(attn.cuda(), 3, 5, resize=(64, 64)).cpu()
    for j in range(3):
        orc = op.video_phasegram(attn[:, :, j:j + 5], resize=(64, 64))
        assert float((st[3 * j:3 * j + 3] - orc).abs().median()) < 2e-6, j


def test_phasegram_sharded_ranks_equal_global_batch():
    """under data parallelism the whole-batch max (utilities.py:226-227) and the batch-axis roll of the
    all-dims fftshift (utilities.py:210) must be those of the GLOBAL batch.  Two emulated ranks (loader remap
    parallel.rolled_clip_indices + batch_roll 0 + max of the per-rank window maxima) == one process on all clips."""
    from maavss_b200 import parallel, phasegram_ops
    from maavss_b200._lib import check, dptr, fptr, lib, stream_ptr
    Bg, world, G, Tw, Tc, R = 6, 2, 4, 8, 12, 64
    attn = synth.attn_clip(Bg, Tc, 128, 128, seed=23).cuda()
    whole = phasegram_ops.video_phasegram_windows(attn, G, Tw, resize=(R, R))            # (G*Bg, 1, Tw, R*R)
    L = lib()
    shards = []
    for r in range(world):
        idx = parallel.rolled_clip_indices(Bg, r, world)
        f = attn[idx, 0].contiguous()
        B = f.shape[0]
        pflat = torch.empty(B, Tc, R * R, device="cuda")
        scratch = torch.empty_like(pflat)
        mx = torch.empty(G, device="cuda", dtype=torch.int32)
        check(L.maavss_phasegram_frames(fptr(f), fptr(pflat), fptr(scratch), B, Tc, 128, 128, R, Tc * 128 * 128, 1, stream_ptr()))
        check(L.maavss_phasegram_windows_max(fptr(pflat), dptr(mx), B, Tc, Tw, G, R, 0, 1, stream_ptr()))
        shards.append((B, pflat, mx))
    gmx = torch.maximum(shards[0][2], shards[1][2])                                      # the MAX all-reduce
    outs = []
    for B, pflat, _ in shards:
        out = torch.empty(G * B, Tw, R * R, device="cuda")
        check(L.maavss_phasegram_windows_write(fptr(pflat), dptr(gmx), fptr(out), B, Tc, Tw, G, R, 0, 1, 1, stream_ptr()))
        outs.append(out.view(G, B, Tw, R * R))
    both = torch.cat(outs, 1).reshape(G * Bg, 1, Tw, R * R)
    assert torch.equal(both, whole)
    # and the per-rank-local variant differs (guards the claim): local roll + local max
    local = phasegram_ops.video_phasegram_windows(attn[:3], G, Tw, resize=(R, R)).view(G, 3, Tw, R * R)
    assert not torch.equal(local, whole.view(G, Bg, Tw, R * R)[:, :3])

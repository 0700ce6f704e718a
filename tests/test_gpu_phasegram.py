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

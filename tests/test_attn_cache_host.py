"""Host logic of the raw attention-frame cache (maavss_b200/attn_cache.py): slot index, validity map, persistence.
No compute calls (those are in test_gpu_attn_cache.py)."""
import numpy as np
import pytest

from maavss_b200.attn_cache import AttnFrameCache


def test_reserve_slots_and_verify(tmp_path):
    c = AttnFrameCache(str(tmp_path / "attn.u8"), capacity=40, frame_hw=(8, 8))
    assert c.reserve("vid_a", 16) == 0 and c.reserve("vid_b", 20) == 16
    assert c.reserve("vid_a", 12) == 0                       # idempotent
    with pytest.raises(ValueError):
        c.reserve("vid_a", 17)
    with pytest.raises(ValueError):
        c.reserve("vid_c", 5)                                # 36 + 5 > 40
    assert not c.has("vid_a", 0, 4) and not c.has("nope", 0, 4)
    c.frames[2:6] = 7
    c.valid[2:6] = 1
    assert c.has("vid_a", 2, 4) and not c.has("vid_a", 2, 5) and not c.has("vid_a", 1, 4)
    with pytest.raises(IndexError):
        c.has("vid_a", 14, 4)                                # clip runs past the video (av_dataset.py:257-258)
    c.flush()

    r = AttnFrameCache(str(tmp_path / "attn.u8"), mode="r")
    assert (r.fmt, r.capacity, r.H, r.W, r.used) == ("uint8", 40, 8, 8, 36)
    assert r.index == {"vid_a": (0, 16), "vid_b": (16, 20)}
    assert r.has("vid_a", 2, 4) and int(r.frames[3, 0, 0]) == 7 and r.frames.dtype == np.uint8


def test_new_cache_arguments(tmp_path):
    with pytest.raises(ValueError):
        AttnFrameCache(str(tmp_path / "x"), capacity=0)
    with pytest.raises(ValueError):
        AttnFrameCache(str(tmp_path / "x"), capacity=4, fmt="jpeg")
    c = AttnFrameCache(str(tmp_path / "h"), capacity=4, frame_hw=(4, 6), fmt="float16")
    assert c.frames.dtype == np.float16 and c.frames.shape == (4, 4, 6)


def test_u8_over_255_without_division_is_exact():
    """attn_cache.cu's u8_over_255 (q = k * fl(1/255); q += fma(-q, 255, k) * fl(1/255)) restated in exact rational
    arithmetic: equals the correctly rounded k / 255 of ToTensor for every byte."""
    from fractions import Fraction
    r = np.float32(1.0) / np.float32(255.0)
    assert float(r) == 0.0039215688593685627

    def rn(fr):                                   # Fraction -> nearest fp32 (float() of a Fraction is correctly rounded to
        d = float(fr)                             # fp64; all values here have <= 53 significant bits before that step)
        assert Fraction(d) == fr
        return np.float32(d)

    for k in range(256):
        kf = np.float32(k)
        q = np.float32(kf * r)
        rem = rn(Fraction(float(kf)) - Fraction(float(q)) * 255)
        assert Fraction(float(rem)) == Fraction(float(kf)) - Fraction(float(q)) * 255      # the remainder is exact in fp32
        exact = Fraction(float(rem)) * Fraction(float(r)) + Fraction(float(q))
        lo = np.float32(float(exact))             # fp64 rounding of a 48+-bit sum may double-round: check both neighbours
        cands = [lo, np.nextafter(lo, np.float32(np.inf)), np.nextafter(lo, np.float32(-np.inf))]
        best = min(cands, key=lambda c: abs(Fraction(float(c)) - exact))
        assert best == kf / np.float32(255.0), k

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


def test_reopen_without_flush_keeps_index(tmp_path):
    """the miss path of AV_Dataset.get_frames_cached calls reserve() + write() and never flush(); the index must
    still be on disk, or a second run hands slot 0 to another video while the old validity bytes say "cached"."""
    path = str(tmp_path / "attn.u8")
    c = AttnFrameCache(path, capacity=32, frame_hw=(4, 4))
    assert c.reserve("vid_a", 8) == 0
    c.frames[0:8] = 11                    # what write() does after the device pack
    c.valid[0:8] = 1
    c.frames.flush()
    c.valid.flush()
    del c                                 # no flush() / close() by the caller

    r = AttnFrameCache(path)
    assert r.index == {"vid_a": (0, 8)} and r.used == 8
    assert r.has("vid_a", 0, 8) and int(r.frames[3, 0, 0]) == 11
    assert not r.has("vid_b", 0, 4)
    assert r.reserve("vid_b", 8) == 8     # a new video gets fresh slots, not slot 0
    assert not r.has("vid_b", 0, 8)


def test_reserve_clears_stale_validity(tmp_path):
    path = str(tmp_path / "attn.u8")
    c = AttnFrameCache(path, capacity=16, frame_hw=(4, 4))
    c.valid[:] = 1                        # stale bytes from a torn earlier run
    assert c.reserve("vid_a", 4) == 0
    assert not c.has("vid_a", 0, 4)


def test_checkpoint_helpers_match_reference_signatures(tmp_path):
    """utilities.save_checkpoint(model_dict, opt_dict, epoch, loss, name, dir) is called positionally by the reference
    trainers (train_audio_net.py:130); save_model always overwrites (utilities.py:162-166); load_checkpoint(model, optimizer,
    dir, auto, path, load_opt) (train.py:101).  A torch.optim.Adam checkpoint round-trips through the same layout."""
    import torch
    from maavss_b200 import utilities
    m = torch.nn.Linear(3, 2)
    opt = torch.optim.Adam(m.parameters(), lr=1e-3)
    m(torch.ones(1, 3)).sum().backward()
    opt.step()
    d = str(tmp_path / "cp")
    utilities.save_checkpoint(m.state_dict(), opt.state_dict(), 3, 0.25, "run-a", d)
    cp = torch.load(f"{d}/run-a.pt")
    assert set(cp.keys()) == {"epoch", "model_state_dict", "optimizer_state_dict", "loss"} and cp["epoch"] == 3
    m2 = torch.nn.Linear(3, 2)
    o2 = torch.optim.Adam(m2.parameters(), lr=1e-3)
    utilities.load_checkpoint(m2, o2, d, True, None, True)
    assert torch.equal(m2.weight, m.weight) and float(o2.state_dict()["state"][0]["step"]) == 1.0
    assert utilities.load_checkpoint(m2, o2, str(tmp_path / "empty"), True, None, False) is None
    p = str(tmp_path / "model.pt")
    utilities.save_model(p, m)
    utilities.save_model(p, m2)           # second save to the same path does not raise
    assert utilities.latest_file(d, "pt").endswith("run-a.pt")

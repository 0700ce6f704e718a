"""GPU parity of the raw attention-frame cache kernels (maavss_attn_cache_pack / _load) through the C ABI against
oracle/attn_cache.py and the torchvision-generated golden: the quantiser is integer work -> bit-exact; the
dequantise + diff + normalise is fp32 elementwise -> bit-exact as well (same operations in the same order)."""
import numpy as np
import pytest
import torch

import synth
from oracle import attn_cache as oc

pytestmark = pytest.mark.gpu


def _maps():
    maps = synth.attn_clip(2, 4, 64, 64, seed=61)[:, 0]
    maps[0, 0, 0, :8] = torch.tensor([0.0, 1.0, -0.3, 1.7, 0.5 / 255, 1.5 / 255, 127.5 / 255, 254.5 / 255])
    return maps


def test_pack_golden(golden):
    from maavss_b200 import attn_cache
    g = golden("attn_cache")
    u8 = attn_cache.pack(_maps().cuda(), "uint8").cpu()
    assert torch.equal(u8, torch.from_numpy(g["u8"]))
    f16 = attn_cache.pack(_maps().cuda(), "float16").cpu()
    assert torch.equal(f16, oc.pack_f16(_maps()))


@pytest.mark.parametrize("n", [0, 1, 3, 4, 5, 1023, 4096 + 7])
def test_pack_ragged(n):
    from maavss_b200 import attn_cache
    x = torch.rand(n, generator=synth.gen(n + 1)) * 1.2 - 0.1
    assert torch.equal(attn_cache.pack(x.cuda(), "uint8").cpu(), oc.pack_u8(x))
    assert torch.equal(attn_cache.pack(x.cuda(), "float16").cpu(), oc.pack_f16(x))


def test_pack_every_tie():
    """k + 0.5 over 255 for every k, and its fp32 neighbours: the rounding of x*255 and of +0.5 is save_image's."""
    from maavss_b200 import attn_cache
    k = (torch.arange(0, 256, dtype=torch.float32) + 0.5) / 255
    x = torch.cat([k, torch.nextafter(k, torch.zeros_like(k)), torch.nextafter(k, torch.ones_like(k) * 2),
                   torch.tensor([float("inf"), -float("inf"), 255.0, -0.0])])
    assert torch.equal(attn_cache.pack(x.cuda(), "uint8").cpu(), oc.pack_u8(x))


@pytest.mark.parametrize("fmt", ["uint8", "float16"])
@pytest.mark.parametrize("diff", [False, True])
@pytest.mark.parametrize("shape", [(2, 4, 64, 64), (3, 5, 7, 9), (1, 1, 16, 16), (1, 12, 256, 256)])
def test_load_matches_oracle(fmt, diff, shape):
    from maavss_b200 import attn_cache
    maps = torch.rand(*shape, generator=synth.gen(sum(shape)))
    stored = oc.pack_u8(maps) if fmt == "uint8" else oc.pack_f16(maps)
    if diff and shape[1] == 1:
        pytest.skip("a one-frame clip differenced is all zeros: 1/max = inf in the reference too")
    out = attn_cache.load(stored.cuda(), diff).cpu()
    ref = oc.load_clips(stored, diff)
    assert torch.equal(out, ref)


@pytest.mark.parametrize("fmt", ["uint8", "float16"])
def test_full_batch_properties(fmt):
    """Config 4's whole batch (256 clips x 12 x 256 x 256, 805 MB fp32: past the oracle's reach) through properties
    that need no second implementation: (1) with a 1.0 in every clip the un-differenced load is the plain dequantiser,
    so packing its output again returns the stored bytes (idempotence); (2) the differenced load telescopes --
    sum over frames = (last - first frame) / max -- and its per-clip maximum is 1 (fp32: v * (1/v) is within one ulp)."""
    from maavss_b200 import attn_cache
    clips, T, H, W = 256, 12, 256, 256
    maps = torch.rand(clips, T, H, W, device="cuda", generator=torch.Generator("cuda").manual_seed(4))
    maps[:, 3, 7, 9] = 1.0
    stored = attn_cache.pack(maps, fmt)
    del maps
    plain = attn_cache.load(stored, False)
    assert torch.equal(attn_cache.pack(plain, fmt), stored)
    assert float(plain.amax()) == 1.0 and float(plain.amin()) >= 0.0
    d = attn_cache.load(stored, True)
    assert float(d[:, 0].abs().max()) == 0.0
    peak = d.amax(dim=(1, 2, 3))
    assert float((peak - 1).abs().max()) <= 2 ** -23
    # telescoping: sum_t d[t] * max = plain[T-1] - plain[0]; recover max from one pixel pair per clip
    tele = d.sum(dim=1)
    span = plain[:, -1] - plain[:, 0]
    del d
    big = span.flatten(1).abs().argmax(dim=1)
    idx = torch.arange(clips, device="cuda")
    scale = span.flatten(1)[idx, big] / tele.flatten(1)[idx, big]           # = the clip's max of the differenced frames
    assert float((tele * scale.view(-1, 1, 1) - span).abs().max()) < 2e-5


def test_load_unaligned_view():
    """A clip batch that starts at an odd byte of the stored buffer takes the scalar path; same numbers."""
    from maavss_b200 import _lib, attn_cache
    maps = torch.rand(2, 3, 8, 8, generator=synth.gen(5))
    stored = oc.pack_u8(maps)
    buf = torch.zeros(stored.numel() + 1, dtype=torch.uint8, device="cuda")
    buf[1:] = stored.flatten().cuda()
    out = torch.empty(2, 3, 8, 8, device="cuda")
    ws = torch.empty(2, dtype=torch.int32, device="cuda")
    _lib.check(_lib.lib().maavss_attn_cache_load(buf.data_ptr() + 1, 0, out.data_ptr(), 2, 3, 64, 1, ws.data_ptr(),
                                                 _lib.stream_ptr()))
    assert torch.equal(out.cpu(), oc.load_clips(stored, True))


def test_load_equals_uncached_path_on_fp16_exact_maps():
    """Maps that fp16 holds exactly: the cached reader returns what attn_ops.diff_normalise returns on the raw maps."""
    from maavss_b200 import attn_cache, attn_ops
    maps = (torch.randint(0, 1024, (8, 1, 32, 32), generator=synth.gen(9)).float() / 1024).cuda()
    ref = attn_ops.diff_normalise(maps, True)
    out = attn_cache.load(attn_cache.pack(maps, "float16").reshape(1, 8, 32, 32), True)
    assert torch.equal(out.reshape(ref.shape), ref)


def test_errors():
    from maavss_b200 import _lib, attn_cache
    x = torch.rand(16, device="cuda")
    with pytest.raises(RuntimeError, match="unknown format"):
        _lib.check(_lib.lib().maavss_attn_cache_pack(x.data_ptr(), x.data_ptr(), 16, 7, _lib.stream_ptr()))
    with pytest.raises(RuntimeError):
        attn_cache.load(torch.zeros(2, 2, 2, 2, device="cuda"), False)           # fp32 is not a cache format
    with pytest.raises(RuntimeError):
        attn_cache.pack(torch.rand(4), "uint8")                                  # CPU tensor: no fallback


def test_cache_write_read_roundtrip(tmp_path):
    from maavss_b200.attn_cache import AttnFrameCache
    c = AttnFrameCache(str(tmp_path / "attn.u8"), capacity=64, frame_hw=(32, 32))
    vid = torch.rand(20, 1, 32, 32, generator=synth.gen(21))
    c.reserve("v0", 20)
    c.write("v0", 0, vid[:12].cuda())
    c.write("v0", 12, vid[12:].cuda())
    assert c.has("v0", 4, 12)
    got = c.read([("v0", 0, 12), ("v0", 8, 12)], attn_diff=True).cpu()      # overlapping clips share stored frames
    stored = oc.pack_u8(vid[:, 0])
    ref = oc.load_clips(torch.stack([stored[0:12], stored[8:20]]), True)
    assert got.shape == (2, 1, 12, 32, 32) and torch.equal(got[:, 0], ref)
    assert float((got[:, 0] - oc.load_clips(torch.stack([vid[0:12, 0], vid[8:20, 0]]).half(), True)).abs().max()) < 2e-2
    c.flush()
    r = AttnFrameCache(str(tmp_path / "attn.u8"), mode="r")
    assert torch.equal(r.read([("v0", 8, 12)], attn_diff=False).cpu()[0, 0], oc.load_clips(stored[None, 8:20], False)[0])
    with pytest.raises(KeyError):
        r.read([("v1", 0, 12)], attn_diff=False)


def test_dataset_cached_example(tmp_path):
    """AV_Dataset.get_frames_cached (av_dataset.py:251-278): a miss runs the extractor and fills the cache, the next call
    is a hit; the hit equals the miss up to the uint8 step, cache_ratio counts both."""
    from maavss_b200.attn_cache import AttnFrameCache
    from maavss_b200.av_dataset import AV_Dataset

    class Extractor:                                   # stands in for VideoAttention: maps = a fixed function of the frames
        calls = 0

        def inference_device(self, video):
            Extractor.calls += 1
            m = video.cuda().mean(dim=1, keepdim=True)
            return m / m.amax(dim=(2, 3), keepdim=True)

    def source(idx):
        g = synth.gen(100 + idx)
        return torch.rand(6, 3, 32, 32, generator=g), torch.randn(8 * 66 * 6, generator=g) * 0.1

    cache = AttnFrameCache(str(tmp_path / "a.u8"), capacity=32, frame_hw=(32, 32))
    ds = AV_Dataset(num_frames=6, fft_len=256, attn_diff=True, clip_source=source, num_clips=2,
                    attention_extractor=Extractor(), attn_cache=cache)
    miss, video = ds.get_frames_cached(1)
    hit, zeros = ds.get_frames_cached(1)
    assert ds.cache_ratio == [1, 1] and Extractor.calls == 1
    assert miss.shape == hit.shape == (1, 6, 32, 32) and float(zeros.abs().max()) == 0.0
    assert float((miss - hit).abs().max()) < 4e-2 and float(hit.max()) == pytest.approx(1.0, abs=1e-6)
    x, y, attn, audio, _ = ds.gen_av_example_cached(1)
    assert ds.cache_ratio == [2, 1] and torch.equal(attn, hit) and x.shape == y.shape

"""GPU parity: audio front end (STFT / iSTFT / mag-phase / generator helpers) through the C ABI vs the oracle."""
import numpy as np
import pytest
import torch

import synth
from oracle import audio as oa

pytestmark = pytest.mark.gpu


def t(a):
    return torch.from_numpy(np.asarray(a))


@pytest.fixture(scope="module")
def ops():
    from maavss_b200 import audio_ops
    return audio_ops


def fft_tol(ref, n_fft):
    # fp32 FFT error grows ~ log2(N) * eps, stated relative to the largest bin
    return 2e-6 * float(ref.abs().max()) * np.log2(n_fft)


@pytest.mark.parametrize("tag,n,n_fft,hop", [("default", 6336, 256, 66), ("n512", 4000, 512, 128), ("odd", 3001, 256, 100)])
def test_stft_golden(ops, golden, tag, n, n_fft, hop):
    g = golden("audio")
    audio = synth.audio_clip(n, seed=11).cuda().reshape(1, -1)
    w = oa.hamming_window(n_fft)
    for trim, key in ((True, "_trim"), (False, "_notrim")):
        ref = t(g[tag + key])
        out = ops.stft_planes(audio, w, n_fft, hop, trim)[0].cpu()
        assert out.shape == ref.shape
        assert float((out - ref).abs().max()) < fft_tol(ref, n_fft)
        ref_i = t(g[tag + "_istft" + key])
        aud = ops.istft_audio(ref.cuda().unsqueeze(0), w, n_fft, hop, trim)[0].cpu()
        assert aud.shape == ref_i.shape
        assert float((aud - ref_i).abs().max()) < 3e-6 * np.log2(n_fft) * max(1.0, float(ref_i.abs().max()))
    y = ops.stft_planes(audio, w, n_fft, hop, True)
    ops.absmax_normalise_(y)
    torch.testing.assert_close(y[0].cpu(), t(g[tag + "_norm"]), rtol=0, atol=2e-5)
    pol = ops.stft_planes(audio, w, n_fft, hop, True, polar=True)[0].cpu()
    refp = t(g[tag + "_polar"]).transpose(1, 2)            # golden is (2, F, T)
    torch.testing.assert_close(pol[0], refp[0], rtol=0, atol=1e-5)
    big = refp[0] > 1e-3 * refp[0].max()                    # the angle of a ~zero bin is ill-conditioned
    d = (pol[1] - refp[1])[big]
    d = torch.minimum(d.abs(), (d.abs() - 2 * np.pi).abs())
    assert float(d.max()) < 1e-3


@pytest.mark.parametrize("n_fft", [64, 128, 256, 512, 1024, 2048, 4096])
@pytest.mark.parametrize("hop_div", [8, 4, 2])
def test_stft_istft_sweep_vs_oracle(ops, n_fft, hop_div):
    hop = n_fft // hop_div
    B, n = 3, 16000
    g = torch.Generator().manual_seed(n_fft * 7 + hop_div)
    audio = torch.randn(B, n, generator=g)
    w = oa.hamming_window(n_fft)
    for trim in (True, False):
        out = ops.stft_planes(audio.cuda(), w, n_fft, hop, trim).cpu()
        ref = torch.stack([oa.stft_planes(audio[b], w, n_fft, hop, trim) for b in range(B)])
        assert out.shape == ref.shape
        assert float((out - ref).abs().max()) < fft_tol(ref, n_fft)
        inv = ops.istft_audio(ref.cuda(), w, n_fft, hop, trim).cpu()
        ref_i = torch.stack([oa.istft(ref[b], w, n_fft, hop, trim) for b in range(B)])
        assert inv.shape == ref_i.shape
        assert float((inv - ref_i).abs().max()) < 3e-6 * np.log2(n_fft) * float(ref_i.abs().max())


def test_stft_index_exact_impulses(ops):
    """Frame placement / reflect padding: unit impulses at chosen positions give exactly
    window-valued spectra; compare against the float64 naive DFT of the oracle."""
    n_fft, hop, n = 256, 66, 1500
    w = oa.hamming_window(n_fft)
    for pos in (0, 1, 65, 66, 127, 128, 700, n - 2, n - 1):
        x = torch.zeros(n)
        x[pos] = 1.0
        out = ops.stft_planes(x.cuda().reshape(1, -1), w, n_fft, hop, False)[0].cpu()
        ref = oa.stft_naive(x, w, n_fft, hop)[:, :-1].permute(2, 1, 0).float()
        assert float((out - ref).abs().max()) < 2e-6
        # frames that do not cover the impulse must be exactly zero in both
        assert torch.equal(out.abs().amax(dim=(0, 2)) == 0, ref.abs().amax(dim=(0, 2)) == 0)


@pytest.mark.parametrize("n_fft,hop,B", [(1024, 256, 64), (4096, 512, 128), (4096, 2048, 96), (512, 64, 256), (256, 66, 300)])
def test_stft_large_batch_roundtrip(ops, n_fft, hop, B):
    """Full-size property test (config 5 shape: 10 s clips, batches large enough that a CTA walks several frame
    batches): istft(stft(x)) == gain * x away from the edges."""
    n = 160000
    audio = torch.randn(B, n, generator=torch.Generator().manual_seed(1)).cuda()
    w = oa.hamming_window(n_fft)
    planes = ops.stft_planes(audio, w, n_fft, hop, False)
    assert planes.shape == (B, 2, n // hop, n_fft // 2 + 1)
    rec = ops.istft_audio(planes, w, n_fft, hop, False)
    gain = (n_fft ** 0.5) / float(w.pow(2).sum().sqrt())
    L = rec.shape[1]
    s = slice(n_fft, L - n_fft)
    assert float((rec[:, s] - gain * audio[:, :L][:, s]).abs().max()) < 5e-5 * gain * float(audio.abs().max())
    # linearity
    p2 = ops.stft_planes(2.5 * audio[:4], w, n_fft, hop, False)
    torch.testing.assert_close(p2, 2.5 * planes[:4], rtol=1e-5, atol=1e-5)


def test_empty_and_tiny(ops):
    w = oa.hamming_window(256)
    out = ops.stft_planes(torch.zeros(0, 1000).cuda(), w, 256, 66, True)
    assert out.shape == (0, 2, 15, 128)
    with pytest.raises(RuntimeError):
        ops.stft_planes(torch.zeros(1, 100).cuda(), w, 256, 66, True)     # reflect pad needs n > n_fft/2
    with pytest.raises(RuntimeError):
        ops.stft_planes(torch.zeros(1, 1000).cuda(), oa.hamming_window(100), 100, 25, True)   # not a power of two
    with pytest.raises(RuntimeError):
        ops.stft_planes(torch.zeros(1, 1000), w, 256, 66, True)           # CPU tensor: no fallback


def test_magphase_and_noise(ops):
    x, y = synth.stft_batch(3, 40, 129, seed=5)
    mp = ops.magphase_split(x.cuda()).cpu()
    ref = oa.magphase_split(x)
    torch.testing.assert_close(mp[:, 0], ref[:, 0], rtol=1e-6, atol=1e-7)
    torch.testing.assert_close(mp[:, 1], ref[:, 1], rtol=0, atol=2e-6)
    torch.testing.assert_close(ops.magphase_combine(mp.cuda()).cpu(), x, rtol=1e-5, atol=2e-6)
    noise = torch.randn(x.shape, generator=synth.gen(9))
    torch.testing.assert_close(ops.add_noise(y.cuda(), noise.cuda(), 0.1).cpu(), oa.add_noise(y, noise, 0.1), rtol=0, atol=1e-7)


def test_generator_helpers(golden):
    from maavss_b200.generator import DataGenerator
    g = golden("generator")
    dg = DataGenerator(batch_size=2)
    assert dg.fft_len == 2133
    audio = torch.randn(3, 2133, generator=synth.gen(23)) * 0.1
    f = dg.fft(audio)
    ref = torch.view_as_complex(t(g["fft"]).contiguous())
    assert f.shape == ref.shape
    assert float((f - ref).abs().max()) < 1e-5 * float(ref.abs().max())
    ri = dg.complex_to_ri(ref)
    assert torch.equal(ri, t(g["ri"]))
    cen = dg.center_fft_bins(ri)
    assert torch.equal(cen, t(g["centered"]))            # pure index permutation: bit exact
    dec = dg.decenter_fft_bins(cen)
    assert torch.equal(dec, t(g["decentered"]))
    back = dg.ifft(dg.ri_to_complex(dec))
    refi = t(g["ifft"])
    assert back.shape == refi.shape == (3, 2132)
    assert float((back - refi).abs().max()) < 1e-5 * float(refi.abs().max())
    pol = dg.cartesian_to_polar(ref)
    refp = t(g["polar"])
    torch.testing.assert_close(pol[:, :1066], refp[:, :1066], rtol=1e-5, atol=1e-8)
    d = (pol[:, 1066:] - refp[:, 1066:]).abs()
    assert float(torch.minimum(d, (d - 2 * np.pi).abs()).max()) < 1e-4
    # odd K centre: bit exact vs oracle
    x = torch.randn(2, 2, 7)
    assert torch.equal(dg.center_fft_bins(x), oa.center_fft_bins(x))


def test_av_dataset_methods(golden):
    from maavss_b200.av_dataset import AV_Dataset
    g = golden("audio")
    ds = AV_Dataset(num_frames=12, fft_len=256, hops_per_frame=8, noise_std=0.1, normalize_output_fft=False)
    assert (ds.hop, ds.audio_sample_len) == (66, 6336)
    audio = synth.audio_clip(6336, seed=11)
    spec = ds.stft(audio)                                    # reference layout (F', T', 2), CPU in -> CPU out
    assert spec.shape == (128, 96, 2) and not spec.is_cuda
    y = spec.permute(2, 1, 0)
    assert y.is_contiguous()
    ref = t(g["default_trim"])
    assert float((y - ref).abs().max()) < fft_tol(ref, 256)
    aud = ds.istft(ref)
    assert not aud.is_cuda and aud.shape == (6270,)
    assert float((aud - t(g["default_istft_trim"])).abs().max()) < 3e-5
    x, yb = ds.gen_stft_batch(audio.reshape(1, -1), noise=torch.zeros(1, 2, 96, 128))
    assert torch.equal(x, yb)


def test_attn_diff_normalise():
    from maavss_b200 import attn_ops
    from oracle import vit as ov
    a = torch.rand(5, 1, 32, 48, generator=synth.gen(3))
    for diff in (False, True):
        out = attn_ops.diff_normalise(a.cuda(), diff).cpu()
        ref = ov.attn_diff_normalise(a, diff).permute(1, 0, 2, 3)
        torch.testing.assert_close(out, ref, rtol=1e-6, atol=1e-7)
    # batched clips: per-clip max, signed-max quirk (negative max flips sign) reproduced
    b = torch.rand(3 * 4, 1, 8, 8, generator=synth.gen(4))
    out = attn_ops.diff_normalise(b.cuda(), True, clips=3).cpu()
    for c in range(3):
        ref = ov.attn_diff_normalise(b[4 * c:4 * c + 4], True).permute(1, 0, 2, 3)
        torch.testing.assert_close(out[4 * c:4 * c + 4], ref, rtol=1e-6, atol=1e-7)
    dec = torch.stack([torch.full((1, 4, 4), 1.0 - 0.1 * i) for i in range(3)])   # strictly decreasing clip
    dec[0, 0, 0, 0] = 5.0
    out = attn_ops.diff_normalise(dec.cuda(), True).cpu()
    ref = ov.attn_diff_normalise(dec, True).permute(1, 0, 2, 3)
    torch.testing.assert_close(out, ref, rtol=1e-6, atol=1e-7, equal_nan=True)


@pytest.mark.parametrize("n_fft,hop,clips,n", [(4096, 512, 128, 40000), (4096, 2048, 160, 50001), (2048, 256, 128, 30000)])
def test_stft_large_nfft_many_clips_canary(n_fft, hop, clips, n):
    """Round-1 "known issue" (n_fft 4096 forward kernel, illegal address at >= 128 clips): the multi-batch-per-CTA loop with
    two CTAs per SM.  Root cause and fix are in stft.cu (group_sync: immediate named-barrier ids).  The outputs sit between
    canaries, every clip is checked against torch.stft, and the inverse round trip must reproduce the 1.5863... gain."""
    from maavss_b200 import audio_ops
    g = torch.Generator().manual_seed(n_fft + clips)
    audio = (torch.randn(clips, n, generator=g) * 0.1).cuda()
    w = torch.hamming_window(n_fft)
    T, F = audio_ops.stft_out_dims(n, n_fft, hop, False)
    pad = 4096
    raw = torch.full((clips * 2 * T * F + 2 * pad,), float("nan"), device="cuda")
    planes = raw[pad:-pad].view(clips, 2, T, F)
    for _ in range(3):                                   # repeated launches: the fault was timing dependent
        audio_ops.stft_planes(audio, w, n_fft, hop, False, out=planes)
    torch.cuda.synchronize()
    assert bool(torch.isnan(raw[:pad]).all()) and bool(torch.isnan(raw[-pad:]).all())
    assert not bool(torch.isnan(planes).any())
    wd = w.cuda()
    ref = torch.stft(audio, n_fft, hop, n_fft, wd, center=True, pad_mode="reflect", return_complex=True) / wd.pow(2).sum().sqrt()
    ref = torch.view_as_real(ref)[:, :, :-1].permute(0, 3, 2, 1)          # (clips, 2, T, F)
    err = (planes - ref).abs().amax(dim=(1, 2, 3)) / ref.abs().amax(dim=(1, 2, 3))
    assert float(err.max()) < 2e-5, (int(err.argmax()), float(err.max()))
    rec = audio_ops.istft_audio(planes, w, n_fft, hop, False)
    gain = float(n_fft ** 0.5 / wd.pow(2).sum().sqrt())
    m = n_fft
    assert float((rec[:, m:-m] - gain * audio[:, m:rec.shape[1] - m]).abs().max()) < 2e-5 * gain

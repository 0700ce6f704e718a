"""Pin the oracle (CPU restatement) against outputs of the reference itself
(tests/golden/*.npz, made by tests/golden/make_golden.py from /root/reference)."""
import numpy as np
import pytest
import torch

import synth
from oracle import audio as oa
from oracle import models as om
from oracle import phasegram as op
from oracle import vit as ov


def t(a):
    return torch.from_numpy(np.asarray(a))


def test_hop_bit_exact(golden):
    for nf, hpf, fps, sr, hop, alen, nfr in golden("hop")["table"]:
        assert oa.calc_hop_size(int(nf), int(hpf), int(fps), int(sr)) == (hop, alen, nfr)
    assert oa.calc_hop_size(8, 8, 30, 16000) == (66, 4224, 64)
    assert oa.calc_hop_size(12, 8, 30, 16000) == (66, 6336, 96)


@pytest.mark.parametrize("tag,n,n_fft,hop", [("default", 6336, 256, 66), ("n512", 4000, 512, 128), ("odd", 3001, 256, 100)])
def test_stft_istft(golden, tag, n, n_fft, hop):
    g = golden("audio")
    audio, w = synth.audio_clip(n, seed=11), oa.hamming_window(n_fft)
    y_trim = oa.stft_planes(audio, w, n_fft, hop, True)
    y_notrim = oa.stft_planes(audio, w, n_fft, hop, False)
    assert torch.equal(y_trim, t(g[tag + "_trim"]))
    assert torch.equal(y_notrim, t(g[tag + "_notrim"]))
    assert torch.equal(oa.stft_planes(audio, w, n_fft, hop, True, normalize_output=True), t(g[tag + "_norm"]))
    assert torch.equal(oa.istft(y_trim, w, n_fft, hop, True), t(g[tag + "_istft_trim"]))
    assert torch.equal(oa.istft(y_notrim, w, n_fft, hop, False), t(g[tag + "_istft_notrim"]))
    pol = oa.stft(audio, w, n_fft, hop, True, use_polar=True)
    torch.testing.assert_close(pol, t(g[tag + "_polar"]), rtol=0, atol=1e-6)
    # independent float64 O(N^2) DFT: frame placement, reflect padding, normalisation
    naive = oa.stft_naive(audio, w, n_fft, hop)[:, :-1].permute(2, 1, 0)
    torch.testing.assert_close(naive.float(), y_notrim, rtol=0, atol=2e-6)
    # round-trip gain sqrt(n_fft)/sqrt(sum w^2), away from the clip edges
    rt = oa.istft(y_notrim, w, n_fft, hop, False)
    gain = (n_fft ** 0.5) / float(w.pow(2).sum().sqrt())
    s = slice(n_fft, rt.numel() - n_fft)
    torch.testing.assert_close(rt[s], audio[: rt.numel()][s] * gain, rtol=0, atol=2e-5)


def test_magphase_roundtrip():
    x, _ = synth.stft_batch(2, 16, 33, seed=3)
    mp = oa.magphase_split(x)
    torch.testing.assert_close(oa.magphase_combine(mp), x, rtol=1e-5, atol=1e-6)


def test_generator_helpers(golden):
    g = golden("generator")
    audio = torch.randn(3, 2133, generator=synth.gen(23)) * 0.1
    f = oa.gen_fft(audio)
    assert torch.equal(torch.view_as_real(f), t(g["fft"]))
    ri = oa.complex_to_ri(f)
    assert torch.equal(ri, t(g["ri"]))
    cen = oa.center_fft_bins(ri)
    assert torch.equal(cen, t(g["centered"]))
    dec = oa.decenter_fft_bins(cen)
    assert torch.equal(dec, t(g["decentered"])) and torch.equal(dec, ri)
    assert torch.equal(oa.gen_ifft(oa.ri_to_complex(dec)), t(g["ifft"]))
    assert torch.equal(oa.cartesian_to_polar(f), t(g["polar"]))


@pytest.mark.parametrize("tag,depth,T,H,W", [("d12_64", 12, 3, 64, 72), ("d2_256", 2, 2, 256, 256), ("d12_256", 12, 4, 256, 256)])
def test_vit_attention_maps(golden, tag, depth, T, H, W):
    g = golden("vit_" + tag)
    sd = synth.vit_state_dict(depth=depth)
    fr = synth.frames(T, H, W, seed=13)
    with torch.no_grad():
        maps = ov.attention_maps(sd, fr)
        last = ov.last_selfattention(sd, fr[:1, :, : H - H % 8, : W - W % 8])
    torch.testing.assert_close(maps[:, :, ::8, ::8], t(g["maps"]), rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(maps[0, 0], t(g["maps_full_frame0"]).float(), rtol=2e-3, atol=1e-3)
    torch.testing.assert_close(last[0, :, 0, :], t(g["cls_row"]), rtol=1e-4, atol=1e-6)
    # the x8 nearest upsample really is blockwise constant
    assert torch.equal(maps[:, :, ::8, ::8].repeat_interleave(8, 2).repeat_interleave(8, 3), maps)
    d = ov.attn_diff_normalise(maps, True).permute(1, 0, 2, 3)
    torch.testing.assert_close(d[:, :, ::8, ::8], t(g["diffnorm"]), rtol=1e-3, atol=2e-4)
    p = ov.attn_diff_normalise(maps, False).permute(1, 0, 2, 3)
    torch.testing.assert_close(p[:, :, ::8, ::8], t(g["plainnorm"]), rtol=1e-4, atol=1e-5)


def test_phasegram(golden):
    g = golden("phasegram")
    attn = synth.attn_clip(4, 8, 256, 256, seed=17)
    pg = op.video_phasegram(attn, resize=(64, 64))
    # atan2 of near-zero imaginary parts can flip by 2*pi between runs; compare robustly
    ref = t(g["pgram"])
    assert pg.shape == ref.shape
    assert float((pg - ref).abs().median()) < 1e-6
    assert float(((pg - ref).abs() > 1e-3).float().mean()) < 1e-3
    pn = op.video_phasegram(attn[:2], resize=(64, 64), diff=False, cumulative=False)
    refn = t(g["pgram_nodiff"])
    assert float(((pn - refn).abs() > 1e-4).float().mean()) < 1e-3


def _check_model(g, model, x_a, x_v, y_a, y_v, names_file, sub=None):
    model.eval()
    with torch.no_grad():
        a, v, fz = model(x_a, x_v)
    torch.testing.assert_close(a, t(g["eval_a"]), rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(v if sub is None else v[..., ::sub], t(g["eval_v"]), rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(fz, t(g["eval_fused"]), rtol=1e-4, atol=1e-5)
    model.train()
    loss, a_loss, v_loss = om.micro_step_loss(model, x_a, x_v, y_a, y_v)
    loss.backward()
    np.testing.assert_allclose([float(loss), float(a_loss), float(v_loss)], g["losses"], rtol=1e-5)
    names, norms, samples = synth.grad_digest(model)
    assert names == open(names_file).read().split()
    np.testing.assert_allclose(norms, g["grad_norms"], rtol=2e-3, atol=1e-9)
    np.testing.assert_allclose(samples, g["grad_samples"], rtol=5e-3, atol=1e-7)
    bn = np.concatenate([x.flatten().numpy() for k, x in model.state_dict().items()
                         if k.endswith("running_mean") or k.endswith("running_var")])
    np.testing.assert_allclose(bn, g["bn_running"], rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("variant", ["fc", "conv"])
def test_fusion_model(golden, variant):
    import os
    from conftest import GOLDEN
    g = golden("fusion_" + variant)
    x_a, y_a = synth.stft_batch(2, 64, 128, seed=19)
    x_v = torch.randn(2, 1, 8, 4096, generator=synth.gen(29)) * 0.3
    m = om.AVFusionOracle([2, 2, 64, 128], [2, 1, 8, 4096], 8, variant=variant)
    m.load_state_dict(synth.fill_state_dict(m.state_dict(), seed=31))
    _check_model(g, m, x_a, x_v, y_a, x_v, os.path.join(GOLDEN, f"fusion_{variant}_params.txt"))


def test_fusion_train_step(golden):
    g = golden("fusion_fc_step")
    m = om.AVFusionOracle([2, 2, 64, 128], [2, 1, 8, 4096], 8)
    m.load_state_dict(synth.fill_state_dict(m.state_dict(), seed=31))
    m.train()
    opt = torch.optim.Adam(m.parameters(), lr=1e-5)
    xs, ys = synth.stft_batch(2, 96, 128, seed=37)
    pg = torch.randn(2, 1, 12, 4096, generator=synth.gen(41)) * 0.3
    losses = []
    for _ in range(2):
        for j in range(4):
            loss, _, _ = om.micro_step_loss(m, xs[:, :, 8 * j:8 * j + 64], pg[:, :, j:j + 8],
                                            ys[:, :, 8 * j:8 * j + 64], pg[:, :, j:j + 8])
            loss.backward()
            losses.append(float(loss))
        opt.step()
        opt.zero_grad()
    np.testing.assert_allclose(losses, g["losses"], rtol=1e-5)
    np.testing.assert_allclose(synth.param_digest(m), g["params"], rtol=1e-5, atol=1e-7)


def test_frames_model(golden):
    import os
    from conftest import GOLDEN
    g = golden("frames")
    xf, yf = synth.stft_batch(2, 64, 129, seed=43)
    fr = synth.attn_clip(2, 8, 256, 256, seed=47)
    m = om.AVFusionFramesOracle([2, 2, 64, 129], [2, 1, 8, 256, 256], 8)
    m.load_state_dict(synth.fill_state_dict(m.state_dict(), seed=53))
    _check_model(g, m, xf, fr, yf[:, :, 32:40], fr[:, :, 4], os.path.join(GOLDEN, "frames_params.txt"))


def _attn_cache_maps():
    maps = synth.attn_clip(2, 4, 64, 64, seed=61)[:, 0]
    maps[0, 0, 0, :8] = torch.tensor([0.0, 1.0, -0.3, 1.7, 0.5 / 255, 1.5 / 255, 127.5 / 255, 254.5 / 255])
    return maps


def test_attn_cache_quantiser(golden):
    """oracle/attn_cache.py against torchvision's own save_image -> PNG -> ToTensor run by make_golden.py: bit-exact,
    including the clamp and the x.5 ties."""
    from oracle import attn_cache as oc
    g = golden("attn_cache")
    maps = _attn_cache_maps()
    u8 = oc.pack_u8(maps)
    assert torch.equal(u8, torch.from_numpy(g["u8"]))
    assert torch.equal(oc.unpack(u8), torch.from_numpy(g["totensor"]))
    assert float(g["raw_max_err"]) <= 0.5 / 255 + 1e-7 < float(g["jpeg_max_err"])    # what dropping JPEG buys
    out = oc.load_clips(u8, True)
    assert out.shape == (2, 4, 64, 64) and float(out[:, 0].abs().max()) == 0.0
    assert float(out.amax(dim=(1, 2, 3)).min()) == pytest.approx(1.0, abs=1e-6)


def test_flow_corner_restatement_matches_cv2_goldens(golden):
    """oracle/flow.py's numpy restatement of cv2.goodFeaturesToTrack (Shi-Tomasi, flow.py:11-14,28) against the cv2 outputs in
    tests/golden/flow.npz: identical corner lists (coordinates AND order) at the reference's quality 0.3 and at 0.01, where the
    100-corner cap and the minimum-distance rule decide; and against cv2 itself when it is importable."""
    from oracle import flow as of
    g = golden("flow")
    for seed in (1, 2, 3):
        a, b = synth.flow_frames(seed)
        eig = of.min_eig_map(a)
        assert abs(float(eig.max()) - float(g[f"s{seed}_eig_max"])) < 1e-6 * float(g[f"s{seed}_eig_max"])
        assert np.array_equal(of.corners_from_eig(eig), g[f"s{seed}_p0"])
        assert np.array_equal(of.corners_from_eig(eig, qualityLevel=0.01), g[f"s{seed}_dense"])
        try:
            import cv2  # noqa: F401
        except ImportError:
            continue
        assert np.array_equal(of.good_features(a), g[f"s{seed}_p0"])
        p1, st, err = of.pyr_lk(a, b, g[f"s{seed}_p0"])
        assert np.array_equal(st, g[f"s{seed}_st"]) and np.allclose(p1, g[f"s{seed}_p1"], atol=1e-4)

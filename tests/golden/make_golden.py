"""Generate tests/golden/*.npz by running the REFERENCE code (/root/reference) on CPU.

Authoring-container only (the reference is not present on the GPU box).  Run:
    python tests/golden/make_golden.py
The reference modules import under a shim: torchsummary / matplotlib / wandb / cv2 stubbed,
"cuda" remapped to cpu in Module.to / Tensor.to (model constructors hard-code .to("cuda"),
avse_model.py:446 ff.).  Inputs and weights come from tests/golden/synth.py (seeded), so the
fixtures hold only reference OUTPUTS.
"""
import contextlib
import io
import os
import sys
import types
from functools import partial

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import synth  # noqa: E402

REF = "/root/reference"


def install_shim():
    for name in ["torchsummary", "matplotlib", "matplotlib.pyplot", "wandb"] + ([] if "flow" in sys.argv else ["cv2"]):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    sys.modules["torchsummary"].summary = lambda *a, **k: None
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    sys.path.insert(0, REF)
    sys.path.insert(0, os.path.join(REF, "dino"))

    def remap(args):
        return tuple("cpu" if (isinstance(x, str) and x.startswith("cuda")) else x for x in args)
    _mto, _tto = torch.nn.Module.to, torch.Tensor.to
    torch.nn.Module.to = lambda self, *a, **k: _mto(self, *remap(a), **k)
    torch.Tensor.to = lambda self, *a, **k: _tto(self, *remap(a), **k)


def quiet(fn, *a, **k):
    with contextlib.redirect_stdout(io.StringIO()):
        return fn(*a, **k)


def save(name, **arrs):
    out = {}
    for k, v in arrs.items():
        out[k] = v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else np.asarray(v)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, {k: v.shape for k, v in out.items()})


from synth import grad_digest, param_digest  # noqa: E402


def golden_hop():
    import utilities
    rows = []
    for nf in (1, 4, 8, 12, 64):
        for hpf in (1, 4, 8, 16):
            for fps in (24, 25, 30, 60):
                for sr in (8000, 16000, 22050, 44100, 48000):
                    rows.append((nf, hpf, fps, sr) + tuple(utilities.calc_hop_size(nf, hpf, fps, sr)))
    save("hop", table=np.array(rows, dtype=np.int64))


def golden_audio():
    import torchaudio
    cases = {}
    for tag, n, n_fft, hop in (("default", 6336, 256, 66), ("n512", 4000, 512, 128), ("odd", 3001, 256, 100)):
        audio = synth.audio_clip(n, seed=11)
        window = torch.hamming_window(n_fft)
        # the literal call of av_dataset.py:160-168
        spec = torchaudio.functional.spectrogram(audio, pad=0, window=window, n_fft=n_fft,
                                                 hop_length=hop, win_length=window.shape[0],
                                                 power=None, normalized=True, onesided=True)
        spec = torch.view_as_real(spec)                       # the layout the reference assumed
        trim = spec[:-1, :-1, :]                              # av_dataset.py:171-172
        notrim = spec[:, :-1, :]                              # av_dataset.py:174
        y_trim = trim.permute(2, 1, 0).contiguous()           # av_dataset.py:338
        y_notrim = notrim.permute(2, 1, 0).contiguous()
        y_norm = y_trim * (1 / torch.max(torch.abs(y_trim) + 1e-7))   # av_dataset.py:339-340
        # av_dataset.py:183-200 on the complex API
        import torch.nn.functional as F
        def inv(y, trimmed):
            s = F.pad(y, (0, 1)).permute(2, 1, 0) if trimmed else y.permute(2, 1, 0)
            return torch.istft(torch.view_as_complex(s.contiguous()), n_fft=n_fft, hop_length=hop,
                               win_length=n_fft, window=window, normalized=True, onesided=True)
        cases[tag + "_trim"] = y_trim
        cases[tag + "_notrim"] = y_notrim
        cases[tag + "_norm"] = y_norm
        cases[tag + "_istft_trim"] = inv(y_trim, True)
        cases[tag + "_istft_notrim"] = inv(y_notrim, False)
        c = torch.view_as_complex(trim.contiguous())
        cases[tag + "_polar"] = torch.stack((c.abs(), c.angle()), 0)  # magphase formulas :176-178
    save("audio", **cases)


def golden_generator():
    import generator
    dg = object.__new__(generator.DataGenerator)
    dg.normalize_input_fft = True
    audio = torch.randn(3, 2133, generator=synth.gen(23)) * 0.1
    f = dg.fft(audio)
    ri = dg.complex_to_ri(f)
    cen = dg.center_fft_bins(ri)
    dec = dg.decenter_fft_bins(cen)
    back = dg.ri_to_complex(dec)
    aud = dg.ifft(back)
    pol = dg.cartesian_to_polar(f)
    save("generator", fft=torch.view_as_real(f), ri=ri, centered=cen, decentered=dec,
         ifft=aud, polar=pol)


def golden_vit():
    import vision_transformer as vits
    import video_attention
    only = os.environ.get("GOLDEN_VIT_ONLY")
    for tag, depth, T, H, W in (("d12_64", 12, 3, 64, 72), ("d2_256", 2, 2, 256, 256), ("d12_256", 12, 4, 256, 256)):
        if only and tag != only:
            continue
        sd = synth.vit_state_dict(depth=depth)
        model = vits.VisionTransformer(patch_size=8, embed_dim=384, depth=depth, num_heads=6,
                                       mlp_ratio=4, qkv_bias=True, num_classes=0,
                                       norm_layer=partial(torch.nn.LayerNorm, eps=1e-6))
        model.load_state_dict(sd, strict=True)
        model.eval()
        va = object.__new__(video_attention.VideoAttention)
        va.patch_size, va.threshold, va.resize, va.model = 8, 0.6, None, model
        fr = synth.frames(T, H, W, seed=13)
        with torch.no_grad():
            maps = va._inference(fr)
            last = model.get_last_selfattention(fr[:1, :, : H - H % 8, : W - W % 8])
        # av_dataset.py:323-328
        attn = maps.clone()
        pad = torch.zeros_like(attn[0:1])
        d = torch.cat((pad, torch.diff(attn, dim=0)), 0)
        d *= 1 / torch.max(d)
        plain = maps.clone()
        plain *= 1 / torch.max(plain)
        save("vit_" + tag, maps=maps[:, :, ::8, ::8], maps_full_frame0=maps[0, 0].half(),
             cls_row=last[0, :, 0, :], diffnorm=d[:, :, ::8, ::8], plainnorm=plain[:, :, ::8, ::8])


def golden_phasegram():
    import utilities
    attn = synth.attn_clip(4, 8, 256, 256, seed=17)
    pg = utilities.video_phasegram(attn, resize=(64, 64))
    pg_nodiff = utilities.video_phasegram(attn[:2], resize=(64, 64), diff=False, cumulative=False)
    save("phasegram", pgram=pg, pgram_nodiff=pg_nodiff)


def golden_attn_cache():
    """Pins oracle/attn_cache.py: torchvision's own save_image quantiser (through lossless PNG) and ToTensor, plus
    the size of the JPEG error of the reference's writer (save_attn_videos.save_frames) that the raw cache removes."""
    import tempfile
    import torchvision
    from PIL import Image
    import save_attn_videos
    maps = synth.attn_clip(2, 4, 64, 64, seed=61)[:, 0]                    # (2, 4, 64, 64) in [0, 1)
    maps[0, 0, 0, :8] = torch.tensor([0.0, 1.0, -0.3, 1.7, 0.5 / 255, 1.5 / 255, 127.5 / 255, 254.5 / 255])
    to_tensor = torchvision.transforms.ToTensor()
    u8, back, jpg = [], [], []
    with tempfile.TemporaryDirectory() as d:
        for c, clip in enumerate(maps):
            for i, frame in enumerate(clip):
                path = os.path.join(d, f"f_{c}_{i}.png")
                torchvision.utils.save_image(frame[None].repeat(3, 1, 1), path)      # av_dataset.py:275-276
                img = Image.open(path)
                u8.append(np.asarray(img)[:, :, 0].copy())
                back.append(to_tensor(img)[0])
            os.mkdir(os.path.join(d, str(c)))
            quiet(save_attn_videos.save_frames, clip[None], os.path.join(d, str(c)))  # the reference writer (JPEG)
            for i in range(clip.shape[0]):
                jpg.append(to_tensor(Image.open(os.path.join(d, str(c), f"img_{i:05d}.jpg")))[0])
    u8 = np.stack(u8).reshape(2, 4, 64, 64)
    back = torch.stack(back).reshape(2, 4, 64, 64)
    jpg = torch.stack(jpg).reshape(2, 4, 64, 64)
    save("attn_cache", u8=u8, totensor=back, jpeg_max_err=float((jpg - maps.clamp(0, 1)).abs().max()),
         raw_max_err=float((back - maps.clamp(0, 1)).abs().max()))


def golden_flow():
    """cv2 (the arithmetic flow.py calls) on seeded synthetic frame pairs: corners, tracked points, status, error."""
    import cv2
    out = {}
    for seed in (1, 2, 3):
        a, b = synth.flow_frames(seed)
        p0 = cv2.goodFeaturesToTrack(a, mask=None, maxCorners=100, qualityLevel=0.3, minDistance=7, blockSize=7)       # flow.py:11-14,28
        p1, st, err = cv2.calcOpticalFlowPyrLK(a, b, p0, None, winSize=(15, 15), maxLevel=2,
                                               criteria=(cv2.TERM_CRITERIA_EPS | cv2.TERM_CRITERIA_COUNT, 10, 0.03))   # flow.py:17-19,37
        dense = cv2.goodFeaturesToTrack(a, mask=None, maxCorners=100, qualityLevel=0.01, minDistance=7, blockSize=7)
        out.update({f"s{seed}_p0": p0, f"s{seed}_p1": p1, f"s{seed}_st": st, f"s{seed}_err": err, f"s{seed}_dense": dense,
                    f"s{seed}_eig_max": np.float32(cv2.cornerMinEigenVal(a, 7, ksize=3).max())})
    rgb = np.random.RandomState(5).randint(0, 256, (32, 40, 3), dtype=np.uint8)
    out["gray"] = cv2.cvtColor(rgb, cv2.COLOR_BGR2GRAY)
    save("flow", **out)


def run_model_case(name, model, x_a, x_v, y_a, y_v, loss_coeff=0.001):
    import torch.nn.functional as F
    out = {}
    model.eval()
    with torch.no_grad():
        a, v, fz = model(x_a, x_v)
    out.update(eval_a=a, eval_v=v[..., ::4] if v.numel() > 200000 else v, eval_fused=fz)
    model.train()
    a, v, fz = model(x_a, x_v)
    a_loss, v_loss = F.mse_loss(a, y_a), F.mse_loss(v, y_v)
    loss = (a_loss + loss_coeff * v_loss) / 4
    loss.backward()
    names, norms, samples = grad_digest(model)
    out.update(train_a=a, train_fused=fz, losses=np.array([float(loss), float(a_loss), float(v_loss)]),
               grad_norms=norms, grad_samples=samples)
    bn = [(k, t) for k, t in model.state_dict().items() if k.endswith("running_mean") or k.endswith("running_var")]
    out["bn_running"] = np.concatenate([t.flatten().numpy() for _, t in bn])
    save(name, **out)
    with open(os.path.join(HERE, name + "_params.txt"), "w") as f:
        f.write("\n".join(names) + "\n")


def golden_models():
    import avse_model, avse_model_conv, avse_model_final
    B = 2
    x_a, y_a = synth.stft_batch(B, 64, 128, seed=19)
    x_v = torch.randn(B, 1, 8, 4096, generator=synth.gen(29)) * 0.3
    for tag, mod in (("fusion_fc", avse_model), ("fusion_conv", avse_model_conv)):
        m = quiet(mod.AV_Fusion_Model, [B, 2, 64, 128], [B, 1, 8, 4096], 8)
        m.load_state_dict(synth.fill_state_dict(m.state_dict(), seed=31))
        run_model_case(tag, m, x_a, x_v, y_a, x_v)
    # one full optimizer step (train.py:136-162) on the fc model, B=2, num_seq=4
    m = quiet(avse_model.AV_Fusion_Model, [B, 2, 64, 128], [B, 1, 8, 4096], 8)
    m.load_state_dict(synth.fill_state_dict(m.state_dict(), seed=31))
    m.train()
    opt = torch.optim.Adam(m.parameters(), lr=1e-5)
    xs, ys = synth.stft_batch(B, 96, 128, seed=37)
    pg = torch.randn(B, 1, 12, 4096, generator=synth.gen(41)) * 0.3
    losses = []
    mse = torch.nn.functional.mse_loss
    for step in range(2):
        for j in range(4):
            xb, yb = xs[:, :, 8 * j:8 * j + 64], ys[:, :, 8 * j:8 * j + 64]
            pb = pg[:, :, j:j + 8]
            a, v, _ = m(xb, pb)
            loss = (mse(a, yb) + 0.001 * mse(v, pb)) / 4
            loss.backward()
            losses.append(float(loss))
        opt.step()
        opt.zero_grad()
    save("fusion_fc_step", losses=np.array(losses), params=param_digest(m))

    xf, yf = synth.stft_batch(B, 64, 129, seed=43)
    fr = synth.attn_clip(B, 8, 256, 256, seed=47)
    m = quiet(avse_model_final.AV_Fusion_Model_Frames, [B, 2, 64, 129], [B, 1, 8, 256, 256], 8, latent_channels=16)
    m.load_state_dict(synth.fill_state_dict(m.state_dict(), seed=53))
    y_mid_a = yf[:, :, 32:40]
    y_mid_v = fr[:, :, 4]
    run_model_case("frames", m, xf, fr, y_mid_a, y_mid_v)


if __name__ == "__main__":
    torch.set_num_threads(8)
    install_shim()
    which = sys.argv[1:] or ["hop", "audio", "generator", "vit", "phasegram", "attn_cache", "models"]
    for w in which:
        globals()["golden_" + w]()

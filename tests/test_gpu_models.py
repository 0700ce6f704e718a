"""GPU parity of the fusion models (drop-in classes of maavss_b200) against the golden vectors produced by the
REFERENCE classes (tests/golden/make_golden.py) and against the oracle on the same seeded inputs: outputs, losses,
gradients of every parameter, BatchNorm running statistics, and two full optimizer steps (train.py:136-162)."""
import os

import numpy as np
import pytest
import torch

import synth
from conftest import GOLDEN
from oracle import models as om

pytestmark = pytest.mark.gpu


def t(a):
    return torch.from_numpy(np.asarray(a))


def grad_digest(model):
    names, norms, samples, seen = [], [], [], set()
    for k, p in model.named_parameters():
        if p.data_ptr() in seen:
            continue
        seen.add(p.data_ptr())
        names.append(k)
        if p.grad is None:
            norms.append(-1.0)
            samples.append(np.zeros(8, np.float32))
            continue
        gr = p.grad.detach().flatten().cpu()
        norms.append(float(gr.double().norm()))
        samples.append(gr[synth.sample_idx(gr.numel())].numpy().astype(np.float32))
    return names, np.array(norms), np.stack(samples)


def param_digest(model):
    out, seen = [], set()
    for k, p in model.named_parameters():
        if p.data_ptr() in seen:
            continue
        seen.add(p.data_ptr())
        f = p.detach().flatten().cpu()
        out.append(np.concatenate(([float(f.double().sum())], f[synth.sample_idx(f.numel())].double().numpy())))
    return np.stack(out)


def assert_digest_close(names, dig, ref, rtol, atol, cols=slice(None)):
    """np.testing.assert_allclose over param_digest rows, naming the offending parameters on failure."""
    a, b = np.asarray(dig)[:, cols], np.asarray(ref)[:, cols]
    bad = np.argwhere(~np.isclose(a, b, rtol=rtol, atol=atol))
    msg = "; ".join(f"{names[r]}[col {c}] {a[r, c]!r} vs {b[r, c]!r}" for r, c in bad[:12])
    assert len(bad) == 0, f"{len(bad)} digest entries differ: {msg}"


def micro_step_loss(model, x_a, x_v, y_a, y_v, loss_coeff=0.001, num_seq=4):
    from maavss_b200.nn_ops import mse_loss
    yh_a, yh_v, _ = model(x_a, x_v)
    a_loss, v_loss = mse_loss(yh_a, y_a), mse_loss(yh_v, y_v)
    return (a_loss + loss_coeff * v_loss) / num_seq, a_loss, v_loss


def check_model(g, model, x_a, x_v, y_a, y_v, names_file, sub=None, unused_zero=True, sample_rtol=5e-3, norm_rtol=2e-3):
    x_a, x_v, y_a, y_v = (z.cuda() for z in (x_a, x_v, y_a, y_v))
    model.eval()
    with torch.no_grad():
        a, v, fz = model(x_a, x_v)
    torch.testing.assert_close(a.cpu(), t(g["eval_a"]), rtol=1e-4, atol=1e-5)
    torch.testing.assert_close((v if sub is None else v[..., ::sub]).cpu(), t(g["eval_v"]), rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(fz.cpu(), t(g["eval_fused"]), rtol=1e-4, atol=1e-5)
    model.train()
    loss, a_loss, v_loss = micro_step_loss(model, x_a, x_v, y_a, y_v)
    loss.backward()
    np.testing.assert_allclose([float(loss), float(a_loss), float(v_loss)], g["losses"], rtol=1e-5)
    names, norms, samples = grad_digest(model)
    assert names == open(names_file).read().split()
    # conv biases that feed a BatchNorm have an exactly-zero gradient: both sides hold only rounding noise (~1e-9) there
    np.testing.assert_allclose(norms, g["grad_norms"], rtol=norm_rtol, atol=2e-8)
    if sample_rtol <= 5e-3:
        np.testing.assert_allclose(samples, g["grad_samples"], rtol=sample_rtol, atol=1e-7)
    else:           # relative to the largest sampled entry of each parameter
        ref_s = np.asarray(g["grad_samples"])
        assert (np.abs(samples - ref_s).max(axis=1) <= sample_rtol * np.abs(ref_s).max(axis=1) + 1e-7).all()
    bn = np.concatenate([x.flatten().cpu().numpy() for k, x in model.state_dict().items()
                         if k.endswith("running_mean") or k.endswith("running_var")])
    np.testing.assert_allclose(bn, g["bn_running"], rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("variant", ["fc", "conv"])
def test_fusion_model_golden(golden, variant):
    from maavss_b200 import avse_model, avse_model_conv
    g = golden("fusion_" + variant)
    x_a, y_a = synth.stft_batch(2, 64, 128, seed=19)
    x_v = torch.randn(2, 1, 8, 4096, generator=synth.gen(29)) * 0.3
    ref = om.AVFusionOracle([2, 2, 64, 128], [2, 1, 8, 4096], 8, variant=variant)
    sd = synth.fill_state_dict(ref.state_dict(), seed=31)
    cls = avse_model.AV_Fusion_Model if variant == "fc" else avse_model_conv.AV_Fusion_Model
    m = cls([2, 2, 64, 128], [2, 1, 8, 4096], 8)
    assert list(m.state_dict().keys()) == list(sd.keys())           # the state_dict layout IS the API
    m.load_state_dict(sd)
    assert next(m.parameters()).is_cuda
    check_model(g, m, x_a, x_v, y_a, x_v, os.path.join(GOLDEN, f"fusion_{variant}_params.txt"))


def test_frames_model_golden(golden):
    """avse_model_final.AV_Fusion_Model_Frames (Conv3d + BN3d + MaxPool3d + LeakyReLU visual encoder) vs the reference goldens."""
    from maavss_b200 import avse_model_final
    g = golden("frames")
    xf, yf = synth.stft_batch(2, 64, 129, seed=43)
    fr = synth.attn_clip(2, 8, 256, 256, seed=47)
    ref = om.AVFusionFramesOracle([2, 2, 64, 129], [2, 1, 8, 256, 256], 8)
    sd = synth.fill_state_dict(ref.state_dict(), seed=53)
    m = avse_model_final.AV_Fusion_Model_Frames([2, 2, 64, 129], [2, 1, 8, 256, 256], 8)
    assert list(m.state_dict().keys()) == list(sd.keys())
    m.load_state_dict(sd)
    # The CPU goldens of the three 2x2-pooled Conv3d blocks carry ~5e-3 of backend noise: the same oracle run through
    # torch on CUDA differs from them by the same amount (benchmarks/debug_frames2.py), so the golden check is held to
    # 1e-2 on gradient samples (5e-3 on gradient norms: 2.0e-3 was measured for one early visual-encoder parameter after
    # a change of summation order) and the tight check is made against the oracle executed on this GPU.
    check_model(g, m, xf, fr, yf[:, :, 32:40], fr[:, :, 4], os.path.join(GOLDEN, "frames_params.txt"), sample_rtol=1e-2, norm_rtol=5e-3)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    ref.load_state_dict(sd)
    ref = ref.cuda().train()
    loss, _, _ = om.micro_step_loss(ref, xf.cuda(), fr.cuda(), yf[:, :, 32:40].cuda(), fr[:, :, 4].cuda())
    loss.backward()
    for (k, p), (_, q) in zip(m.named_parameters(), ref.named_parameters()):
        if q.grad is not None:
            scale = max(float(q.grad.abs().max()), 1e-12)
            # arg-max flips of the 2x2 pools under 1e-6 forward noise move up to ~1e-2 of the early-layer gradients between backends
            assert float((p.grad - q.grad).abs().max()) < (2e-2 if k.startswith("visual_encoder") else 2e-4) * scale, k
    with pytest.raises(RuntimeError):
        avse_model_final.AV_Fusion_Model_Frames([2, 2, 64, 129], [2, 1, 8, 256, 256], 8, latent_channels=64)


def test_fusion_train_step_golden(golden):
    """Two optimizer steps of the reference loop (4 accumulation micro-steps each) with the flat-arena Adam."""
    from maavss_b200 import avse_model, optim
    g = golden("fusion_fc_step")
    ref = om.AVFusionOracle([2, 2, 64, 128], [2, 1, 8, 4096], 8)
    sd = synth.fill_state_dict(ref.state_dict(), seed=31)
    m = avse_model.AV_Fusion_Model([2, 2, 64, 128], [2, 1, 8, 4096], 8)
    m.load_state_dict(sd)
    m.train()
    opt = optim.Adam(m.parameters(), lr=1e-5)
    xs, ys = (z.cuda() for z in synth.stft_batch(2, 96, 128, seed=37))
    pg = (torch.randn(2, 1, 12, 4096, generator=synth.gen(41)) * 0.3).cuda()
    losses = []
    for _ in range(2):
        for j in range(4):
            loss, _, _ = micro_step_loss(m, xs[:, :, 8 * j:8 * j + 64].contiguous(), pg[:, :, j:j + 8].contiguous(),
                                         ys[:, :, 8 * j:8 * j + 64].contiguous(), pg[:, :, j:j + 8].contiguous())
            loss.backward()
            losses.append(float(loss))
        opt.step()
        opt.zero_grad()
    np.testing.assert_allclose(losses, g["losses"], rtol=1e-5)
    # Conv biases that feed a BatchNorm have an exactly-zero gradient; what reaches Adam is rounding noise (~1e-9, below
    # eps = 1e-8), which Adam turns into a noise-dependent move of up to lr per step -- in the reference as well.  Those
    # rows get an absolute tolerance of 2 steps * lr; every other parameter is held to 1e-5 relative.
    names = [k for k, _ in m.named_parameters()]
    names = [k for i, k in enumerate(names) if k not in names[:i]]
    mods = dict(m.named_modules())
    noisy = np.array([k.endswith(".bias") and type(mods[k.rsplit(".", 1)[0]]).__name__.startswith("Conv") for k in names])
    dig, ref_dig = param_digest(m), g["params"]
    assert dig.shape[0] == len(names) == noisy.shape[0]
    np.testing.assert_allclose(dig[~noisy], ref_dig[~noisy], rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(dig[noisy][:, 1:], ref_dig[noisy][:, 1:], rtol=0, atol=2.5e-5)


def test_batched_windows_equal_sequential():
    """FusionTrainStep(batch_windows=True) == the reference's sequential accumulation loop (train.py:136-162): losses,
    parameters after two optimizer steps and BatchNorm running statistics."""
    from maavss_b200 import avse_model, optim, train_step
    ref = om.AVFusionOracle([2, 2, 64, 128], [2, 1, 8, 4096], 8)
    sd = synth.fill_state_dict(ref.state_dict(), seed=31)
    xs, ys = (z.cuda() for z in synth.stft_batch(2, 96, 128, seed=37))
    attn = synth.attn_clip(2, 12, 256, 256, seed=17).cuda()
    res = []
    for batched in (False, True):
        m = avse_model.AV_Fusion_Model([2, 2, 64, 128], [2, 1, 8, 4096], 8)
        m.load_state_dict(sd)
        m.train()
        opt = optim.Adam(m.parameters(), lr=1e-5)
        st = train_step.FusionTrainStep(m, opt, graph=False, batch_windows=batched)
        losses = [[float(v) for v in st(xs, ys, attn)] for _ in range(2)]
        res.append((losses, param_digest(m), {k: v.clone() for k, v in m.state_dict().items() if "running" in k or "tracked" in k}, m))
    (l0, p0, bn0, m0), (l1, p1, bn1, m1) = res
    np.testing.assert_allclose(l1, l0, rtol=2e-5)
    names = [k for k, _ in m0.named_parameters()]
    names = [k for i, k in enumerate(names) if k not in names[:i]]
    mods = dict(m0.named_modules())
    noisy = np.array([k.endswith(".bias") and type(mods[k.rsplit(".", 1)[0]]).__name__.startswith("Conv") for k in names])
    nm = np.array(names)
    # sampled entries tight.  The per-tensor sums (column 0) collect the entries whose gradient is rounding noise -- Adam moves
    # them a full +-lr whatever the magnitude, and the batch-2 and batch-8 passes use different kernels for some layers (the
    # tensor-core convolutions need >= 128 positions) -- so they are held to 1e-4 of the largest possible drift 2 steps x lr x numel
    assert_digest_close(nm[~noisy], p1[~noisy], p0[~noisy], 1e-5, 1e-7, slice(1, None))
    numel = {k: p.numel() for k, p in m0.named_parameters()}
    for k, a, b in zip(nm[~noisy], p1[~noisy][:, 0], p0[~noisy][:, 0]):
        assert abs(a - b) <= 1e-5 * abs(b) + 1e-4 * 2 * 1e-5 * numel[k] + 1e-7, (k, a, b)
    assert_digest_close(nm[noisy], p1[noisy], p0[noisy], 0, 2.5e-5, slice(1, None))      # zero-gradient conv biases: Adam on noise
    # running_var = E[x^2] - mean^2 amplifies the ~1e-7 summation-order differences between the batch-2 and batch-8
    # split-reduction convolutions when mean^2 >> var (the phasegram encoder's first layers): 1e-3 here, 1e-5 vs the goldens
    for k in bn0:
        torch.testing.assert_close(bn1[k].float(), bn0[k].float(), rtol=1e-3, atol=1e-6)


def test_sub_forwards_and_toggles():
    """audio_ae_forward / visual_ae_forward (ConvTranspose decoders) vs the oracle; toggle_* freeze what they say."""
    from maavss_b200 import avse_model
    ref = om.AVFusionOracle([2, 2, 64, 128], [2, 1, 8, 4096], 8)
    sd = synth.fill_state_dict(ref.state_dict(), seed=5)
    ref.load_state_dict(sd)
    m = avse_model.AV_Fusion_Model([2, 2, 64, 128], [2, 1, 8, 4096], 8)
    m.load_state_dict(sd)
    x_a, _ = synth.stft_batch(2, 64, 128, seed=6)
    x_v = torch.randn(2, 1, 8, 4096, generator=synth.gen(7)) * 0.3
    ref.train()
    m.train()
    ya, yv = m.audio_ae_forward(x_a.cuda()), m.visual_ae_forward(x_v.cuda())
    ra, rv = ref.audio_ae_forward(x_a), ref.visual_ae_forward(x_v)
    # 10 + 10 conv / BatchNorm(batch of 2) layers deep: the split-reduction convolutions add in atomicAdd order, and that
    # run-to-run noise alone reaches 8e-6 at the decoder output (benchmarks/debug_flaky.py, 300 runs) -> atol 5e-5
    torch.testing.assert_close(ya.cpu(), ra, rtol=1e-4, atol=5e-5)
    torch.testing.assert_close(yv.cpu(), rv, rtol=1e-4, atol=5e-5)
    (ya.square().mean() + yv.square().mean()).backward()
    (ra.square().mean() + rv.square().mean()).backward()
    for (k, p), (_, q) in zip(m.named_parameters(), ref.named_parameters()):
        if q.grad is None:
            assert p.grad is None, k
        else:
            scale = max(float(q.grad.abs().max()), 1e-12)
            # 1e-6: the conv biases in front of a train-mode BatchNorm have an exactly-zero gradient; both sides hold
            # summation noise ~ eps * sum|dy| there, and ours depends on the atomicAdd order of the split reductions
            assert float((p.grad.cpu() - q.grad).abs().max()) < 2e-4 * scale + 1e-6, k
    m.toggle_enc_grads(False)
    assert not any(p.requires_grad for p in m.stft_encoder.parameters())
    assert all(p.requires_grad for p in m.fc1.parameters())
    m.toggle_fusion_grads(False)
    assert not any(p.requires_grad for p in m.lstm.parameters())


def _cos(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a @ b) / (a.norm() * b.norm()).clamp_min(1e-300))


def test_frames_model_tensor_core_golden(golden):
    """The tcgen05 (bf16 operand / fp32 accumulate) Conv3d path of AV_Fusion_Model_Frames -- the path every quoted Frames
    throughput number runs -- at MODEL level: outputs against the reference goldens and the fp32 run on this GPU at the
    SURVEY 8c tensor-core tolerance (<= 2e-2 of the reference maximum), per-parameter gradient cosine >= 0.999, and the
    training loss within 1 % of the fp32 path over 20 optimizer steps (train_avse_frames.py:150-181)."""
    from maavss_b200 import avse_model_final, nn_ops, optim, train_step
    g = golden("frames")
    xf, yf = synth.stft_batch(2, 64, 129, seed=43)
    fr = synth.attn_clip(2, 8, 256, 256, seed=47)
    ref = om.AVFusionFramesOracle([2, 2, 64, 129], [2, 1, 8, 256, 256], 8)
    sd = synth.fill_state_dict(ref.state_dict(), seed=53)
    xa, xv, ya, yv = xf.cuda(), fr.cuda(), yf[:, :, 32:40].cuda(), fr[:, :, 4].cuda()

    def run(tc):
        m = avse_model_final.AV_Fusion_Model_Frames([2, 2, 64, 129], [2, 1, 8, 256, 256], 8)
        m.load_state_dict(sd)
        nn_ops.TENSOR_CORE_CONV = tc
        try:
            m.eval()
            with torch.no_grad():
                ev = [o.clone() for o in m(xa, xv)]
            m.train()
            loss, a_loss, v_loss = micro_step_loss(m, xa, xv, ya, yv)
            loss.backward()
        finally:
            nn_ops.TENSOR_CORE_CONV = False
        return m, ev, [float(loss), float(a_loss), float(v_loss)]

    m32, ev32, l32 = run(False)
    mtc, evtc, ltc = run(True)
    rel = lambda a, b: float((a.float() - b.float()).abs().max()) / max(float(b.float().abs().max()), 1e-12)
    # eval outputs: vs the reference goldens and vs the fp32 kernels
    assert rel(evtc[0].cpu(), t(g["eval_a"])) < 2e-2
    assert rel(evtc[1].cpu(), t(g["eval_v"])) < 2e-2
    assert rel(evtc[2].cpu(), t(g["eval_fused"])) < 2e-2
    for a, b in zip(evtc, ev32):
        assert rel(a, b) < 2e-2
    np.testing.assert_allclose(ltc, g["losses"], rtol=1e-2)
    np.testing.assert_allclose(ltc, l32, rtol=1e-2)
    cosines = {}
    for (k, p), (_, q) in zip(mtc.named_parameters(), m32.named_parameters()):
        if q.grad is None:
            assert p.grad is None, k
            continue
        cosines[k] = _cos(p.grad, q.grad)
    print("frames tensor-core path: per-parameter gradient cosine vs the fp32 kernels")
    for k, c in cosines.items():
        print(f"   {k:40s} {c:.6f}")
    print("   losses tc", ltc, "fp32", l32)
    # Everything outside the visual encoder sees the bf16 convolutions only through a 3e-3 perturbation of its input:
    # cosine >= 0.999.  Inside the visual encoder the gradient passes 2-4 hard MaxPool3d layers BELOW a bf16
    # convolution: a 3e-3 forward perturbation flips the arg-max of a few per cent of the 2x2 / 3x3 windows, and each flip
    # re-routes a whole gradient element to a neighbouring pixel (the fp32 backends already disagree by ~1e-2 there under
    # 1e-6 noise, test_frames_model_golden).  That is a property of max-pooling, not of the kernels -- the per-layer tests
    # hold the tcgen05 forward / dgrad / wgrad to 2e-5 against torch on the same bf16-rounded operands -- so those
    # parameters are held to the direction only.
    for k, c in cosines.items():
        assert c >= (0.95 if k.startswith("visual_encoder") else 0.999), (k, c)
    # gradient norms against the reference goldens (fp32 path is held to 5e-3; bf16 operands: 2e-2)
    gnames, norms, _ = grad_digest(mtc)
    vis = np.array([k.startswith("visual_encoder") for k in gnames])
    np.testing.assert_allclose(norms[~vis], np.asarray(g["grad_norms"])[~vis], rtol=5e-3, atol=2e-8)
    np.testing.assert_allclose(norms[vis], np.asarray(g["grad_norms"])[vis], rtol=5e-2)       # re-routed by pool arg-max flips, see above

    # loss trajectory over 20 optimizer steps, 4 accumulation windows each
    xs, ys = (z.cuda() for z in synth.stft_batch(2, 96, 129, seed=71))
    attn = synth.attn_clip(2, 12, 256, 256, seed=73).cuda()
    traj = []
    for tc in (False, True):
        m = avse_model_final.AV_Fusion_Model_Frames([2, 2, 64, 129], [2, 1, 8, 256, 256], 8)
        m.load_state_dict(sd)
        m.train()
        opt = optim.Adam(m.parameters(), lr=1e-4)
        st = train_step.FramesTrainStep(m, opt, graph=False, tensor_core=tc)
        traj.append(np.array([[float(v) for v in st(xs, ys, attn)] for _ in range(20)]))
    np.testing.assert_allclose(traj[1], traj[0], rtol=1e-2)
    assert traj[0][-1, 1] < traj[0][0, 1]                 # the audio loss does go down over the 20 steps
    print("frames loss trajectory (a_loss) fp32 vs tcgen05:", traj[0][::5, 1], traj[1][::5, 1])


def test_config1_chain_stft_model_istft():
    """BASELINE configs[0]: one synthetic clip at run_config.py defaults -- STFT -> avse_model forward (eval and train
    mode) -> iSTFT of the predicted 64 frames -- chained on the GPU against the oracle chained on the CPU."""
    from maavss_b200 import av_dataset, avse_model, run_config, utilities
    from oracle import audio as oa
    from oracle import phasegram as oph
    import sys
    argv, sys.argv = sys.argv, ["x"]
    try:
        cfg = run_config.model_args()
    finally:
        sys.argv = argv
    hop, audio_len, n_frames = utilities.calc_hop_size(cfg.num_frames + cfg.num_seq, cfg.hops_per_frame, cfg.framerate, cfg.samplerate)
    assert (hop, audio_len, n_frames) == (66, 6336, 96)
    audio = synth.audio_clip(audio_len, seed=81)
    attn = synth.attn_clip(1, cfg.num_frames + cfg.num_seq, cfg.framesize, cfg.framesize, seed=83)
    noise = torch.randn(1, 2, 96, 128, generator=synth.gen(85))
    # the dataset exactly as train.py:31-48 builds it from the flags
    ds = av_dataset.AV_Dataset(num_frames=cfg.num_seq + cfg.num_frames, frame_hop=cfg.frame_hop, framerate=cfg.framerate,
                               samplerate=cfg.samplerate, fft_len=cfg.fft_len, hops_per_frame=cfg.hops_per_frame,
                               noise_std=cfg.noise_scalar, use_polar=cfg.use_polar, normalize_input_fft=cfg.normalize_fft,
                               normalize_output_fft=cfg.normalize_output_fft)
    ref = om.AVFusionOracle([1, 2, 64, 128], [1, 1, 8, 4096], 8)
    sd = synth.fill_state_dict(ref.state_dict(), seed=87)
    ref.load_state_dict(sd)
    m = avse_model.AV_Fusion_Model([1, 2, 64, 128], [1, 1, 8, 4096], 8, latent_channels=cfg.latent_chan, fc_size=cfg.fc_size)
    m.load_state_dict(sd)
    # oracle chain
    w = oa.hamming_window(cfg.fft_len)
    y_ref = oa.stft_planes(audio, w, cfg.fft_len, hop, True, cfg.normalize_fft, normalize_output=cfg.normalize_output_fft)[None]
    x_ref = oa.add_noise(y_ref, noise, cfg.noise_scalar)
    pg_ref = oph.video_phasegram(attn[:, :, :8], resize=(cfg.p_size, cfg.p_size))
    # device chain
    x, y = ds.gen_stft_batch(audio.cuda()[None], noise.cuda())
    pg = utilities.video_phasegram(attn.cuda()[:, :, :8], resize=(cfg.p_size, cfg.p_size))
    torch.testing.assert_close(y.cpu(), y_ref, rtol=0, atol=2e-5 * float(y_ref.abs().max()))
    torch.testing.assert_close(x.cpu(), x_ref, rtol=0, atol=2e-5 * float(x_ref.abs().max()))
    for mode in ("eval", "train"):
        getattr(ref, mode)()
        getattr(m, mode)()
        with torch.no_grad():
            # same phasegram into both models: the ill-conditioned angle of near-zero bins is pinned separately
            # (test_gpu_phasegram.py); here the chain is about STFT -> model -> iSTFT
            ra, rv, _ = ref(x_ref[:, :, :64], pg.cpu())
            a, v, _ = m(x[:, :, :64].contiguous(), pg)
        torch.testing.assert_close(a.cpu(), ra, rtol=1e-4, atol=2e-5)
        torch.testing.assert_close(v.cpu(), rv, rtol=1e-4, atol=2e-5)
        rec = ds.istft(a[0])
        rec_ref = oa.istft(ra[0], w, cfg.fft_len, hop, True, True)
        assert rec.shape == rec_ref.shape == (hop * 63,)
        torch.testing.assert_close(rec.cpu(), rec_ref, rtol=0, atol=1e-4 * max(float(rec_ref.abs().max()), 1e-3))
    assert float((pg.cpu() - pg_ref).abs().median()) < 2e-6


def test_fusion_step_graph_b16_vs_oracle():
    """BASELINE configs[1] exactly as bench.py runs it (B = 16, CUDA graph, batched windows, STFT front end on the device)
    against the oracle step bench.py times as the reference arm: first-step losses, then parameters after 3 steps."""
    import bench
    from maavss_b200 import av_dataset, avse_model, optim, train_step
    B = bench.B_PER_GPU
    torch.set_num_threads(max(1, min(16, (os.cpu_count() or 1))))
    audio, noise, attn = bench.synth_batch(torch, B, 4321)
    torch.manual_seed(0)
    ref = om.AVFusionOracle([B, 2, 64, 128], [B, 1, 8, 4096], 8)
    sd = {k: v.clone() for k, v in ref.state_dict().items()}
    ref.train()
    ropt = torch.optim.Adam(ref.parameters(), lr=bench.LR)
    from oracle import audio as oa
    from oracle import phasegram as oph
    hop = oa.calc_hop_size(bench.NUM_FRAMES, bench.HOPS_PER_FRAME, bench.FPS, bench.SR)[0]
    w = oa.hamming_window(bench.N_FFT)

    m = avse_model.AV_Fusion_Model([B, 2, 64, 128], [B, 1, 8, 4096], 8)
    m.load_state_dict(sd)
    m.train()
    opt = optim.Adam(m.parameters(), lr=bench.LR)
    ds = av_dataset.AV_Dataset(num_frames=bench.CLIP_FRAMES, fft_len=bench.N_FFT, hops_per_frame=bench.HOPS_PER_FRAME,
                               noise_std=bench.NOISE_STD, framerate=bench.FPS, samplerate=bench.SR)
    step = train_step.FusionTrainStep(m, opt, num_seq=bench.NUM_SEQ, num_frames=bench.NUM_FRAMES, hops_per_frame=bench.HOPS_PER_FRAME,
                                      p_size=64, loss_coeff=bench.LOSS_COEFF, graph=True)
    x, y = ds.gen_stft_batch(audio.cuda(), noise.cuda())
    attn_d = attn.cuda()
    # the device phasegrams of the 4 windows feed the oracle too (their near-zero-bin angle sensitivity is pinned elsewhere)
    from maavss_b200 import phasegram_ops
    pgs = phasegram_ops.video_phasegram_windows(attn_d, bench.NUM_SEQ, bench.NUM_FRAMES, resize=(64, 64)).cpu().view(bench.NUM_SEQ, B, 1, 8, 4096)
    yr = torch.stack([oa.stft_planes(audio[b], w, bench.N_FFT, hop, True, True, normalize_output=True) for b in range(B)])
    xr = oa.add_noise(yr, noise, bench.NOISE_STD)
    ours, refs = [], []
    for it in range(3):
        out = step(x, y, attn_d)
        ours.append([float(v) for v in out])
        for j in range(bench.NUM_SEQ):
            s0, s1 = bench.HOPS_PER_FRAME * j, bench.HOPS_PER_FRAME * (j + bench.NUM_FRAMES)
            loss, a_l, v_l = om.micro_step_loss(ref, xr[:, :, s0:s1], pgs[j], yr[:, :, s0:s1], pgs[j], bench.LOSS_COEFF, bench.NUM_SEQ)
            loss.backward()
        ropt.step()
        ropt.zero_grad()
        refs.append([float(loss), float(a_l), float(v_l)])
    np.testing.assert_allclose(ours[0], refs[0], rtol=1e-5)
    np.testing.assert_allclose(ours, refs, rtol=1e-4)
    assert opt.step_count == 3 and int(opt.dev_state[0]) == 3
    names = [k for k, _ in m.named_parameters()]
    names = [k for i, k in enumerate(names) if k not in names[:i]]
    mods = dict(m.named_modules())
    noisy = np.array([k.endswith(".bias") and type(mods[k.rsplit(".", 1)[0]]).__name__.startswith("Conv") for k in names])
    # Every entry of every parameter after the 3 steps (both models are in this process; stronger than the digest), compared
    # as UPDATES: delta = p - p_initial.  One Adam step moves an entry by at most lr = 1e-5 and, where |g| >> eps, by exactly
    # lr * sign-ish -- so an entry whose tiny gradient differs in the last bits between the two backends (ten train-mode
    # BatchNorm layers deep, |g| ~ eps = 1e-8) can legitimately differ by a sizeable fraction of lr.  Gate per parameter:
    # relative L2 error of the update <= 2 %, and no entry further apart than the 3 steps can move it (3 lr + slack).
    worst = ("", 0.0)
    refp = dict(ref.named_parameters())
    ours = dict(m.named_parameters())
    for k, is_noisy in zip(names, noisy):
        p, q, p0 = ours[k].detach().cpu().double(), refp[k].detach().double(), sd[k].double()
        d = float((p - q).abs().max())
        assert d <= 3.5e-5, (k, d)
        if is_noisy:
            continue
        du, dr = p - p0, q - p0
        rel = float((du - dr).norm() / dr.norm().clamp_min(1e-30))
        if rel > worst[1]:
            worst = (k, rel)
        assert rel <= 2e-2, (k, rel, d)
    print("B16 graph step vs oracle: worst per-parameter relative update error after 3 steps", worst)
    bn_ours = {k: v for k, v in m.state_dict().items() if "running" in k}
    for k, v in ref.state_dict().items():
        if "running" in k:
            torch.testing.assert_close(bn_ours[k].cpu(), v, rtol=1e-3, atol=1e-6)


def test_graph_step_checkpoint_roundtrip():
    """A CUDA-graph train step advances Adam's step counter on the device; state_dict() must carry it so
    that a checkpoint saved after N replays resumes with the right bias correction: graph x3 -> save -> load into a fresh
    model/optimizer -> one eager step == four eager steps from the start."""
    from maavss_b200 import avse_model, optim, train_step
    ref = om.AVFusionOracle([2, 2, 64, 128], [2, 1, 8, 4096], 8)
    sd = synth.fill_state_dict(ref.state_dict(), seed=31)
    xs, ys = (z.cuda() for z in synth.stft_batch(2, 96, 128, seed=37))
    attn = synth.attn_clip(2, 12, 256, 256, seed=17).cuda()

    def fresh():
        m = avse_model.AV_Fusion_Model([2, 2, 64, 128], [2, 1, 8, 4096], 8)
        m.load_state_dict(sd)
        m.train()
        return m, optim.Adam(m.parameters(), lr=1e-4)

    m0, o0 = fresh()
    eager = train_step.FusionTrainStep(m0, o0, graph=False)
    l_eager = [[float(v) for v in eager(xs, ys, attn)] for _ in range(4)]
    m1, o1 = fresh()
    gstep = train_step.FusionTrainStep(m1, o1, graph=True)
    l_graph = [[float(v) for v in gstep(xs, ys, attn)] for _ in range(3)]
    np.testing.assert_allclose(l_graph, l_eager[:3], rtol=2e-5)          # warm-up steps left no trace in the state
    ck = {"model": {k: v.clone() for k, v in m1.state_dict().items()}, "opt": o1.state_dict()}
    assert all(float(st["step"]) == 3.0 for st in ck["opt"]["state"].values())
    m2, o2 = fresh()
    m2.load_state_dict(ck["model"])
    o2.load_state_dict(ck["opt"])
    assert int(o2.dev_state[0]) == 3
    resumed = train_step.FusionTrainStep(m2, o2, graph=False)
    l4 = [float(v) for v in resumed(xs, ys, attn)]
    np.testing.assert_allclose(l4, l_eager[3], rtol=2e-5)
    # one Adam step at lr 1e-4 moves a weight by <= 1e-4; a wrong bias correction (step 1 instead of 4) changes that move by ~2x
    # (conv biases in front of a BatchNorm see Adam on rounding noise, see test_fusion_train_step_golden: skipped)
    mods = dict(m0.named_modules())
    # A wrong step count moves (nearly) EVERY entry; what the two runs may legitimately differ in is a handful of entries whose
    # gradient is rounding noise: the thin-layer weight gradients and the skinny-Linear data gradients meet through atomics, so
    # their last bits depend on CTA scheduling, and Adam turns a flipped sign into a full +-lr move per step (seen: one entry of
    # a_fc1.0.weight 1.4e-4 apart in some runs, 0 in others).  So: >= 99.9 % of every tensor within 2e-5, nothing beyond 4 steps x lr.
    for (k, p2), (_, p0) in zip(m2.named_parameters(), m0.named_parameters()):
        if k.endswith(".bias") and type(mods[k.rsplit(".", 1)[0]]).__name__.startswith("Conv"):
            continue
        d = (p2.detach() - p0.detach()).abs()
        assert (d < 2e-5).float().mean().item() >= 0.999, (k, (d < 2e-5).float().mean().item())
        assert d.max().item() <= 4.5e-4, (k, d.max().item())

"""GPU parity of the tcgen05 ViT front end: the bf16 tensor-core GEMM and flash attention against fp32 PyTorch on the
same bf16-rounded operands, and the full attention-map extractor against the goldens generated from the reference
VideoAttention / dino ViT (tests/golden/make_golden.py).  Tolerances are the bf16-operand / fp32-accumulate ones of
SURVEY 8c: <= 2e-2 of the reference maximum."""
import numpy as np
import pytest
import torch

import synth

pytestmark = pytest.mark.gpu


def rel(a, b):
    return float((a.float() - b.float()).abs().max()) / max(float(b.float().abs().max()), 1e-12)


@pytest.mark.parametrize("M,N,K", [(300, 384, 384), (2050, 1152, 384), (257, 1536, 384), (2050, 384, 1536), (1024, 384, 192),
                                   (128, 64, 64), (77, 40, 72), (5, 384, 384)])
@pytest.mark.parametrize("mode", ["f32", "bf16_bias", "gelu_bf16", "residual_f32"])
def test_tc_gemm(M, N, K, mode):
    from maavss_b200 import vit_ops
    g = synth.gen(3)
    a = (torch.randn(M, K, generator=g) * 0.5).cuda().bfloat16()
    b = (torch.randn(N, K, generator=g) * K ** -0.5).cuda().bfloat16()
    bias = torch.randn(N, generator=g).cuda()
    res = torch.randn(M, N, generator=g).cuda()
    ref = a.float() @ b.float().t()
    if mode == "f32":
        out = vit_ops.gemm_bf16(a, b, out_dtype=torch.float32)
        tol = 2e-5
    elif mode == "bf16_bias":
        out = vit_ops.gemm_bf16(a, b, bias=bias)
        ref = ref + bias
        tol = 6e-3
    elif mode == "gelu_bf16":
        out = vit_ops.gemm_bf16(a, b, bias=bias, act=vit_ops.ACT_GELU)
        ref = torch.nn.functional.gelu(ref + bias)
        tol = 6e-3
    else:
        x = res.clone()
        out = vit_ops.gemm_bf16(a, b, bias=bias, residual=x, out=x)      # in-place residual stream update
        ref = ref + bias + res
        tol = 2e-5
    assert out.shape == ref.shape
    assert rel(out, ref) < tol


@pytest.mark.parametrize("T,N,heads", [(2, 1025, 6), (3, 73, 6), (1, 128, 2), (1, 300, 1)])
def test_vit_attention(T, N, heads):
    from maavss_b200 import vit_ops
    g = synth.gen(5)
    D = heads * 64
    qkv = (torch.randn(T * N, 3 * D, generator=g)).cuda().bfloat16()
    out = vit_ops.attention(qkv, T, N, heads).float().view(T, N, heads, 64)
    q, k, v = qkv.float().view(T, N, 3, heads, 64).permute(2, 0, 3, 1, 4)
    p = ((q @ k.transpose(-2, -1)) * 0.125).softmax(-1)
    ref = (p @ v).permute(0, 2, 1, 3)
    assert rel(out, ref) < 2e-2


def cos(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a @ b) / (a.norm() * b.norm()))


@pytest.mark.parametrize("tag,depth,T,H,W", [("d12_64", 12, 3, 64, 72), ("d2_256", 2, 2, 256, 256), ("d12_256", 12, 4, 256, 256)])
def test_vit_attention_maps_golden(golden, tag, depth, T, H, W):
    """d12_256 is BASELINE configs[2]'s own geometry: ViT-S/8 depth 12, 256 x 256 frames, N = 1025 tokens (9 masked
    128-key tiles per query tile), goldens from the reference VideoAttention / dino ViT on the CPU."""
    from maavss_b200 import vit_ops
    g = golden("vit_" + tag)
    plan = vit_ops.ViTPlan(synth.vit_state_dict(depth=depth), torch.device("cuda"))
    fr = synth.frames(T, H, W, seed=13).cuda()
    rows, gh, gw = plan.cls_attention_rows(fr)
    ref_row = torch.from_numpy(g["cls_row"])
    # bf16 activations through `depth` blocks with the deliberately peaky synthetic weights (qkv std sqrt(2/dim):
    # softmax rows concentrate on a few keys, so a bf16 rounding of a logit moves a probability by up to ~4 %).
    # Gate: 2e-2 of the reference maximum at depth 2; at depth 12 the max-error gate is 5e-2 AND the
    # shape of every row must agree: cosine >= 0.999 and the same arg-max key per head.
    tol = 2e-2 if depth <= 2 else 5e-2
    r0 = rows[0].cpu()
    print(f"{tag}: cls_row max-rel {rel(r0, ref_row):.4f} cosine {cos(r0, ref_row):.6f} "
          f"argmax agree {(r0[:, 1:].argmax(-1) == ref_row[:, 1:].argmax(-1)).float().mean():.2f}")
    assert rel(r0, ref_row) < tol
    assert cos(r0, ref_row) >= 0.999
    maps = plan.attention_maps(fr).cpu()
    assert maps.shape == (T, 1, H - H % 8, W - W % 8)
    ref = torch.from_numpy(g["maps"])
    print(f"{tag}: maps max-rel {rel(maps[:, :, ::8, ::8], ref):.4f} cosine {cos(maps[:, :, ::8, ::8], ref):.6f}")
    assert rel(maps[:, :, ::8, ::8], ref) < tol
    assert cos(maps[:, :, ::8, ::8], ref) >= 0.999
    assert rel(maps[0, 0], torch.from_numpy(g["maps_full_frame0"]).float()) < tol
    # per-frame arg-max patch of the head-summed map (what the motion-attention front end localises)
    ours_peak = maps[:, 0, ::8, ::8].flatten(1).argmax(-1)
    ref_peak = ref[:, 0].flatten(1).argmax(-1)
    assert torch.equal(ours_peak, ref_peak)


def test_video_attention_class_golden(golden):
    """The drop-in class itself (video_attention.VideoAttention): `_inference` keeps the reference contract -- float32 CPU
    tensor (T, 1, H, W), per-frame max-normalised (video_attention.py:38-103) -- and `clip_attention` adds the dataset's
    frame differencing + clip normalise (av_dataset.py:323-328), both against the reference goldens at 256 x 256, depth 12."""
    from maavss_b200 import video_attention
    g = golden("vit_d12_256")
    va = video_attention.VideoAttention(patch_size=8, threshold=0.6, state_dict=synth.vit_state_dict(depth=12))
    fr = synth.frames(4, 256, 256, seed=13)
    maps = va._inference(fr)
    assert maps.device.type == "cpu" and maps.dtype == torch.float32 and maps.shape == (4, 1, 256, 256)
    assert float(maps.flatten(1).max(-1).values.min()) == pytest.approx(1.0, abs=1e-6)      # each frame divided by its max
    ref = torch.from_numpy(g["maps"])
    assert rel(maps[:, :, ::8, ::8], ref) < 5e-2 and cos(maps[:, :, ::8, ::8], ref) >= 0.999
    # nearest x8 upsample: every 8x8 block is constant
    assert torch.equal(maps[:, :, ::8, ::8].repeat_interleave(8, -2).repeat_interleave(8, -1), maps)
    plain = va.clip_attention(fr, attn_diff=False).cpu()
    diff = va.clip_attention(fr, attn_diff=True).cpu()
    assert rel(plain[:, :, ::8, ::8], torch.from_numpy(g["plainnorm"])) < 5e-2
    # the differenced clip is normalised by its (small) signed max: errors of two maps add and are scaled up by 1/max
    d_ref = torch.from_numpy(g["diffnorm"])
    assert cos(diff[:, :, ::8, ::8], d_ref) >= 0.995
    assert float(diff[0].abs().max()) == 0.0
    # crop to a multiple of the patch size (video_attention.py:43-47)
    odd = va._inference(synth.frames(1, 70, 77, seed=5))
    assert odd.shape == (1, 1, 64, 72)


def test_vit_native_plan_matches_composed_and_torch():
    """The one-call C-ABI extractor (maavss_vit_create / _prepare / _attn_maps, csrc/vit_plan.cu): identical to the
    call-by-call composition, its bicubic position-embedding resize equals torch's interpolate_pos_encoding arithmetic
    (dino/vision_transformer.py:174-194), and the fused post-processing equals attn_ops.diff_normalise."""
    from maavss_b200 import attn_ops, vit_ops
    from oracle import vit as ov
    sd = synth.vit_state_dict(depth=2)
    plan = vit_ops.ViTPlan(sd, torch.device("cuda"))
    for H, W in ((256, 256), (64, 72), (224, 224)):
        ref = ov.pos_embed_for(sd, H, W)[0]
        got = plan.native_pos_embed(H, W).cpu()
        assert got.shape == ref.shape
        assert float((got - ref).abs().max()) < 4e-6 * float(ref.abs().max()) + 1e-7, (H, W, float((got - ref).abs().max()))
    fr = synth.frames(6, 64, 72, seed=21).cuda()
    one = plan.attention_maps(fr)
    two = plan.attention_maps_composed(fr)
    assert torch.equal(one, two)
    for post, diff in ((1, False), (2, True)):
        fused = plan.attention_maps(fr, post=post, clips=2)
        sep = attn_ops.diff_normalise(two, diff, clips=2)
        assert torch.equal(fused, sep)

"""Oracle (test infrastructure): DINO ViT motion-attention front end restated on torch-CPU.

Follows /root/reference:
  dino/vision_transformer.py:116-131  PatchEmbed (8x8/8 conv == GEMM over 192-vectors)
  dino/vision_transformer.py:174-194  interpolate_pos_encoding (bicubic, the +0.1 trick)
  dino/vision_transformer.py:196-207  prepare_tokens
  dino/vision_transformer.py:80-92    Attention.forward
  dino/vision_transformer.py:59-65    Mlp            :107-113 Block.forward
  dino/vision_transformer.py:216-223  get_last_selfattention
  video_attention.py:38-103           VideoAttention._inference (dead threshold code dropped)
  av_dataset.py:323-328               attn_diff + clip max-normalise

Functional form over a plain state_dict (keys as produced by dino's VisionTransformer), batched
over frames (the reference loops at batch 1; equivalence is checked by make_golden.py).
"""
import math

import torch
import torch.nn.functional as F

LN_EPS = 1e-6  # vit_small: partial(nn.LayerNorm, eps=1e-6), dino/vision_transformer.py:243-247


def vit_config(sd):
    dim = sd["cls_token"].shape[-1]
    depth = 1 + max(int(k.split(".")[1]) for k in sd if k.startswith("blocks."))
    patch = sd["patch_embed.proj.weight"].shape[-1]
    heads = dim // 64  # vit_tiny 192/3, vit_small 384/6, vit_base 768/12: head_dim 64
    return dim, depth, heads, patch


def pos_embed_for(sd, w, h):
    """interpolate_pos_encoding; (1, 1+npatch, dim)."""
    pos = sd["pos_embed"]
    N = pos.shape[1] - 1
    dim = pos.shape[-1]
    patch = sd["patch_embed.proj.weight"].shape[-1]
    npatch = (w // patch) * (h // patch)
    if npatch == N and w == h:
        return pos
    w0, h0 = w // patch + 0.1, h // patch + 0.1
    s = int(math.sqrt(N))
    pp = F.interpolate(pos[:, 1:].reshape(1, s, s, dim).permute(0, 3, 1, 2),
                       scale_factor=(w0 / math.sqrt(N), h0 / math.sqrt(N)), mode="bicubic")
    assert int(w0) == pp.shape[-2] and int(h0) == pp.shape[-1]
    pp = pp.permute(0, 2, 3, 1).reshape(1, -1, dim)
    return torch.cat((pos[:, :1], pp), dim=1)


def prepare_tokens(sd, x):
    B, _, w, h = x.shape
    t = F.conv2d(x, sd["patch_embed.proj.weight"], sd["patch_embed.proj.bias"],
                 stride=sd["patch_embed.proj.weight"].shape[-1])
    t = t.flatten(2).transpose(1, 2)
    t = torch.cat((sd["cls_token"].expand(B, -1, -1), t), dim=1)
    return t + pos_embed_for(sd, w, h)


def block(sd, i, x, heads, return_attention=False):
    p = f"blocks.{i}."
    B, N, C = x.shape
    y = F.layer_norm(x, (C,), sd[p + "norm1.weight"], sd[p + "norm1.bias"], LN_EPS)
    qkv = F.linear(y, sd[p + "attn.qkv.weight"], sd.get(p + "attn.qkv.bias"))
    qkv = qkv.reshape(B, N, 3, heads, C // heads).permute(2, 0, 3, 1, 4)
    q, k, v = qkv[0], qkv[1], qkv[2]
    attn = (q @ k.transpose(-2, -1)) * ((C // heads) ** -0.5)
    attn = attn.softmax(dim=-1)
    if return_attention:
        return attn
    y = (attn @ v).transpose(1, 2).reshape(B, N, C)
    x = x + F.linear(y, sd[p + "attn.proj.weight"], sd[p + "attn.proj.bias"])
    y = F.layer_norm(x, (C,), sd[p + "norm2.weight"], sd[p + "norm2.bias"], LN_EPS)
    y = F.gelu(F.linear(y, sd[p + "mlp.fc1.weight"], sd[p + "mlp.fc1.bias"]))
    return x + F.linear(y, sd[p + "mlp.fc2.weight"], sd[p + "mlp.fc2.bias"])


def last_selfattention(sd, x):
    dim, depth, heads, _ = vit_config(sd)
    t = prepare_tokens(sd, x)
    for i in range(depth - 1):
        t = block(sd, i, t, heads)
    return block(sd, depth - 1, t, heads, return_attention=True)


def attention_maps(sd, frames, chunk=4):
    """VideoAttention._inference: frames (T,3,H,W) -> (T,1,H',W') with H',W' cropped to /patch."""
    _, _, _, patch = vit_config(sd)
    T, _, H, W = frames.shape
    w, h = H - H % patch, W - W % patch
    frames = frames[:, :, :w, :h]
    out = torch.zeros(T, 1, w, h)
    for s in range(0, T, chunk):
        a = last_selfattention(sd, frames[s:s + chunk])            # (b, heads, N, N)
        a = a[:, :, 0, 1:].reshape(a.shape[0], a.shape[1], w // patch, h // patch)
        a = F.interpolate(a, scale_factor=patch, mode="nearest")  # per-head nearest x patch
        m = a.sum(dim=1)                                           # sum over heads
        m = m * (1 / m.flatten(1).max(dim=1).values)[:, None, None]
        out[s:s + chunk, 0] = m
    return out


def attn_diff_normalise(attn, attn_diff):
    """av_dataset.py:323-328 then :332 permute -> (1,T,H,W)."""
    if attn_diff:
        attn = torch.cat((torch.zeros_like(attn[0:1]), torch.diff(attn, dim=0)), 0)
    attn = attn * (1 / torch.max(attn))
    return attn.permute(1, 0, 2, 3)

"""ViT-S/8 last-layer attention-map extractor on tcgen05 tensor cores (dino/vision_transformer.py + video_attention.py).

`ViTPlan` owns the packed weights (bf16 GEMM operands, fp32 biases / LayerNorm parameters, the bicubically resized
position embedding cached per input size) and the activation workspaces; `attention_maps` runs all frames of a clip as
ONE batch (the reference loops over frames at batch 1 with a host round trip per frame, video_attention.py:41-96).
The last block computes only what the CLS->patch attention row needs: K for all tokens, Q for the CLS row.
"""
import math

import torch
import torch.nn.functional as F

from ._lib import check, dptr, fptr, lib, stream_ptr

ACT_NONE, ACT_GELU = 0, 4
LN_EPS = 1e-6          # vit_small uses partial(nn.LayerNorm, eps=1e-6), dino/vision_transformer.py:243-247


def _bf16(t):
    t = t.contiguous()
    out = torch.empty(t.shape, device=t.device, dtype=torch.bfloat16)
    check(lib().maavss_f32_to_bf16(fptr(t), dptr(out), t.numel(), stream_ptr()))
    return out


def gemm_bf16(a, b, bias=None, residual=None, out=None, out_dtype=torch.bfloat16, act=ACT_NONE):
    """a (M, K) bf16 (row pitch a.stride(0)), b (N, K) bf16 -> (M, N) = a b^T (+bias) (->act) (+residual)."""
    M, K = a.shape
    N = b.shape[0]
    assert a.dtype == b.dtype == torch.bfloat16 and a.stride(1) == 1 and b.stride(1) == 1 and b.shape[1] == K
    if out is None:
        out = torch.empty(M, N, device=a.device, dtype=out_dtype)
    check(lib().maavss_tc_gemm_bf16(a.data_ptr(), b.data_ptr(), M, N, K, a.stride(0), b.stride(0),
                                    fptr(bias) if bias is not None else None,
                                    residual.data_ptr() if residual is not None else None,
                                    residual.stride(0) if residual is not None else 0,
                                    out.data_ptr(), out.stride(0), int(out.dtype == torch.bfloat16), act, stream_ptr()))
    return out


def attention(qkv, T, N, heads):
    """qkv (T*N, 3*heads*64) bf16 contiguous -> (T*N, heads*64) bf16."""
    assert qkv.is_contiguous() and qkv.dtype == torch.bfloat16
    out = torch.empty(T * N, heads * 64, device=qkv.device, dtype=torch.bfloat16)
    check(lib().maavss_vit_attention(dptr(qkv), dptr(out), T, N, heads, 0.125, stream_ptr()))
    return out


def layernorm_bf16(x, g, b):
    rows, D = x.shape
    y = torch.empty(rows, D, device=x.device, dtype=torch.bfloat16)
    check(lib().maavss_layernorm_bf16(fptr(x), fptr(g), fptr(b), dptr(y), rows, D, LN_EPS, stream_ptr()))
    return y


class ViTPlan:
    def __init__(self, state_dict, device):
        sd = {k: v.detach().to(device=device, dtype=torch.float32).contiguous() for k, v in state_dict.items()}
        self.device = device
        self.dim = sd["cls_token"].shape[-1]
        self.depth = 1 + max(int(k.split(".")[1]) for k in sd if k.startswith("blocks."))
        self.patch = sd["patch_embed.proj.weight"].shape[-1]
        if self.dim % 64:
            raise ValueError("head_dim must be 64 (vit_tiny / vit_small / vit_base)")
        self.heads = self.dim // 64
        self.cls = sd["cls_token"].reshape(-1).contiguous()
        self.pos = sd["pos_embed"]
        self.w_pe = _bf16(sd["patch_embed.proj.weight"].reshape(self.dim, -1))
        self.b_pe = sd["patch_embed.proj.bias"]
        self.blocks = []
        for i in range(self.depth):
            p = f"blocks.{i}."
            qkv_b = sd.get(p + "attn.qkv.bias")
            if qkv_b is None:
                qkv_b = torch.zeros(3 * self.dim, device=device)
            self.blocks.append(dict(
                ln1=(sd[p + "norm1.weight"], sd[p + "norm1.bias"]), w_qkv=_bf16(sd[p + "attn.qkv.weight"]), b_qkv=qkv_b,
                w_proj=_bf16(sd[p + "attn.proj.weight"]), b_proj=sd[p + "attn.proj.bias"],
                ln2=(sd[p + "norm2.weight"], sd[p + "norm2.bias"]), w_fc1=_bf16(sd[p + "mlp.fc1.weight"]), b_fc1=sd[p + "mlp.fc1.bias"],
                w_fc2=_bf16(sd[p + "mlp.fc2.weight"]), b_fc2=sd[p + "mlp.fc2.bias"]))
        self._pos_cache = {}
        self._native = self._create_native(sd)
        self._prepared = set()

    # ---- the one-call C-ABI plan (maavss_vit_create / maavss_vit_attn_maps, csrc/vit_plan.cu) ----
    _BLOCK_KEYS = ("norm1.weight", "norm1.bias", "attn.qkv.weight", "attn.qkv.bias", "attn.proj.weight", "attn.proj.bias",
                   "norm2.weight", "norm2.bias", "mlp.fc1.weight", "mlp.fc1.bias", "mlp.fc2.weight", "mlp.fc2.bias")

    def _create_native(self, sd):
        import ctypes
        n0 = sd["pos_embed"].shape[1] - 1
        grid = int(round(math.sqrt(n0)))
        if grid * grid != n0:
            raise ValueError("pos_embed is not a square grid")
        zeros3 = torch.zeros(3 * self.dim, device=self.device)
        ptrs = []
        keep = [zeros3]
        for i in range(self.depth):
            for k in self._BLOCK_KEYS:
                t = sd.get(f"blocks.{i}.{k}")
                if t is None and k == "attn.qkv.bias":
                    t = zeros3                                   # qkv_bias=False models
                keep.append(t)
                ptrs.append(t.data_ptr())
        arr = (ctypes.c_void_p * len(ptrs))(*ptrs)
        plan = ctypes.c_void_p()
        pw = sd["patch_embed.proj.weight"].reshape(self.dim, -1).contiguous()
        torch.cuda.synchronize(self.device)
        check(lib().maavss_vit_create(ctypes.byref(plan), self.dim, self.depth, self.patch, sd["patch_embed.proj.weight"].shape[1], grid,
                                      fptr(sd["cls_token"].reshape(-1).contiguous()), fptr(sd["pos_embed"].reshape(-1, self.dim).contiguous()),
                                      fptr(pw), fptr(sd["patch_embed.proj.bias"]), arr))
        del keep
        return plan

    def __del__(self):
        try:
            if getattr(self, "_native", None):
                lib().maavss_vit_destroy(self._native)
                self._native = None
        except Exception:
            pass

    def native_pos_embed(self, H, W):
        """The plan's interpolated position embedding for an (H, W) input (tests: equals interpolate_pos_encoding)."""
        self._prepare(H, W)
        gh, gw = H // self.patch, W // self.patch
        out = torch.empty(1 + gh * gw, self.dim, device=self.device)
        check(lib().maavss_vit_pos_embed(self._native, H, W, fptr(out)))
        return out

    def _prepare(self, H, W):
        if (H, W) not in self._prepared:
            check(lib().maavss_vit_prepare(self._native, H, W))
            self._prepared.add((H, W))

    def pos_embed_for(self, w, h):
        """interpolate_pos_encoding (dino/vision_transformer.py:174-194) for a (w, h) = (rows, columns) input, evaluated once
        per size by the plan's bicubic kernel."""
        key = (w, h)
        if key not in self._pos_cache:
            self._pos_cache[key] = self.native_pos_embed(w, h)
        return self._pos_cache[key]

    @torch.no_grad()
    def cls_attention_rows(self, frames):
        """frames (T, 3, H, W) fp32 on the device -> (T, heads, N) softmax row of the CLS query in the last block."""
        L = lib()
        frames = frames.contiguous()
        T, C, H, W = frames.shape
        P, D = self.patch, self.dim
        gh, gw = H // P, W // P
        np_, N = gh * gw, gh * gw + 1
        dev = frames.device
        patches = torch.empty(T * np_, C * P * P, device=dev, dtype=torch.bfloat16)
        check(L.maavss_vit_patchify(fptr(frames), dptr(patches), T, C, H, W, gh, gw, P, stream_ptr()))
        tok = gemm_bf16(patches, self.w_pe, bias=self.b_pe, out_dtype=torch.float32)
        x = torch.empty(T * N, D, device=dev, dtype=torch.float32)
        check(L.maavss_vit_assemble_tokens(fptr(tok), fptr(self.cls), fptr(self.pos_embed_for(gh * P, gw * P)), fptr(x), T, np_, D,
                                           stream_ptr()))
        for blk in self.blocks[:-1]:
            y = layernorm_bf16(x, *blk["ln1"])
            qkv = gemm_bf16(y, blk["w_qkv"], bias=blk["b_qkv"])
            o = attention(qkv, T, N, self.heads)
            gemm_bf16(o, blk["w_proj"], bias=blk["b_proj"], residual=x, out=x)
            y = layernorm_bf16(x, *blk["ln2"])
            hdn = gemm_bf16(y, blk["w_fc1"], bias=blk["b_fc1"], act=ACT_GELU)
            gemm_bf16(hdn, blk["w_fc2"], bias=blk["b_fc2"], residual=x, out=x)
        blk = self.blocks[-1]
        y = layernorm_bf16(x, *blk["ln1"])
        k = gemm_bf16(y, blk["w_qkv"][D:2 * D], bias=blk["b_qkv"][D:2 * D].contiguous(), out_dtype=torch.float32)   # K of every token
        y_cls = y.view(T, N, D)[:, 0, :]                                                           # CLS rows, pitch N*D
        q = gemm_bf16(y_cls, blk["w_qkv"][:D], bias=blk["b_qkv"][:D].contiguous(), out_dtype=torch.float32)
        rows = torch.empty(T, self.heads, N, device=dev, dtype=torch.float32)
        check(L.maavss_vit_cls_attn_row(fptr(q), q.stride(0), fptr(k), k.stride(0), fptr(rows), T, self.heads, N, 0.125, stream_ptr()))
        return rows, gh, gw

    @torch.no_grad()
    def attention_maps(self, frames, post=0, clips=1):
        """VideoAttention._inference on the device in ONE C-ABI call: (T, 3, H, W) -> (T, 1, gh*P, gw*P), per-frame
        max-normalised (post 0); post 1 / 2 adds the dataset's clip normalise / frame difference + clip normalise
        (av_dataset.py:323-328) over `clips` equal groups of frames."""
        import ctypes
        frames = frames.contiguous()
        T, C, H, W = frames.shape
        self._prepare(H, W)
        gh, gw = H // self.patch, W // self.patch
        out = torch.empty(T, 1, gh * self.patch, gw * self.patch, device=frames.device, dtype=torch.float32)
        nbytes = lib().maavss_vit_workspace_bytes(self._native, T, H, W, clips)
        ws = torch.empty(nbytes, device=frames.device, dtype=torch.uint8)
        check(lib().maavss_vit_attn_maps(self._native, fptr(frames), T, H, W, fptr(out), int(post), int(clips), dptr(ws),
                                         ctypes.c_longlong(nbytes), stream_ptr()))
        return out

    @torch.no_grad()
    def attention_maps_composed(self, frames):
        """The same result composed call by call from Python (kept for tests of the individual entry points)."""
        rows, gh, gw = self.cls_attention_rows(frames)
        T, heads, N = rows.shape
        out = torch.empty(T, 1, gh * self.patch, gw * self.patch, device=frames.device, dtype=torch.float32)
        check(lib().maavss_attn_headsum_upsample(fptr(rows), fptr(out), T, heads, N, gh, gw, self.patch, stream_ptr()))
        return out

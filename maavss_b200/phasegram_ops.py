"""utilities.video_phasegram (utilities.py:206-228) on the device through the C ABI."""
import ctypes

import torch

from ._lib import check, dptr, fptr, lib, stream_ptr


def video_phasegram(frames, resize=None, diff=True, cumulative=True, normalize=True):
    """frames (B, 1, T, H, W) cuda fp32 -> (B, 1, T, R*R).  No gradient flows through it (the reference feeds detached
    attention maps, train.py:138-139)."""
    f = torch.squeeze(frames.detach(), 1)
    if f.dim() != 4:
        raise RuntimeError(f"video_phasegram expects (B, 1, T, H, W), got {tuple(frames.shape)}")
    B, T, H, W = f.shape
    if not f.is_cuda or f.dtype != torch.float32:
        raise RuntimeError("video_phasegram needs a CUDA fp32 tensor (no CPU fallback)")
    if f.stride()[1:] != (H * W, W, 1) or (B > 1 and f.stride(0) < T * H * W):
        f = f.contiguous()            # a time-slice view attn[:, :, j:j+T] is consumed in place (train.py:138)
    bstride = f.stride(0) if B > 1 else T * H * W
    if resize is None:
        if H != W:
            raise RuntimeError("video_phasegram without resize needs square frames")
        R = H
    else:
        if tuple(resize)[0] != tuple(resize)[1]:
            raise RuntimeError("video_phasegram: only square resize targets are supported")
        R = int(tuple(resize)[0])
    out = torch.empty(B, T, R * R, device=f.device, dtype=torch.float32)
    scratch = torch.empty_like(out)
    mx = torch.empty(1, device=f.device, dtype=torch.int32)
    check(lib().maavss_video_phasegram(ctypes.c_void_p(f.data_ptr()), fptr(out), fptr(scratch), dptr(mx), B, T, H, W, R, bstride,
                                       int(bool(diff)),
                                       int(bool(cumulative)), int(bool(normalize)), stream_ptr()))
    return out.unsqueeze(1)

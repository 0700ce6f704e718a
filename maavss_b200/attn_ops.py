"""Python surface of the attention-map post-processing kernels."""
import torch

from ._lib import check, dptr, fptr, lib, stream_ptr


def diff_normalise(attn, attn_diff, clips=1):
    """attn (T,1,H,W) [one clip] or (clips*T, ...) -> same shape; av_dataset.py:323-328 per clip."""
    a = attn.contiguous()
    T = a.shape[0] // clips
    hw = a[0].numel()
    out = torch.empty_like(a)
    ws = torch.empty(max(clips, 1), device=a.device, dtype=torch.int32)
    check(lib().maavss_attn_diff_normalise(fptr(a), fptr(out), clips, T, hw, int(bool(attn_diff)), dptr(ws), stream_ptr()))
    return out


def headsum_upsample(cls_rows, grid_h, grid_w, patch):
    """cls_rows (frames, heads, stride) -> (frames, 1, grid_h*patch, grid_w*patch)."""
    fr, heads, stride = cls_rows.shape
    out = torch.empty(fr, 1, grid_h * patch, grid_w * patch, device=cls_rows.device, dtype=torch.float32)
    check(lib().maavss_attn_headsum_upsample(fptr(cls_rows), fptr(out), fr, heads, stride, grid_h, grid_w, patch, stream_ptr()))
    return out

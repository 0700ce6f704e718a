"""utilities.video_phasegram (utilities.py:206-228) on the device through the C ABI."""
import torch

from ._lib import check, dptr, fptr, lib, stream_ptr


def _frames_view(frames):
    """(B, 1, T, H, W) -> (tensor whose clips are `bstride` elements apart and whose frames are contiguous, B, T, H, W, bstride).
    A time-slice view attn[:, :, j:j+T] is consumed in place (train.py:138)."""
    f = torch.squeeze(frames.detach(), 1)
    if f.dim() != 4:
        raise RuntimeError(f"video_phasegram expects (B, 1, T, H, W), got {tuple(frames.shape)}")
    B, T, H, W = f.shape
    if not f.is_cuda or f.dtype != torch.float32:
        raise RuntimeError("video_phasegram needs a CUDA fp32 tensor (no CPU fallback)")
    if f.stride()[1:] != (H * W, W, 1) or (B > 1 and f.stride(0) < T * H * W):
        f = f.contiguous()
    return f, B, T, H, W, (f.stride(0) if B > 1 else T * H * W)


def _target(resize, H, W):
    if resize is None:
        if H != W:
            raise RuntimeError("video_phasegram without resize needs square frames")
        return H
    if tuple(resize)[0] != tuple(resize)[1]:
        raise RuntimeError("video_phasegram: only square resize targets are supported")
    return int(tuple(resize)[0])


def video_phasegram_windows(frames, num_windows, window_frames, resize=(64, 64), diff=True, cumulative=True, normalize=True,
                            batch_roll=None, max_group=None):
    """The num_seq calls `video_phasegram(attn[:, :, j:j+Tw])`, j = 0..G-1, of one accumulation loop (train.py:136-139)
    as one pass: frames (B, 1, Tc, H, W) -> (G*B, 1, Tw, R*R), window j in rows [j*B, (j+1)*B).  Each of the Tc distinct
    frames is resized / transformed once (the windows overlap); results equal the per-window calls bit for bit.

    batch_roll: positions the all-dims fftshift rolls the batch axis (default B // 2, utilities.py:210).
    max_group: a torch.distributed group -> the per-window max |.| is all-reduced over it (the reference normalises by the
    max of the WHOLE batch, utilities.py:226-227); the caller then feeds each rank the attention frames of the clips the
    GLOBAL roll pairs with its audio (parallel.rolled_clip_indices) and passes batch_roll=0."""
    f, B, Tc, H, W, bstride = _frames_view(frames)
    G, Tw = int(num_windows), int(window_frames)
    if G < 1 or G - 1 + Tw > Tc:
        raise RuntimeError(f"{G} windows of {Tw} frames do not fit a {Tc}-frame clip")
    R = _target(resize, H, W)
    pflat = torch.empty(B, Tc, R * R, device=f.device, dtype=torch.float32)
    mx = torch.empty(G, device=f.device, dtype=torch.int32)
    out = torch.empty(G * B, Tw, R * R, device=f.device, dtype=torch.float32)
    # the resize scratch (B*Tc*R*R floats) is dead before the window kernel writes `out`: share the allocation when it fits
    scratch = out if G * Tw >= Tc else torch.empty_like(pflat)
    roll = B // 2 if batch_roll is None else int(batch_roll)
    L = lib()
    check(L.maavss_phasegram_frames(fptr(f) if f.is_contiguous() else f.data_ptr(), fptr(pflat), fptr(scratch), B, Tc, H, W, R,
                                    bstride, int(bool(cumulative)), stream_ptr()))
    check(L.maavss_phasegram_windows_max(fptr(pflat), dptr(mx), B, Tc, Tw, G, R, roll, int(bool(diff)), stream_ptr()))
    if max_group is not None and normalize:
        from . import parallel
        parallel.allreduce_max_(mx)                                      # non-negative float bits order like integers
    check(L.maavss_phasegram_windows_write(fptr(pflat), dptr(mx), fptr(out), B, Tc, Tw, G, R, roll, int(bool(diff)),
                                           int(bool(normalize)), stream_ptr()))
    return out.unsqueeze(1)


def video_phasegram(frames, resize=None, diff=True, cumulative=True, normalize=True, batch_roll=None, max_group=None):
    """frames (B, 1, T, H, W) cuda fp32 -> (B, 1, T, R*R).  No gradient flows through it (the reference feeds detached
    attention maps, train.py:138-139).  batch_roll / max_group: see video_phasegram_windows (data-parallel callers)."""
    T = frames.shape[2] if frames.dim() == 5 else frames.shape[1]
    return video_phasegram_windows(frames, 1, T, resize=resize, diff=diff, cumulative=cumulative, normalize=normalize,
                                   batch_roll=batch_roll, max_group=max_group)

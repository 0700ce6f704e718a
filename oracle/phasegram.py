"""Oracle (test infrastructure): utilities.video_phasegram (utilities.py:206-228) on torch-CPU.

Quirks reproduced, not fixed: fftshift over ALL dims (rolls batch and time axes);
resize = torchvision TF.resize semantics of the installed version (bilinear, antialias=True)
restated with F.interpolate(mode='bilinear', antialias=True, align_corners=False).
"""
import math

import torch
import torch.nn.functional as F


def video_phasegram(frames, resize=None, diff=True, cumulative=True, normalize=True):
    frames = torch.squeeze(frames, 1)
    if resize is not None:
        frames = F.interpolate(frames, size=tuple(resize), mode="bilinear", antialias=True,
                               align_corners=False)
    fft = torch.fft.fftshift(torch.fft.fft2(frames))
    p = torch.angle(fft)
    p_flat = torch.flatten(p, start_dim=-2, end_dim=-1)
    if cumulative:
        p_flat = torch.cumsum(p_flat, dim=-1)
        p_flat = p_flat / (2.0 * math.pi * p_flat.shape[-1])
    else:
        p_flat = (p_flat + math.pi) / (math.pi * 2.0)
    if diff:
        p_diff = torch.diff(p_flat, dim=-2)
        phasegram = torch.cat((torch.zeros_like(p_diff[:, 0:1, :]), p_diff), dim=1)
    else:
        phasegram = p_flat
    phasegram = torch.unsqueeze(phasegram, 1)
    if normalize:
        phasegram = phasegram * (1 / torch.max(torch.abs(phasegram)))
    return phasegram

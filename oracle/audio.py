"""Oracle (test infrastructure): audio front end of MAAVSS restated on torch-CPU.

Follows /root/reference:
  utilities.py:24-28     calc_hop_size
  av_dataset.py:106      window = torch.hamming_window(fft_len)   (periodic)
  av_dataset.py:157-179  AV_Dataset.stft   (== STFT_Dataset.stft :450-467)
  av_dataset.py:181-201  AV_Dataset.istft  (== STFT_Dataset.istft :431-448)
  av_dataset.py:217-220  add_noise
  av_dataset.py:335-342  gen_stft_example (permute + optional 1/max|.| normalise)
  generator.py:60-133    DataGenerator.fft / ifft / cartesian_to_polar / complex_to_ri /
                         ri_to_complex / center_fft_bins / decenter_fft_bins

The two AV_Dataset methods no longer run as written on torchaudio >= 2 (they index
a complex tensor as if it were the old real (F,T,2) layout and call the removed
``magphase``); they are restated on the complex API exactly as SURVEY.md section 8c
prescribes.  ``tests/golden/make_golden.py`` confirms the restatement against
``torchaudio.functional.spectrogram`` called with the reference's own arguments.
"""
import math

import torch
import torch.nn.functional as F


def calc_hop_size(num_frames, hops_per_frame, fps, sr):
    # utilities.py:24-28 -- float division then int() truncation.
    hop = int((sr / fps) / hops_per_frame)
    audio_sample_len = int(hops_per_frame * hop * num_frames)
    num_fft_frames = audio_sample_len // hop
    return hop, audio_sample_len, num_fft_frames


def hamming_window(n_fft):
    return torch.hamming_window(n_fft)  # periodic=True default, av_dataset.py:106


def stft_complex(audio, window, n_fft, hop, normalized=True):
    """torchaudio.functional.spectrogram(pad=0, power=None, normalized, onesided) with its
    defaults center=True / pad_mode='reflect': torch.stft then division by sqrt(sum w^2)."""
    spec = torch.stft(audio, n_fft=n_fft, hop_length=hop, win_length=window.shape[0],
                      window=window, center=True, pad_mode="reflect", normalized=False,
                      onesided=True, return_complex=True)
    if normalized:
        spec = spec / window.pow(2.0).sum().sqrt()
    return spec


def stft(audio, window, n_fft, hop, trim_stft_end=True, normalized=True, use_polar=False):
    """AV_Dataset.stft (av_dataset.py:157-179) for a 1-D clip.  Returns the old real layout
    (F', T-1, 2) (or (2, F', T-1) = [mag; phase] when use_polar)."""
    spec = torch.view_as_real(stft_complex(audio, window, n_fft, hop, normalized))
    if trim_stft_end:
        spec = spec[:-1, :-1, :]
    else:
        spec = spec[:, :-1, :]
    if use_polar:
        c = torch.view_as_complex(spec.contiguous())
        spec = torch.cat((c.abs().unsqueeze(0), c.angle().unsqueeze(0)), dim=0)
    return spec


def stft_planes(audio, window, n_fft, hop, trim_stft_end=True, normalized=True,
                normalize_output=False):
    """gen_stft_example's y_stft (av_dataset.py:337-340): (2, T-1, F') real/imag planes."""
    y = stft(audio, window, n_fft, hop, trim_stft_end, normalized).permute(2, 1, 0).contiguous()
    if normalize_output:
        y = y * (1 / torch.max(torch.abs(y) + 1e-7))
    return y


def add_noise(tensor, noise, noise_std):
    # av_dataset.py:217-220; the gaussian draw itself is passed in (RNG is never restated).
    return tensor + noise * noise_std


def istft(planes, window, n_fft, hop, trim_stft_end=True, normalized=True, use_polar=False):
    """AV_Dataset.istft (av_dataset.py:181-201): planes (2, T, F') -> audio (hop*(T-1),)."""
    if trim_stft_end:
        x = F.pad(planes, (0, 1)).permute(2, 1, 0)
    else:
        x = planes.permute(2, 1, 0)
    if use_polar:
        # av_dataset.py:187-191 is broken as written (calls a tensor); formula semantics:
        mag, phase = x[:, :, 0], x[:, :, 1]
        c = mag * (torch.cos(phase) + 1j * torch.sin(phase))
    else:
        c = torch.view_as_complex(x.contiguous())
    return torch.istft(c, n_fft=n_fft, hop_length=hop, win_length=n_fft, window=window,
                       normalized=normalized, onesided=True)


def magphase_split(planes):
    """(.., 2, T, F) real/imag -> (.., 2, T, F) [|z|; atan2(im, re)]  (av_dataset.py:176-178)."""
    re, im = planes[..., 0, :, :], planes[..., 1, :, :]
    return torch.stack((torch.sqrt(re * re + im * im), torch.atan2(im, re)), dim=-3)


def magphase_combine(planes):
    mag, ph = planes[..., 0, :, :], planes[..., 1, :, :]
    return torch.stack((mag * torch.cos(ph), mag * torch.sin(ph)), dim=-3)


# ---------------- DataGenerator tensor helpers (generator.py:60-133) ----------------

def gen_fft(audio, normalize=True):
    f = torch.fft.fft(audio, dim=-1, norm="forward" if normalize else None)
    return f[:, : f.shape[-1] // 2]


def gen_ifft(fft, normalize=True):
    # generator.py:71-81: zero-fills the mirrored half (2*K bins, not the original N) and
    # uses norm="forward" on the inverse too -- reproduced, not fixed.
    fft = torch.cat((fft, torch.zeros_like(fft)), -1)
    audio = torch.fft.ifft(fft, dim=-1, norm="forward" if normalize else None)
    return torch.real(audio)


def cartesian_to_polar(cart):
    return torch.cat((torch.abs(cart), torch.angle(cart)), -1)


def complex_to_ri(t):
    return torch.view_as_real(t).permute(0, 2, 1)


def ri_to_complex(t):
    return torch.view_as_complex(t.permute(0, 2, 1).contiguous())


def center_fft_bins(t):
    left = torch.flip(t[:, :, ::2], [-1])
    return torch.cat((left, t[:, :, 1::2]), -1)


def decenter_fft_bins(t):
    out = torch.zeros_like(t)
    half = t.shape[-1] // 2
    out[:, :, ::2] = torch.flip(t[:, :, :half], [-1])
    out[:, :, 1::2] = t[:, :, half:]
    return out


# ---------------- index-exact framing (independent of any FFT library) ----------------

def reflect_index(i, n):
    """Index into the un-padded clip for padded position i - pad (torch 'reflect' mode)."""
    if i < 0:
        return -i
    if i >= n:
        return 2 * (n - 1) - i
    return i


def frame_sample_indices(t, n_samples, n_fft, hop):
    """Source sample index of every tap of frame t under center=True reflect padding."""
    pad = n_fft // 2
    return [reflect_index(t * hop + j - pad, n_samples) for j in range(n_fft)]


def stft_naive(audio, window, n_fft, hop, normalized=True):
    """O(N^2) float64 DFT restatement: pins frame placement/normalisation independently of
    torch.stft.  Small inputs only."""
    x = audio.double()
    w = window.double()
    n = x.shape[-1]
    T = 1 + n // hop
    k = torch.arange(n_fft // 2 + 1, dtype=torch.float64)[:, None]
    j = torch.arange(n_fft, dtype=torch.float64)[None, :]
    ang = -2.0 * math.pi * k * j / n_fft
    cosm, sinm = torch.cos(ang), torch.sin(ang)
    out = torch.zeros(n_fft // 2 + 1, T, 2, dtype=torch.float64)
    for t in range(T):
        idx = torch.tensor(frame_sample_indices(t, n, n_fft, hop))
        fr = x[idx] * w
        out[:, t, 0] = cosm @ fr
        out[:, t, 1] = sinm @ fr
    if normalized:
        out = out / w.pow(2).sum().sqrt()
    return out

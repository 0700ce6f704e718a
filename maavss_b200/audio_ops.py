"""Python surface of the audio front-end kernels (thin wrappers over the C ABI)."""
import ctypes

import torch

from ._lib import check, fptr, lib, stream_ptr, dptr

_stft_plans = {}
_dft_plans = {}


class _Plan:
    def __init__(self, handle, destroy):
        self.handle, self._destroy = handle, destroy

    def __del__(self):
        try:
            self._destroy(self.handle)
        except Exception:
            pass


def stft_plan(n_fft, hop, window):
    w = window.detach().to("cpu", torch.float32).contiguous()
    key = (n_fft, hop, torch.cuda.current_device(), w.numpy().tobytes())
    p = _stft_plans.get(key)
    if p is None:
        h = ctypes.c_void_p()
        check(lib().maavss_stft_plan_create(ctypes.byref(h), n_fft, hop, ctypes.c_void_p(w.data_ptr())))
        p = _stft_plans[key] = _Plan(h, lib().maavss_stft_plan_destroy)
    return p


def stft_out_dims(n_samples, n_fft, hop, trim_stft_end):
    T, F = ctypes.c_int(), ctypes.c_int()
    check(lib().maavss_stft_out_dims(n_samples, n_fft, hop, int(trim_stft_end), ctypes.byref(T), ctypes.byref(F)))
    return T.value, F.value


def stft_planes(audio, window, n_fft, hop, trim_stft_end=True, normalized=True, polar=False, out=None):
    """audio (B, N) cuda fp32 -> (B, 2, T_out, F_out)."""
    assert audio.dim() == 2
    B, N = audio.shape
    T, F = stft_out_dims(N, n_fft, hop, trim_stft_end)
    if out is None:
        out = torch.empty(B, 2, T, F, device=audio.device, dtype=torch.float32)
    plan = stft_plan(n_fft, hop, window)
    check(lib().maavss_stft_fwd(plan.handle, fptr(audio), fptr(out), B, N, int(trim_stft_end), int(normalized),
                                int(polar), stream_ptr()))
    return out


def istft_audio(planes, window, n_fft, hop, trim_stft_end=True, normalized=True, polar=False, out=None):
    """planes (B, 2, T, F_in) cuda fp32 -> (B, hop*(T-1))."""
    assert planes.dim() == 4 and planes.shape[1] == 2
    B, _, T, F = planes.shape
    exp_f = n_fft // 2 if trim_stft_end else n_fft // 2 + 1
    if F != exp_f:
        raise RuntimeError(f"istft: expected {exp_f} frequency bins, got {F}")
    if out is None:
        out = torch.empty(B, hop * max(T - 1, 0), device=planes.device, dtype=torch.float32)
    plan = stft_plan(n_fft, hop, window)
    check(lib().maavss_istft_fwd(plan.handle, fptr(planes), fptr(out), B, T, int(trim_stft_end), int(normalized),
                                 int(polar), stream_ptr()))
    return out


def absmax_normalise_(x, eps=1e-7):
    """In place: x[b] *= 1/max(|x[b]| + eps) for every leading index b."""
    B = x.shape[0]
    ws = torch.empty(max(B, 1), device=x.device, dtype=torch.int32)
    check(lib().maavss_absmax_normalise(fptr(x), B, x[0].numel() if B else 0, eps, dptr(ws), stream_ptr()))
    return x


def add_noise(y, noise, noise_std):
    x = torch.empty_like(y)
    check(lib().maavss_add_noise(fptr(y), fptr(noise), float(noise_std), fptr(x), y.numel(), stream_ptr()))
    return x


def magphase_split(planes):
    out = torch.empty_like(planes)
    plane = planes.shape[-1] * planes.shape[-2]
    check(lib().maavss_magphase_split(fptr(planes), fptr(out), planes.numel() // (2 * plane) if plane else 0, plane, stream_ptr()))
    return out


def magphase_combine(planes):
    out = torch.empty_like(planes)
    plane = planes.shape[-1] * planes.shape[-2]
    check(lib().maavss_magphase_combine(fptr(planes), fptr(out), planes.numel() // (2 * plane) if plane else 0, plane, stream_ptr()))
    return out


def bins_center(x):
    out = torch.empty_like(x)
    K = x.shape[-1]
    check(lib().maavss_bins_center(fptr(x), fptr(out), x.numel() // K if K else 0, K, stream_ptr()))
    return out


def bins_decenter(x):
    out = torch.empty_like(x)
    K = x.shape[-1]
    check(lib().maavss_bins_decenter(fptr(x), fptr(out), x.numel() // K if K else 0, K, stream_ptr()))
    return out


def dft_plan(n):
    key = (n, torch.cuda.current_device())
    p = _dft_plans.get(key)
    if p is None:
        h = ctypes.c_void_p()
        check(lib().maavss_dft_plan_create(ctypes.byref(h), n))
        p = _dft_plans[key] = _Plan(h, lib().maavss_dft_plan_destroy)
    return p


def dft_fwd(audio, n_keep, norm_forward=True):
    """audio (B, N) real -> complex64 (B, n_keep)."""
    B, N = audio.shape
    out = torch.empty(B, n_keep, 2, device=audio.device, dtype=torch.float32)
    check(lib().maavss_dft_fwd(dft_plan(N).handle, fptr(audio), fptr(out), B, n_keep, int(norm_forward), stream_ptr()))
    return torch.view_as_complex(out)


def dft_half_inv(spec, norm_forward=True):
    """complex64 (B, K) -> real (B, 2K): generator.ifft semantics."""
    B, K = spec.shape
    x = torch.view_as_real(spec.contiguous())
    out = torch.empty(B, 2 * K, device=spec.device, dtype=torch.float32)
    check(lib().maavss_dft_half_inv(dft_plan(2 * K).handle, fptr(x), fptr(out), B, int(norm_forward), stream_ptr()))
    return out

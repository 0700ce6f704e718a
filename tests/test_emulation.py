"""CPU emulation of the STFT/iSTFT kernel phases (the same .cuh code compiled with g++ and run
thread-by-thread) against the oracle.  This validates FFT passes, twiddles, reflect framing,
Hermitian pack/unpack and the gather overlap-add without a GPU."""
import ctypes
import os
import subprocess

import numpy as np
import pytest
import torch

from oracle import audio as oa

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EMU = os.path.join(ROOT, "tests", "emu")


@pytest.fixture(scope="module")
def emu():
    so = os.path.join(EMU, "_build", "libemu_stft.so")
    os.makedirs(os.path.dirname(so), exist_ok=True)
    src = os.path.join(EMU, "emu_stft.cpp")
    subprocess.check_call(["g++", "-O1", "-std=c++17", "-shared", "-fPIC", "-I", EMU, "-include", "cuda_emu.h",
                           src, "-o", so])
    return ctypes.CDLL(so)


def fptr(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_float))


@pytest.mark.parametrize("n_fft,hop,n", [(64, 16, 500), (128, 32, 700), (256, 66, 6336), (512, 128, 3000),
                                         (1024, 256, 5000), (2048, 512, 9000), (4096, 1024, 12000), (256, 100, 3001),
                                         (256, 77, 3000), (128, 128, 1000), (512, 300, 4000), (4096, 512, 30000),
                                         (4096, 2048, 30000), (2048, 1024, 9000), (1024, 64, 5000), (512, 8, 2000), (1024, 1024, 9000)])
@pytest.mark.parametrize("trim", [1, 0])
def test_stft_istft_emulated(emu, n_fft, hop, n, trim):
    g = torch.Generator().manual_seed(n_fft + hop)
    audio = torch.randn(n, generator=g) * 0.1
    w = oa.hamming_window(n_fft)
    ref = oa.stft_planes(audio, w, n_fft, hop, bool(trim))
    out = np.zeros(tuple(ref.shape), np.float32)
    a_np, w_np = audio.numpy().copy(), w.numpy().copy()
    assert emu.emu_stft(fptr(a_np), n, n_fft, hop, fptr(w_np), trim, 1, 0, fptr(out)) == 0
    tol = 2e-6 * float(ref.abs().max()) * np.log2(n_fft)
    assert np.abs(out - ref.numpy()).max() < tol
    # inverse
    ref_i = oa.istft(ref, w, n_fft, hop, bool(trim))
    T = ref.shape[1]
    for G in (3, 64):
        aud = np.zeros(hop * (T - 1), np.float32)
        planes = ref.numpy().copy()
        assert emu.emu_istft(fptr(planes), T, n_fft, hop, fptr(w_np), trim, 1, 0, G, fptr(aud)) == 0
        assert np.abs(aud - ref_i.numpy()).max() < 3e-6 * max(1.0, float(ref_i.abs().max())) * np.log2(n_fft)


def test_polar_emulated(emu):
    n_fft, hop, n = 256, 66, 2000
    audio = torch.randn(n, generator=torch.Generator().manual_seed(5)) * 0.1
    w = oa.hamming_window(n_fft)
    ref = oa.stft_planes(audio, w, n_fft, hop, True)
    pol = oa.magphase_split(ref)
    out = np.zeros(tuple(ref.shape), np.float32)
    a_np, w_np = audio.numpy().copy(), w.numpy().copy()
    emu.emu_stft(fptr(a_np), n, n_fft, hop, fptr(w_np), 1, 1, 1, fptr(out))
    assert np.abs(out[0] - pol[0].numpy()).max() < 1e-5
    rec = oa.magphase_combine(torch.from_numpy(out))
    assert (rec - ref).abs().max() < 1e-5
    aud = np.zeros(hop * (ref.shape[1] - 1), np.float32)
    emu.emu_istft(fptr(out), ref.shape[1], n_fft, hop, fptr(w_np), 1, 1, 1, 64, fptr(aud))
    assert np.abs(aud - oa.istft(ref, w, n_fft, hop, True).numpy()).max() < 1e-5

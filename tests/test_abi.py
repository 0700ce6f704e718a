"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports every symbol that
include/maavss_b200.h declares; host-side integer logic is bit-exact vs the reference golden table."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def built():
    from maavss_b200 import build
    return build.build()


def test_library_exports_every_declared_symbol(built):
    from maavss_b200 import _lib
    protos = _lib.parse_header()
    assert len(protos) >= 20
    L = ctypes.CDLL(built)
    for name in protos:
        assert hasattr(L, name), f"{name} declared in include/maavss_b200.h but not exported"
    out = subprocess.run(["nm", "-D", "--defined-only", built], capture_output=True, text=True).stdout
    exported = {l.split()[-1] for l in out.splitlines() if " T " in l and "maavss_" in l}
    assert set(protos) <= exported
    assert _lib.lib().maavss_abi_version() == 1


def test_sass_is_sm100a(built):
    out = subprocess.run(["cuobjdump", "-lelf", built], capture_output=True, text=True).stdout
    assert "sm_100a" in out


def test_host_logic_bit_exact(golden, built):
    from maavss_b200 import utilities, audio_ops
    for nf, hpf, fps, sr, hop, alen, nfr in golden("hop")["table"]:
        assert utilities.calc_hop_size(int(nf), int(hpf), int(fps), int(sr)) == (hop, alen, nfr)
    assert audio_ops.stft_out_dims(6336, 256, 66, True) == (96, 128)
    assert audio_ops.stft_out_dims(6336, 256, 66, False) == (96, 129)
    assert audio_ops.stft_out_dims(160000, 4096, 512, False) == (312, 2049)
    with pytest.raises(RuntimeError):
        audio_ops.stft_out_dims(100, 256, 66, True)


def test_run_config_defaults():
    from maavss_b200.run_config import model_args
    a = model_args([])
    assert (a.batch_size, a.learning_rate, a.num_frames, a.num_seq, a.hops_per_frame, a.fft_len, a.samplerate,
            a.framerate, a.framesize, a.p_size, a.latent_chan, a.fc_size, a.loss_coeff) == \
           (4, 1e-5, 8, 4, 8, 256, 16000, 30, 256, 64, 64, 4096, 0.001)
    assert model_args(["--use_polar", "False"]).use_polar is True   # type=bool quirk


def test_no_cpu_fallback(built):
    import torch
    from maavss_b200 import audio_ops
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(Exception):
        audio_ops.stft_planes(torch.zeros(1, 1000), torch.hamming_window(256), 256, 66)

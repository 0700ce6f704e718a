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


def test_envelopes_of_the_tensor_core_paths(built):
    """Host-side envelope predicates of the C ABI (pure arithmetic, no GPU): which encoder layers of the headline model take
    the 3xTF32 tensor-core path (conv_tc3.cu), which stay on the direct kernels, and the limits the header states."""
    from maavss_b200 import _specs
    from maavss_b200 import nn_ops as ops
    from maavss_b200._lib import lib
    L = lib()
    B = 64                                       # 16 clips x 4 stacked accumulation windows
    sp = _specs.fusion_specs([B, 2, 64, 128], [B, 1, 8, 4096], 64, 4096, "fc")
    took = {}
    for name, specs, hw in (("pg", sp["pg_enc"], (8, 4096)), ("st", sp["st_enc"], (64, 128))):
        h, w = hw
        for i, s in enumerate(specs):
            oh, ow = _specs.conv_out(h, s["k"][0], s["s"][0], s["p"][0]), _specs.conv_out(w, s["k"][1], s["s"][1], s["p"][1])
            desc = ops.make_conv_desc((B, s["cin"], 1, h, w), (s["cout"], s["cin"], 1, *s["k"]), (1, *s["s"]), (0, *s["p"]), (1, oh, ow))
            took[f"{name}{i + 1}"] = bool(L.maavss_conv_tc3_supported(ctypes.byref(desc)))
            if took[f"{name}{i + 1}"]:
                assert L.maavss_conv_tc3_col_elems(ctypes.byref(desc)) == B * oh * ow * s["cin"] * s["k"][0] * s["k"][1]
                assert all(L.maavss_conv_tc3_workspace_bytes(ctypes.byref(desc), m) >= 256 for m in range(3))
            h, w = oh, ow
    assert [k for k, v in took.items() if v] == ["pg5", "pg6", "pg7", "pg8", "pg9", "pg10", "st2", "st3", "st4", "st5"]
    # 3-D kernels, thin layers and ragged output maps are outside the envelope
    d3 = ops.make_conv_desc((2, 16, 8, 32, 32), (32, 16, 3, 5, 5), (1, 1, 1), (1, 2, 2), (8, 32, 32))
    assert L.maavss_conv_tc3_supported(ctypes.byref(d3)) == 0
    odd = ops.make_conv_desc((2, 32, 1, 9, 11), (32, 32, 1, 3, 3), (1, 1, 1), (0, 0, 0), (1, 7, 9))
    assert L.maavss_conv_tc3_supported(ctypes.byref(odd)) == 0
    # GEMM: pitches in multiples of 4 floats; row-major output needs Q % 4 == 0
    assert L.maavss_gemm_tf32x3_supported(64, 32768, 512, 0, 512, 0, 512, 32768, 0) == 1
    assert L.maavss_gemm_tf32x3_supported(64, 130, 512, 0, 512, 0, 512, 130, 0) == 0
    assert L.maavss_gemm_tf32x3_supported(64, 130, 512, 0, 512, 0, 512, 64, 1) == 1
    assert L.maavss_gemm_tf32x3_supported(64, 128, 510, 0, 510, 0, 510, 128, 0) == 0
    assert L.maavss_gemm_tf32x3_workspace_bytes(32768, 64, 512) == 0            # enough tiles: no split-K
    assert L.maavss_gemm_tf32x3_workspace_bytes(512, 64, 4096) > 0
    assert L.maavss_bilstm_recurrent_supported(64, 8, 256) == 1 and L.maavss_bilstm_recurrent_supported(64, 8, 128) == 0
    assert L.maavss_channel_sum_static_max_channels() == 4096

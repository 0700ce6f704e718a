#!/usr/bin/env python
"""Encoder Conv2d layers of avse_model.AV_Fusion_Model at the headline shapes (16 clips x 4 stacked windows = 64 samples):
the 3xTF32 tensor-core path (csrc/conv_tc3.cu: im2col + tc_gemm3) beside the fp32 SIMT kernels (conv_direct.cu /
conv_igemm.cu), forward / data gradient / weight gradient; device time per call from a CUDA graph of rotating buffers."""
import ctypes
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from benchmarks.gemm3_bench import time_ms  # noqa: E402
from maavss_b200 import _specs  # noqa: E402
from maavss_b200 import nn_ops as ops  # noqa: E402
from maavss_b200._lib import check, dptr, fptr, lib, stream_ptr  # noqa: E402


def layers(B=64):
    sp = _specs.fusion_specs([B, 2, 64, 128], [B, 1, 8, 4096], 64, 4096, "fc")
    out = []
    for name, specs, hw in (("pg", sp["pg_enc"], (8, 4096)), ("st", sp["st_enc"], (64, 128))):
        h, w = hw
        for i, s in enumerate(specs):
            oh, ow = _specs.conv_out(h, s["k"][0], s["s"][0], s["p"][0]), _specs.conv_out(w, s["k"][1], s["s"][1], s["p"][1])
            out.append((f"{name}{i + 1}", s["cin"], s["cout"], s["k"], s["s"], s["p"], (h, w), (oh, ow)))
            h, w = oh, ow
    return out


def main():
    B = 64
    L = lib()
    only = sys.argv[1].split(",") if len(sys.argv) > 1 else None
    for name, cin, cout, k, s, p, hw, ohw in layers(B):
        if only and name not in only:
            continue
        desc = ops.make_conv_desc((B, cin, 1, *hw), (cout, cin, 1, *k), (1, *s), (0, *p), (1, *ohw))
        tc_ok = bool(L.maavss_conv_tc3_supported(ctypes.byref(desc)))
        n_sets = 3
        x = [torch.randn(B, cin, 1, *hw, device="cuda") for _ in range(n_sets)]
        dy = [torch.randn(B, cout, 1, *ohw, device="cuda") for _ in range(n_sets)]
        w = torch.randn(cout, cin, 1, *k, device="cuda") * (cin * k[0] * k[1]) ** -0.5
        b = torch.randn(cout, device="cuda")
        y, dx, dw = torch.empty_like(dy[0]), torch.empty_like(x[0]), torch.zeros_like(w)
        flops = 2.0 * B * ohw[0] * ohw[1] * cout * cin * k[0] * k[1]
        rec = {"layer": name, "cin": cin, "cout": cout, "k": list(k), "s": list(s), "in": list(hw), "out": list(ohw), "tc3": tc_ok,
               "gflop": round(flops / 1e9, 3)}
        null = ctypes.c_void_p(None)
        rec["fwd_simt_us"] = round(1e3 * time_ms(lambda i: check(L.maavss_conv_fwd(ctypes.byref(desc), fptr(x[i]), fptr(w), fptr(b), fptr(y), stream_ptr())), n_sets), 1)
        rec["dgrad_simt_us"] = round(1e3 * time_ms(lambda i: check(L.maavss_conv_dgrad(ctypes.byref(desc), fptr(dy[i]), fptr(w), null, fptr(dx), stream_ptr())), n_sets), 1)
        rec["wgrad_simt_us"] = round(1e3 * time_ms(lambda i: check(L.maavss_conv_wgrad(ctypes.byref(desc), fptr(x[i]), fptr(dy[i]), fptr(dw), 1, stream_ptr())), n_sets), 1)
        if tc_ok:
            col = [torch.empty(L.maavss_conv_tc3_col_elems(ctypes.byref(desc)), device="cuda") for _ in range(n_sets)]
            rows = [torch.empty(B * ohw[0] * ohw[1], cout, device="cuda") for _ in range(n_sets)]
            ws = [torch.empty(max(256, L.maavss_conv_tc3_workspace_bytes(ctypes.byref(desc), m)), dtype=torch.uint8, device="cuda") for m in range(3)]
            n = [ctypes.c_longlong(t.numel()) for t in ws]
            rec["fwd_tc3_us"] = round(1e3 * time_ms(lambda i: check(L.maavss_conv_tc3_fwd(ctypes.byref(desc), fptr(x[i]), fptr(w), fptr(b), fptr(y), fptr(col[i]), dptr(ws[0]), n[0], stream_ptr())), n_sets), 1)
            rec["rows_us"] = round(1e3 * time_ms(lambda i: check(L.maavss_nchw_to_rows(fptr(dy[i]), B, cout, ohw[0] * ohw[1], fptr(rows[i]), stream_ptr())), n_sets), 1)
            wp = torch.empty_like(w)
            check(L.maavss_conv_tc3_permute_weights(ctypes.byref(desc), fptr(w), fptr(wp), stream_ptr()))
            rec["dgrad_tc3_us"] = round(1e3 * time_ms(lambda i: check(L.maavss_conv_tc3_dgrad(ctypes.byref(desc), fptr(rows[i]), fptr(wp), fptr(dx), dptr(ws[1]), n[1], stream_ptr())), n_sets), 1)
            rec["wgrad_tc3_us"] = round(1e3 * time_ms(lambda i: check(L.maavss_conv_tc3_wgrad(ctypes.byref(desc), fptr(col[i]), fptr(rows[i]), fptr(dw), 1, dptr(ws[2]), n[2], stream_ptr())), n_sets), 1)
        print(json.dumps(rec), flush=True)


if __name__ == "__main__":
    main()

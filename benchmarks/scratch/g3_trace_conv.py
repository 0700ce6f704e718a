"""Pipeline trace (MAAVSS_G3_DBG=64) of the tc_gemm3 kernel on one conv_tc3 forward GEMM."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from maavss_b200 import nn_ops as ops
from maavss_b200._lib import check, dptr, fptr, lib, stream_ptr
L = lib()
B, cin, cout, k, s, p, hw, ohw = 64, 32, 64, (1, 9), (1, 2), (0, 4), (8, 128), (8, 64)
desc = ops.make_conv_desc((B, cin, 1, *hw), (cout, cin, 1, *k), (1, *s), (0, *p), (1, *ohw))
x = torch.randn(B, cin, 1, *hw, device="cuda")
w = torch.randn(cout, cin, 1, *k, device="cuda")
b = torch.randn(cout, device="cuda")
y = torch.empty(B, cout, 1, *ohw, device="cuda")
col = torch.empty(L.maavss_conv_tc3_col_elems(ctypes.byref(desc)), device="cuda")
ws = torch.empty(max(256, L.maavss_conv_tc3_workspace_bytes(ctypes.byref(desc), 0)), dtype=torch.uint8, device="cuda")
for i in range(3):
    check(L.maavss_conv_tc3_fwd(ctypes.byref(desc), fptr(x), fptr(w), fptr(b), fptr(y), fptr(col), dptr(ws), ctypes.c_longlong(ws.numel()), stream_ptr()))
torch.cuda.synchronize()
print("---- dgrad gemm", file=sys.stderr, flush=True)
rows = torch.randn(B * ohw[0] * ohw[1], cout, device="cuda")
dx = torch.empty_like(x)
wp = torch.empty_like(w)
check(L.maavss_conv_tc3_permute_weights(ctypes.byref(desc), fptr(w), fptr(wp), stream_ptr()))
ws1 = torch.empty(max(256, L.maavss_conv_tc3_workspace_bytes(ctypes.byref(desc), 1)), dtype=torch.uint8, device="cuda")
for i in range(2):
    check(L.maavss_conv_tc3_dgrad(ctypes.byref(desc), fptr(rows), fptr(wp), fptr(dx), dptr(ws1), ctypes.c_longlong(ws1.numel()), stream_ptr()))
torch.cuda.synchronize()
print("---- wgrad gemm", file=sys.stderr, flush=True)
dw = torch.zeros_like(w)
ws2 = torch.empty(max(256, L.maavss_conv_tc3_workspace_bytes(ctypes.byref(desc), 2)), dtype=torch.uint8, device="cuda")
for i in range(2):
    check(L.maavss_conv_tc3_wgrad(ctypes.byref(desc), fptr(col), fptr(rows), fptr(dw), 1, dptr(ws2), ctypes.c_longlong(ws2.numel()), stream_ptr()))
torch.cuda.synchronize()

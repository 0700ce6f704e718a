"""maavss_channel_sum_static: the allocation-free variant of maavss_channel_sum (static device scratch inside the library;
no counterpart in the reference).  Called through the C ABI and checked against
the existing entry point -- same kernels, so bit-identical -- on two calls in a row that share the static array, plus the
size limit.  Collected last on purpose (file name): `pytest -x` must not lose other tests to a failure here."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _both(L, x, accumulate, seed_out):
    from maavss_b200._lib import check, dptr, fptr, stream_ptr
    N, C = x.shape[0], x.shape[1]
    inner = x[0, 0].numel()
    base = torch.randn(C, generator=torch.Generator().manual_seed(seed_out)).cuda()
    ref, out = base.clone(), base.clone()
    ws = torch.empty((L.maavss_reduce_workspace_bytes(C) + 7) // 8, dtype=torch.float64, device=x.device)
    check(L.maavss_channel_sum(fptr(x), N, C, inner, fptr(ref), accumulate, dptr(ws), stream_ptr()))
    free0 = torch.cuda.mem_get_info()[0]
    check(L.maavss_channel_sum_static(fptr(x), N, C, inner, fptr(out), accumulate, stream_ptr()))
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    return ref, out, base, free0 - free1


def test_channel_sum_static_matches_workspace_variant_twice():
    from maavss_b200._lib import lib
    L = lib()
    g = torch.Generator().manual_seed(5)
    # call 1: many channels, overwrite; call 2: fewer channels of another shape, accumulating -- the array still holds call 1's sums
    x1 = torch.randn(6, 96, 8, 52, generator=g).cuda()
    ref1, out1, _, _ = _both(L, x1, 0, 1)
    x2 = (torch.randn(3, 17, 5, 7, generator=g) * 3).cuda()       # inner % 4 != 0: the scalar path
    ref2, out2, base2, grew = _both(L, x2, 1, 2)
    assert torch.equal(out1, ref1)
    assert torch.equal(out2, ref2)
    # and against the plain torch fp32 op, at the tolerance of the layer tests (relative to the largest |reference|)
    for out, x, add in ((out1, x1, None), (out2, x2, base2)):
        want = x.double().sum(dim=(0, 2, 3))
        if add is not None:
            want = want + add.double()
        assert float((out.double() - want).abs().max()) / float(want.abs().max()) < 1e-6
    assert grew <= 0, f"device memory shrank by {grew} bytes across an allocation-free call"


def test_channel_sum_static_rejects_oversize():
    from maavss_b200._lib import fptr, lib, stream_ptr
    L = lib()
    cmax = L.maavss_channel_sum_static_max_channels()
    assert cmax == 4096
    x = torch.zeros(1, cmax + 1, 4, device="cuda")
    out = torch.zeros(cmax + 1, device="cuda")
    rc = L.maavss_channel_sum_static(fptr(x), 1, cmax + 1, 4, fptr(out), 0, stream_ptr())
    assert rc != 0
    msg = L.maavss_last_error()
    msg = msg.decode() if isinstance(msg, bytes) else str(msg)
    assert "4096" in msg and "channel_sum_static" in msg
    # the limit itself is accepted
    x = torch.ones(2, cmax, 4, device="cuda")
    out = torch.zeros(cmax, device="cuda")
    assert L.maavss_channel_sum_static(fptr(x), 2, cmax, 4, fptr(out), 0, stream_ptr()) == 0
    torch.cuda.synchronize()
    assert torch.equal(out, torch.full((cmax,), 8.0, device="cuda"))

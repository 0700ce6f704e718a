"""GPU parity of every model-layer kernel (through the C ABI / autograd wrappers) against the plain PyTorch fp32 op
it replaces, forward and backward, on seeded inputs.  Tolerances: fp32 kernels, relative to the largest |reference|."""
import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", autouse=True)
def fp32_reference():
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield


def rel_err(a, b):
    return float((a - b).abs().max()) / max(float(b.abs().max()), 1e-12)


def g(seed):
    return torch.Generator().manual_seed(seed)


def pair(ours, ref, seed=0):
    """Copy ref's parameters into ours (same names)."""
    ours.load_state_dict(ref.state_dict())
    return ours.cuda(), ref.cuda()


CONV_CASES = [
    # cin, cout, k, s, p, in_hw, batch
    (2, 8, (5, 5), (2, 2), (2, 2), (64, 128), 3),
    (8, 32, (5, 5), (2, 2), (2, 2), (32, 64), 2),
    (64, 64, (5, 5), (1, 2), (2, 2), (8, 16), 4),
    (1, 2, (1, 9), (1, 2), (0, 4), (8, 512), 3),
    (32, 64, (1, 9), (1, 2), (0, 4), (8, 128), 2),
    (2, 4, (3, 9), (2, 2), (1, 3), (64, 129), 2),
    (5, 7, (3, 3), (1, 1), (0, 0), (9, 11), 2),
]


@pytest.mark.parametrize("cin,cout,k,s,p,hw,B", CONV_CASES)
@pytest.mark.parametrize("bias", [True, False])
def test_conv2d(cin, cout, k, s, p, hw, B, bias):
    from maavss_b200 import layers as L
    torch.manual_seed(1)
    ref = nn.Conv2d(cin, cout, k, s, p, bias=bias)
    ours, ref = pair(L.Conv2d(cin, cout, k, s, p, bias=bias), ref)
    x = torch.randn(B, cin, *hw, generator=g(2)).cuda()
    xo, xr = x.clone().requires_grad_(), x.clone().requires_grad_()
    yo, yr = ours(xo), ref(xr)
    assert yo.shape == yr.shape
    assert rel_err(yo, yr) < 2e-6
    dy = torch.randn(yr.shape, generator=g(3)).cuda()
    yo.backward(dy)
    yr.backward(dy)
    assert rel_err(xo.grad, xr.grad) < 3e-6
    assert rel_err(ours.weight.grad, ref.weight.grad) < 1e-5
    if bias:
        assert rel_err(ours.bias.grad, ref.bias.grad) < 1e-5
    # accumulation into an existing .grad (gradient accumulation over num_seq)
    yo2 = ours(xo)
    yo2.backward(dy)
    assert rel_err(ours.weight.grad, 2 * ref.weight.grad) < 1e-5


# the mid / deep encoder layers of avse_model.AV_Fusion_Model at the headline batch (16 clips x 4 stacked windows): these run on
# the 3xTF32 tensor-core path (conv_tc3.cu); checked against torch fp32 AND against the SIMT kernels they replace
TC3_CASES = [
    (16, 32, (1, 9), (1, 2), (0, 4), (8, 256), 64),
    (64, 64, (1, 9), (1, 2), (0, 4), (8, 32), 64),
    (64, 64, (1, 9), (1, 2), (0, 4), (8, 8), 64),
    (8, 32, (5, 5), (2, 2), (2, 2), (32, 64), 64),
    (32, 64, (5, 5), (2, 2), (2, 2), (16, 32), 64),
    (64, 64, (5, 5), (1, 2), (2, 2), (8, 8), 64),
    (8, 32, (3, 9), (2, 2), (1, 4), (32, 64), 6),          # avse_model_conv / avse_model_final style kernel, ragged batch
]


@pytest.mark.parametrize("cin,cout,k,s,p,hw,B", TC3_CASES)
def test_conv2d_tensor_core_fp32(cin, cout, k, s, p, hw, B):
    import ctypes
    from maavss_b200 import layers as L
    from maavss_b200 import nn_ops as ops
    from maavss_b200._lib import lib
    oh, ow = (hw[0] + 2 * p[0] - k[0]) // s[0] + 1, (hw[1] + 2 * p[1] - k[1]) // s[1] + 1
    desc = ops.make_conv_desc((B, cin, 1, *hw), (cout, cin, 1, *k), (1, *s), (0, *p), (1, oh, ow))
    assert lib().maavss_conv_tc3_supported(ctypes.byref(desc)) == 1
    torch.manual_seed(3)
    ref = nn.Conv2d(cin, cout, k, s, p, bias=True)
    ours, ref = pair(L.Conv2d(cin, cout, k, s, p, bias=True), ref)
    x = torch.randn(B, cin, *hw, generator=g(12)).cuda()
    dy = torch.randn(B, cout, oh, ow, generator=g(13)).cuda()
    res = {}
    for tc in (True, False):
        ops.TENSOR_CORE_CONV_FP32 = tc
        try:
            ours.zero_grad(set_to_none=True)
            xo = x.clone().requires_grad_()
            yo = ours(xo)
            yo.backward(dy)
            yo2 = ours(xo)                 # second pass accumulates into the existing .grad
            yo2.backward(dy)
            res[tc] = (yo.detach(), xo.grad.clone(), ours.weight.grad.clone(), ours.bias.grad.clone())
        finally:
            ops.TENSOR_CORE_CONV_FP32 = True
    xr = x.clone().requires_grad_()
    yr = ref(xr)
    yr.backward(dy)
    for tc in (True, False):
        y, dx, dw, db = res[tc]
        assert rel_err(y, yr) < 3e-6
        assert rel_err(dx, 2 * xr.grad) < 3e-6
        assert rel_err(dw, 2 * ref.weight.grad) < 1e-5
        assert rel_err(db, 2 * ref.bias.grad) < 1e-5
    assert rel_err(res[True][0], res[False][0]) < 3e-6


def test_conv3d():
    from maavss_b200 import layers as L
    torch.manual_seed(1)
    for cin, cout, pad in ((1, 16, (1, 2, 2)), (16, 8, (1, 3, 3))):
        ref = nn.Conv3d(cin, cout, (3, 5, 5), 1, pad, bias=False)
        ours, ref = pair(L.Conv3d(cin, cout, (3, 5, 5), 1, pad, bias=False), ref)
        x = torch.randn(2, cin, 8, 20, 24, generator=g(4)).cuda()
        xo, xr = x.clone().requires_grad_(), x.clone().requires_grad_()
        yo, yr = ours(xo), ref(xr)
        assert yo.shape == yr.shape and rel_err(yo, yr) < 5e-6
        dy = torch.randn(yr.shape, generator=g(5)).cuda()
        yo.backward(dy)
        yr.backward(dy)
        assert rel_err(xo.grad, xr.grad) < 3e-6
        assert rel_err(ours.weight.grad, ref.weight.grad) < 1e-5


@pytest.mark.parametrize("cin,cout,k,s,p,op,hw", [
    (64, 16, (5, 5), (2, 2), (2, 2), (1, 1), (8, 4)),
    (2, 2, (5, 5), (1, 2), (2, 2), (0, 1), (64, 32)),
    (64, 32, (1, 9), (1, 2), (0, 4), (0, 1), (8, 4)),
    (4, 2, (3, 10), (2, 2), (1, 4), (1, 1), (32, 64)),
])
def test_conv_transpose2d(cin, cout, k, s, p, op, hw):
    from maavss_b200 import layers as L
    torch.manual_seed(1)
    ref = nn.ConvTranspose2d(cin, cout, k, s, p, output_padding=op)
    ours, ref = pair(L.ConvTranspose2d(cin, cout, k, s, p, output_padding=op), ref)
    x = torch.randn(3, cin, *hw, generator=g(6)).cuda()
    xo, xr = x.clone().requires_grad_(), x.clone().requires_grad_()
    yo, yr = ours(xo), ref(xr)
    assert yo.shape == yr.shape and rel_err(yo, yr) < 2e-6
    dy = torch.randn(yr.shape, generator=g(7)).cuda()
    yo.backward(dy)
    yr.backward(dy)
    assert rel_err(xo.grad, xr.grad) < 3e-6
    assert rel_err(ours.weight.grad, ref.weight.grad) < 1e-5
    assert rel_err(ours.bias.grad, ref.bias.grad) < 1e-5


@pytest.mark.parametrize("shape", [(4, 8, 32, 64), (16, 64, 8, 4), (3, 2, 8, 2048), (2, 16, 8, 12, 12)])
@pytest.mark.parametrize("act", ["tanh", "none", "leaky"])
def test_batchnorm_act(shape, act):
    from maavss_b200 import layers as L
    C = shape[1]
    cls_o, cls_r = (L.BatchNorm3d, nn.BatchNorm3d) if len(shape) == 5 else (L.BatchNorm2d, nn.BatchNorm2d)
    ref = cls_r(C)
    with torch.no_grad():
        ref.weight.copy_(1 + 0.1 * torch.randn(C, generator=g(8)))
        ref.bias.copy_(0.1 * torch.randn(C, generator=g(9)))
    ours, ref = pair(cls_o(C), ref)
    a_o = {"tanh": L.Tanh(), "none": nn.Identity(), "leaky": L.LeakyReLU(0.3)}[act]
    a_r = {"tanh": nn.Tanh(), "none": nn.Identity(), "leaky": nn.LeakyReLU(0.3)}[act]
    seq_o = L.FusedSequential(ours, a_o)
    x = (torch.randn(shape, generator=g(10)) * 1.5 + 0.3).cuda()
    for train in (True, False):
        seq_o.train(train)
        ref.train(train)
        xo, xr = x.clone().requires_grad_(), x.clone().requires_grad_()
        yo, yr = seq_o(xo), a_r(ref(xr))
        assert rel_err(yo, yr) < 5e-6
        dy = torch.randn(shape, generator=g(11)).cuda()
        yo.backward(dy)
        yr.backward(dy)
        assert rel_err(xo.grad, xr.grad) < 2e-5
        assert rel_err(ours.weight.grad, ref.weight.grad) < 2e-5
        assert rel_err(ours.bias.grad, ref.bias.grad) < 2e-5
        ours.weight.grad = None
        ours.bias.grad = None
        ref.zero_grad()
    assert rel_err(ours.running_mean, ref.running_mean) < 1e-6
    assert rel_err(ours.running_var, ref.running_var) < 1e-6
    assert int(ours.num_batches_tracked) == int(ref.num_batches_tracked) == 1


@pytest.mark.parametrize("M,K,N", [(16, 4096, 2048), (16, 512, 16384), (1, 2048, 512), (128, 512, 1024), (37, 100, 77),
                                   (5, 132, 45), (16, 2048, 512), (32, 8192, 100), (16, 512, 32768), (48, 260, 70),
                                   (64, 512, 16384), (64, 4096, 2048), (63, 2048, 130)])
@pytest.mark.parametrize("act", ["leaky", "tanh", "sigmoid", "none"])
def test_linear(M, K, N, act):
    from maavss_b200 import layers as L
    from maavss_b200 import nn_ops as ops
    torch.manual_seed(2)
    ref = nn.Linear(K, N)
    ours, ref = pair(L.Linear(K, N), ref)
    code = {"leaky": (ops.ACT_LEAKY, 0.3), "tanh": (ops.ACT_TANH, 0.0), "sigmoid": (ops.ACT_SIGMOID, 0.0), "none": (ops.ACT_NONE, 0.0)}[act]
    fr = {"leaky": lambda t: F.leaky_relu(t, 0.3), "tanh": torch.tanh, "sigmoid": torch.sigmoid, "none": lambda t: t}[act]
    x = torch.randn(M, K, generator=g(12)).cuda()
    xo, xr = x.clone().requires_grad_(), x.clone().requires_grad_()
    yo, yr = ours.fused_forward(xo, *code), fr(ref(xr))
    assert rel_err(yo, yr) < 5e-6
    dy = torch.randn(M, N, generator=g(13)).cuda()
    yo.backward(dy)
    yr.backward(dy)
    assert rel_err(xo.grad, xr.grad) < 1e-5
    assert rel_err(ours.weight.grad, ref.weight.grad) < 1e-5
    assert rel_err(ours.bias.grad, ref.bias.grad) < 1e-5
    # second backward accumulates into the existing .grad (num_seq accumulation, train.py:158)
    yo2 = ours.fused_forward(xo, *code)
    yo2.backward(dy)
    assert rel_err(ours.weight.grad, 2 * ref.weight.grad) < 1e-5
    assert rel_err(ours.bias.grad, 2 * ref.bias.grad) < 1e-5


@pytest.mark.parametrize("B,T,I,H", [(16, 8, 512, 256), (2, 16, 256, 256), (3, 5, 40, 64), (64, 8, 512, 256), (13, 6, 64, 256), (70, 3, 64, 96), (9, 4, 32, 32)])
def test_lstm(B, T, I, H):
    """Reference: nn.LSTM in float64 on the CPU (cuDNN's fp32 RNN may use TF32 and is itself ~1e-5 off)."""
    from maavss_b200 import layers as L
    torch.manual_seed(3)
    kw = dict(input_size=I, hidden_size=H, num_layers=1, bias=False, batch_first=True, bidirectional=True)
    ref = nn.LSTM(**kw)
    ours = L.LSTM(**kw)
    ours.load_state_dict(ref.state_dict())
    ours, ref = ours.cuda(), ref.double()
    x = torch.randn(B, T, I, generator=g(14))
    xo, xr = x.cuda().requires_grad_(), x.double().requires_grad_()
    yo, yr = ours(xo)[0], ref(xr)[0]
    assert yo.shape == yr.shape and rel_err(yo.cpu().double(), yr) < 1e-5
    dy = torch.randn(yr.shape, generator=g(15))
    yo.backward(dy.cuda())
    yr.backward(dy.double())
    assert rel_err(xo.grad.cpu().double(), xr.grad) < 2e-5
    for n in ("weight_ih_l0", "weight_hh_l0", "weight_ih_l0_reverse", "weight_hh_l0_reverse"):
        assert rel_err(getattr(ours, n).grad.cpu().double(), getattr(ref, n).grad) < 2e-5, n


@pytest.mark.parametrize("B,T", [(64, 8), (13, 6), (2, 3)])
def test_lstm_recurrent_kernels(B, T):
    """One-launch cluster kernels (lstm_persist.cu) == the per-step kernels (lstm.cu), forward and backward, through the C ABI."""
    from maavss_b200._lib import check, fptr, lib, stream_ptr
    L, H = lib(), 256
    assert L.maavss_bilstm_recurrent_supported(B, T, H) == 1 and L.maavss_bilstm_recurrent_supported(B, T, 96) == 0
    gin = torch.randn(2, B, T, 4 * H, generator=g(40)).cuda()
    w = [(torch.randn(4 * H, H, generator=g(41 + d)) * H ** -0.5).cuda() for d in range(2)]
    dy = torch.randn(B, T, 2 * H, generator=g(43)).cuda()
    out = []
    for one_launch in (False, True):
        gates, cells = torch.empty(2, B, T, 4 * H, device="cuda"), torch.empty(2, B, T, H, device="cuda")
        y, dgates = torch.empty(B, T, 2 * H, device="cuda"), torch.empty(2, B, T, 4 * H, device="cuda")
        if one_launch:
            check(L.maavss_bilstm_recurrent_fwd(fptr(gin), fptr(w[0]), fptr(w[1]), fptr(gates), fptr(cells), fptr(y), B, T, H, stream_ptr()))
            check(L.maavss_bilstm_recurrent_bwd(fptr(dy), fptr(gates), fptr(cells), fptr(w[0]), fptr(w[1]), fptr(dgates), B, T, H, stream_ptr()))
        else:
            for s_ in range(T):
                check(L.maavss_lstm_step_fwd(fptr(gin), fptr(w[0]), fptr(w[1]), fptr(gates), fptr(cells), fptr(y), B, T, H, s_, stream_ptr()))
            dcell = torch.zeros(2, B, H, device="cuda")
            for s_ in range(T - 1, -1, -1):
                check(L.maavss_lstm_step_bwd(fptr(dy), fptr(gates), fptr(cells), fptr(w[0]), fptr(w[1]), fptr(dgates), fptr(dcell), B, T, H, s_,
                                             stream_ptr()))
        torch.cuda.synchronize()
        out.append((gates, cells, y, dgates))
    for a, b in zip(out[1], out[0]):
        assert rel_err(a, b) < 2e-6


def test_mse_and_adam():
    from maavss_b200 import nn_ops as ops
    from maavss_b200 import optim
    a = torch.randn(16, 2, 64, 128, generator=g(16)).cuda().requires_grad_()
    b = torch.randn(16, 2, 64, 128, generator=g(17)).cuda()
    ar = a.detach().clone().requires_grad_()
    lo, lr_ = ops.mse_loss(a, b), F.mse_loss(ar, b)
    assert abs(float(lo) - float(lr_)) < 1e-6 * float(lr_)
    (lo * 0.25).backward()
    (lr_ * 0.25).backward()
    assert rel_err(a.grad, ar.grad) < 1e-6
    # Adam: 5 steps on three odd-sized tensors vs torch.optim.Adam
    shapes = [(33, 7), (5,), (128, 65)]
    po = [nn.Parameter(torch.randn(s, generator=g(18 + i)).cuda()) for i, s in enumerate(shapes)]
    pr = [nn.Parameter(p.detach().clone()) for p in po]
    oo = optim.Adam(po, lr=1e-3)
    orr = torch.optim.Adam(pr, lr=1e-3)
    for step in range(5):
        for i, (p, q) in enumerate(zip(po, pr)):
            gr = torch.randn(p.shape, generator=g(100 + 10 * step + i)).cuda()
            p.grad.add_(gr)
            q.grad = gr.clone()
        oo.step()
        orr.step()
        oo.zero_grad()
        orr.zero_grad()
    for p, q in zip(po, pr):
        assert rel_err(p.detach(), q.detach()) < 1e-6
    sd = oo.state_dict()
    assert set(sd["state"][0].keys()) == {"step", "exp_avg", "exp_avg_sq"}


@pytest.mark.parametrize("shape,pk", [((2, 16, 8, 64, 64), 2), ((2, 4, 3, 32, 32), 3), ((1, 3, 2, 13, 10), 3), ((2, 2, 1, 12, 12), 3)])
def test_maxpool3d(shape, pk):
    from maavss_b200 import layers as L
    x = torch.randn(*shape, generator=g(30)).cuda()
    x[0, 0, 0, :4, :4] = 1.5                      # ties: the first maximum of the window takes the gradient
    xo, xr = x.clone().requires_grad_(), x.clone().requires_grad_()
    yo, yr = L.MaxPool3d((1, pk, pk))(xo), F.max_pool3d(xr, (1, pk, pk))
    assert torch.equal(yo, yr)
    dy = torch.randn(yr.shape, generator=g(31)).cuda()
    yo.backward(dy)
    yr.backward(dy)
    assert torch.equal(xo.grad, xr.grad)


@pytest.mark.parametrize("N,Cin,Cout,D,H,W,k,p", [
    (2, 16, 32, 3, 16, 128, (3, 5, 5), (1, 2, 2)),      # visual encoder layer 2 geometry (avse_model_final.py:39)
    (3, 32, 64, 2, 8, 64, (3, 5, 5), (1, 2, 2)),
    (2, 64, 64, 4, 8, 32, (3, 5, 5), (1, 2, 2)),
    (1, 64, 16, 1, 4, 32, (1, 3, 3), (0, 1, 1)),
    (5, 16, 16, 2, 2, 64, (3, 1, 5), (1, 0, 2)),
    # the real layers 2-4 of the AV_Fusion_Model_Frames visual encoder at full geometry (avse_model_final.py:39-49), B = 2
    (2, 16, 32, 8, 128, 128, (3, 5, 5), (1, 2, 2)),
    (2, 32, 64, 8, 64, 64, (3, 5, 5), (1, 2, 2)),
    (2, 64, 64, 8, 32, 32, (3, 5, 5), (1, 2, 2)),
])
def test_conv_tensor_core(N, Cin, Cout, D, H, W, k, p):
    """tcgen05 implicit-GEMM convolution (forward and data gradient) vs torch on the same bf16-rounded operands
    (tight: only the accumulation order differs) and vs the fp32 result (bf16-operand tolerance 2e-2)."""
    from maavss_b200 import nn_ops as ops
    x = torch.randn(N, Cin, D, H, W, generator=g(31)).cuda()
    w = (torch.randn(Cout, Cin, *k, generator=g(32)) / (Cin * k[0] * k[1] * k[2]) ** 0.5).cuda()
    dy = torch.randn(N, Cout, D, H, W, generator=g(33)).cuda()
    xr, wr, dyr = (t.to(torch.bfloat16).float() for t in (x, w, dy))
    ops.TENSOR_CORE_CONV = True
    try:
        xo = x.clone().requires_grad_()
        wo = w.clone().requires_grad_()
        yo = ops.conv(xo, wo, None, (1, 1, 1), p)
        yo.backward(dy)
    finally:
        ops.TENSOR_CORE_CONV = False
    xb = xr.clone().requires_grad_()
    yb = F.conv3d(xb, wr, None, 1, p)
    assert rel_err(yo, yb) < 2e-5
    assert rel_err(yo, F.conv3d(x, w, None, 1, p)) < 2e-2
    dxb = torch.autograd.grad(F.conv3d(xb, wr, None, 1, p), xb, dyr)[0]
    assert rel_err(xo.grad, dxb) < 2e-5
    x32 = x.clone().requires_grad_()
    dx32 = torch.autograd.grad(F.conv3d(x32, w, None, 1, p), x32, dy)[0]
    assert rel_err(xo.grad, dx32) < 2e-2
    wb = wr.clone().requires_grad_()
    dwb = torch.autograd.grad(F.conv3d(xr, wb, None, 1, p), wb, dyr)[0]
    assert rel_err(wo.grad, dwb) < 5e-5
    w32 = w.clone().requires_grad_()
    dw32 = torch.autograd.grad(F.conv3d(x, w32, None, 1, p), w32, dy)[0]
    assert rel_err(wo.grad, dw32) < 2e-2
    # a second backward accumulates into .grad (gradient accumulation over num_seq)
    ops.TENSOR_CORE_CONV = True
    try:
        ops.conv(xo, wo, None, (1, 1, 1), p).backward(dy)
    finally:
        ops.TENSOR_CORE_CONV = False
    assert rel_err(wo.grad, 2 * dwb) < 5e-5


@pytest.mark.parametrize("P,Q,K,a_mn,b_mn,transposed,extras", [
    # the Linear layers of avse_model.AV_Fusion_Model with the 4 accumulation windows stacked (M = 64): forward ...
    (32768, 64, 512, 0, 0, 1, "bias_p+leaky"), (16384, 64, 512, 0, 0, 1, "bias_p+leaky"), (2048, 64, 4096, 0, 0, 1, "bias_p+leaky"),
    (512, 64, 2048, 0, 0, 1, "bias_p+tanh"),
    # ... data gradient (the weight is read MN-major, contraction over its rows) ...
    (512, 64, 32768, 1, 0, 1, ""), (4096, 64, 2048, 1, 0, 1, ""), (2048, 64, 512, 1, 0, 1, ""),
    # ... weight gradient (both operands MN-major, 64 contraction rows, row-major output accumulated into the arena)
    (32768, 512, 64, 1, 1, 0, "acc"), (2048, 4096, 64, 1, 1, 0, "acc"), (512, 2048, 64, 1, 1, 0, ""),
    # Bi-LSTM input projection (B*T = 512 rows) and ragged shapes: tails in P, Q and K, bias on q, sigmoid
    (1024, 512, 512, 0, 0, 1, ""), (200, 40, 100, 0, 0, 0, "bias_q+sigmoid"), (132, 36, 36, 1, 1, 0, "bias_p"), (96, 36, 64, 0, 1, 1, "bias_q"),
    (1, 64, 512, 0, 0, 1, ""), (300, 260, 8, 1, 0, 0, "acc"),
])
def test_gemm_tf32x3(P, Q, K, a_mn, b_mn, transposed, extras):
    """3xTF32 tensor-core GEMM (csrc/tc_gemm3.cu) against float64: fp32-grade accuracy (<= 2e-6 of the output maximum, the
    size of fp32 summation-order noise) for every operand layout, k-split and epilogue the fusion models use."""
    from maavss_b200 import nn_ops as ops
    gen = g(40)
    A = torch.randn((K, P) if a_mn else (P, K), generator=gen).cuda()
    B = torch.randn((K, Q) if b_mn else (Q, K), generator=gen).cuda()
    # adversarial magnitudes: a wide dynamic range inside every dot product exercises the hi / lo split
    A = A * torch.exp2(torch.randint(-6, 7, A.shape, generator=gen).float()).cuda()
    if "tanh" in extras or "sigmoid" in extras:
        B = B / (40.0 * K ** 0.5)            # pre-activations of order 1: a saturated output would hide the GEMM's error scale
    bias_on_q = "bias_q" in extras
    bias = torch.randn(Q if bias_on_q else P, generator=gen).cuda() if "bias" in extras else None
    act = ops.ACT_LEAKY if "leaky" in extras else ops.ACT_TANH if "tanh" in extras else ops.ACT_SIGMOID if "sigmoid" in extras else ops.ACT_NONE
    acc = "acc" in extras
    shape = (Q, P) if transposed else (P, Q)
    out0 = torch.randn(shape, generator=gen).cuda()
    out = out0.clone()
    ok = ops.gemm_tf32x3(A, a_mn, A.shape[1], B, b_mn, B.shape[1], P, Q, K, out, shape[1], transposed, bias, int(bias_on_q), act, 0.3, int(acc))
    assert ok
    Ad = (A.t() if a_mn else A).double().cpu()
    Bd = (B.t() if b_mn else B).double().cpu()
    ref = Ad @ Bd.t()
    if bias is not None:
        ref = ref + (bias.double().cpu()[None, :] if bias_on_q else bias.double().cpu()[:, None])
    ref = {ops.ACT_LEAKY: lambda t: F.leaky_relu(t, 0.3), ops.ACT_TANH: torch.tanh, ops.ACT_SIGMOID: torch.sigmoid, ops.ACT_NONE: lambda t: t}[act](ref)
    if transposed:
        ref = ref.t()
    if acc:
        ref = ref + out0.double().cpu()
    err = float((out.double().cpu() - ref).abs().max()) / float(ref.abs().max())
    # fp32 reference error for scale: torch's own fp32 GEMM (TF32 off) on the same operands
    print(f"tf32x3 P{P} Q{Q} K{K} a_mn{a_mn} b_mn{b_mn}: rel err {err:.2e}")
    assert err < 2e-6, err


def test_gemm_tf32x3_envelope():
    """Shapes the TMA path cannot address (operand pitch not a multiple of 4 floats, odd Q for a row-major output) are
    reported as unsupported, and the Linear wrappers then fall back to the fp32 SIMT kernels with the same result."""
    from maavss_b200 import nn_ops as ops
    L = ops.lib()
    assert L.maavss_gemm_tf32x3_supported(128, 64, 512, 0, 512, 0, 512, 128, 1) == 1
    assert L.maavss_gemm_tf32x3_supported(130, 64, 64, 1, 130, 0, 64, 130, 1) == 0       # MN-major pitch 130
    assert L.maavss_gemm_tf32x3_supported(128, 33, 64, 0, 64, 0, 64, 33, 0) == 0         # row-major output, Q = 33
    x = torch.randn(40, 37, generator=g(41)).cuda()          # K = 37: pitch not a multiple of 4 -> SIMT path
    w = torch.randn(50, 37, generator=g(42)).cuda()
    y = torch.empty(40, 50, device="cuda")
    ops.linear_fwd_into(x, w, None, y)
    assert rel_err(y, x @ w.t()) < 5e-6

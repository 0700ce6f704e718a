"""torch.autograd.Function wrappers over the model-layer entry points of the C ABI (include/maavss_b200.h).

Every forward / backward here is one or a few kernel launches from libmaavss_b200.so on the current CUDA stream;
there is no PyTorch fallback (CPU tensors raise in _lib.dptr).  Parameter gradients are accumulated IN PLACE into an
existing ``param.grad`` (the flat gradient arena of optim.FlatParams) -- the gradient-accumulation add of
train.py:158 is fused into the weight-gradient kernels -- and a fresh tensor is returned only when there is none.
"""
import ctypes
import os

import torch

from ._lib import check, dptr, fptr, lib, stream_ptr

ACT_NONE, ACT_TANH, ACT_LEAKY, ACT_SIGMOID = 0, 1, 2, 3

# data-parallel context: set by parallel.init(); when not None, train-mode BatchNorm statistics are all-reduced
# over this process group so that N ranks reproduce the single-process global-batch result.
SYNC_BN_GROUP = None
# number of independent sub-batches (accumulation windows batched by train_step) a train-mode BatchNorm normalises separately
BN_GROUPS = 1


# Weight gradients off the critical path.  Inside a train step nothing reads a weight gradient before the gradient exchange /
# optimizer, so with ASYNC_WGRAD set (train_step does, around its micro-steps) every weight-gradient product (conv wgrad, bias
# channel sums, Linear / LSTM wgrad) is issued on a companion stream of the stream its layer runs on; the backward chain of a
# branch is then bn_bwd -> dgrad only.  join_wgrad_streams() is the barrier in front of the all-reduce / Adam.  Off by
# default: a plain `.backward()` followed by a read of `.grad` stays correct without a join.
ASYNC_WGRAD = False
_wgrad_streams = {}
_wgrad_active = set()


class _WgradScope:
    def __init__(self, *tensors):
        self.tensors, self.ctx, self.side = tensors, None, None

    def __enter__(self):
        if not ASYNC_WGRAD:
            return self
        cur = torch.cuda.current_stream()
        key = (cur.device_index, cur.cuda_stream)
        side = _wgrad_streams.get(key)
        if side is None:
            side = _wgrad_streams[key] = torch.cuda.Stream(device=cur.device)
        side.wait_stream(cur)
        for t in self.tensors:
            if t is not None:
                t.record_stream(side)
        _wgrad_active.add(side)
        self.side = side
        self.ctx = torch.cuda.stream(side)
        self.ctx.__enter__()
        return self

    def __exit__(self, *exc):
        if self.ctx is not None:
            self.ctx.__exit__(*exc)
        return False


def join_wgrad_streams(target=None, clear=True):
    """Make `target` (default: the current stream) wait for every outstanding weight-gradient stream."""
    tgt = target if target is not None else torch.cuda.current_stream()
    for s in list(_wgrad_active):
        tgt.wait_stream(s)
    if clear:
        _wgrad_active.clear()


class ConvDesc(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int) for n in "N Cin D H W Cout KD KH KW SD SH SW PD PH PW OD OH OW".split()]


def _triple(v, fill):
    v = tuple(v) if isinstance(v, (tuple, list)) else (v, v)
    return (fill,) + v if len(v) == 2 else v


def _null():
    return ctypes.c_void_p(None)


def _opt(t):
    return fptr(t) if t is not None else _null()


def _ws(nbytes, device):
    return torch.empty((nbytes + 7) // 8, dtype=torch.float64, device=device)


def _grad_slot(param, needs):
    """-> (tensor to write, accumulate flag, value to hand back to autograd)."""
    if not needs:
        return None, 0, None
    g = param.grad
    if g is not None and g.is_cuda and g.is_contiguous() and g.dtype == torch.float32:
        return g, 1, None
    new = torch.empty_like(param, memory_format=torch.contiguous_format)
    return new, 0, new


def make_conv_desc(in_shape5, w_shape5, stride, padding, out_shape3):
    N, Cin, D, H, W = in_shape5
    Cout, _, KD, KH, KW = w_shape5
    return ConvDesc(N, Cin, D, H, W, Cout, KD, KH, KW, *stride, *padding, *out_shape3)


def _as5(t):
    return t if t.dim() == 5 else t.unsqueeze(2)


def channel_sum(x5, out, accumulate):
    N, C = x5.shape[0], x5.shape[1]
    inner = x5[0, 0].numel() if N else 0
    ws = _ws(lib().maavss_reduce_workspace_bytes(C), x5.device)
    check(lib().maavss_channel_sum(fptr(x5), N, C, inner, fptr(out), accumulate, dptr(ws), stream_ptr()))


# Tensor-core (tcgen05, bf16 operands / fp32 accumulate) path for the stride-1 "same" Conv3d / Conv2d layers with
# 16 / 32 / 64 channels (the visual encoder of AV_Fusion_Model_Frames): off by default so that the fp32 parity
# tolerances of the model tests hold; the Frames train step / benchmark switch it on (tolerance 2e-2).
TENSOR_CORE_CONV = False


def _conv_tc(desc, mode, act5, w5, out5):
    """Returns True when the layer ran on the tensor-core kernel."""
    if not TENSOR_CORE_CONV or not lib().maavss_conv_tc_supported(ctypes.byref(desc), mode):
        return False
    nbytes = lib().maavss_conv_tc_workspace_bytes(ctypes.byref(desc), mode)
    ws = torch.empty(nbytes + 1024, device=act5.device, dtype=torch.uint8)
    off = (-ws.data_ptr()) % 1024
    check(lib().maavss_conv_tc(ctypes.byref(desc), mode, fptr(act5), fptr(w5), fptr(out5), ctypes.c_void_p(ws.data_ptr() + off),
                               ctypes.c_longlong(nbytes), stream_ptr()))
    return True


def _conv_wgrad_tc(desc, x5, dy5, tgt, acc):
    if not TENSOR_CORE_CONV or not lib().maavss_conv_wgrad_tc_supported(ctypes.byref(desc)):
        return False
    nbytes = lib().maavss_conv_wgrad_tc_workspace_bytes(ctypes.byref(desc))
    ws = torch.empty(nbytes + 1024, device=x5.device, dtype=torch.uint8)
    off = (-ws.data_ptr()) % 1024
    check(lib().maavss_conv_wgrad_tc(ctypes.byref(desc), fptr(x5), fptr(dy5), fptr(tgt), acc, ctypes.c_void_p(ws.data_ptr() + off),
                                     ctypes.c_longlong(nbytes), stream_ptr()))
    return True


# fp32-accurate tensor-core path (3xTF32) for the mid / deep Conv2d layers of the fusion encoders (conv_tc3.cu): on by default,
# results within the fp32 tolerances of the layer / model tests.
TENSOR_CORE_CONV_FP32 = os.environ.get("MAAVSS_NO_TC3", "0") in ("", "0")


def ctx_desc_ref(desc):
    return ctypes.byref(desc)


def _conv_tc3_ok(desc, x5):
    return TENSOR_CORE_CONV_FP32 and x5.is_cuda and bool(lib().maavss_conv_tc3_supported(ctypes.byref(desc)))


def _ws_bytes(nbytes, device):
    return torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=device)


class ConvFn(torch.autograd.Function):
    """nn.Conv2d / nn.Conv3d forward + backward (zero padding, no dilation / groups)."""

    @staticmethod
    def forward(ctx, x, weight, bias, stride, padding):
        nd = x.dim()
        x5, w5 = _as5(x.contiguous()), _as5(weight)
        s, p = _triple(stride, 1), _triple(padding, 0)
        out3 = tuple((x5.shape[2 + i] + 2 * p[i] - w5.shape[2 + i]) // s[i] + 1 for i in range(3))
        desc = make_conv_desc(x5.shape, w5.shape, s, p, out3)
        y = torch.empty((x5.shape[0], w5.shape[0]) + out3, device=x.device, dtype=torch.float32)
        col = None
        if bias is None and _conv_tc(desc, 0, x5, w5.contiguous(), y):
            pass                                   # opt-in bf16 tensor-core path (TENSOR_CORE_CONV) takes the layers it supports
        elif _conv_tc3_ok(desc, x5):
            L = lib()
            col = torch.empty(L.maavss_conv_tc3_col_elems(ctypes.byref(desc)), device=x.device, dtype=torch.float32)
            nb = L.maavss_conv_tc3_workspace_bytes(ctypes.byref(desc), 0)
            ws = _ws_bytes(nb, x.device)
            check(L.maavss_conv_tc3_fwd(ctypes.byref(desc), fptr(x5), fptr(w5.contiguous()), _opt(bias), fptr(y), fptr(col), dptr(ws),
                                        ctypes.c_longlong(ws.numel()), stream_ptr()))
        else:
            check(lib().maavss_conv_fwd(ctypes.byref(desc), fptr(x5), fptr(w5.contiguous()), _opt(bias), fptr(y), stream_ptr()))
        ctx.desc, ctx.nd = desc, nd
        ctx.tc3 = col is not None
        if col is not None:
            # tap-major copy of the weights for the data-gradient product: off the forward chain (companion stream)
            w_perm = torch.empty_like(w5, memory_format=torch.contiguous_format)
            scope = _WgradScope(w_perm)
            with scope:
                check(lib().maavss_conv_tc3_permute_weights(ctx_desc_ref(desc), fptr(w5.contiguous()), fptr(w_perm), stream_ptr()))
            ctx.perm_stream = scope.side
            ctx.save_for_backward(x5, weight, bias, col, w_perm)
        else:
            ctx.save_for_backward(x5, weight, bias)
        return y if nd == 5 else y.squeeze(2)

    @staticmethod
    def backward(ctx, dy):
        if ctx.tc3:
            return ConvFn._backward_tc3(ctx, dy)
        x5, weight, bias = ctx.saved_tensors
        dy5 = _as5(dy.contiguous())
        desc = ctx.desc
        dx = gw = gb = None
        if ctx.needs_input_grad[0]:
            dx5 = torch.empty_like(x5)
            if not _conv_tc(desc, 1, dy5, _as5(weight).contiguous(), dx5):
                check(lib().maavss_conv_dgrad(ctypes.byref(desc), fptr(dy5), fptr(weight), _null(), fptr(dx5), stream_ptr()))
            dx = dx5 if ctx.nd == 5 else dx5.squeeze(2)
        with _WgradScope(dy5, x5):
            tgt, acc, gw = _grad_slot(weight, ctx.needs_input_grad[1])
            if tgt is not None and not _conv_wgrad_tc(desc, x5, dy5, tgt, acc):
                check(lib().maavss_conv_wgrad(ctypes.byref(desc), fptr(x5), fptr(dy5), fptr(tgt), acc, stream_ptr()))
            if bias is not None:
                tgt, acc, gb = _grad_slot(bias, ctx.needs_input_grad[2])
                if tgt is not None:
                    channel_sum(dy5, tgt, acc)
        return dx, gw, gb, None, None


def _conv_backward_tc3(ctx, dy):
    """Backward of a layer whose forward ran on conv_tc3: dy with the channel axis last feeds both GEMMs."""
    x5, weight, bias, col, w_perm = ctx.saved_tensors
    if ctx.perm_stream is not None:
        torch.cuda.current_stream().wait_stream(ctx.perm_stream)
    dy5 = _as5(dy.contiguous())
    desc, L, dev = ctx.desc, lib(), dy.device
    N, Cout = dy5.shape[0], dy5.shape[1]
    inner = dy5[0, 0].numel()
    dx = gw = gb = None
    dy_rows = torch.empty(N * inner, Cout, device=dev, dtype=torch.float32)
    check(L.maavss_nchw_to_rows(fptr(dy5), N, Cout, inner, fptr(dy_rows), stream_ptr()))
    if ctx.needs_input_grad[0]:
        dx5 = torch.empty_like(x5)
        ws = _ws_bytes(L.maavss_conv_tc3_workspace_bytes(ctypes.byref(desc), 1), dev)
        check(L.maavss_conv_tc3_dgrad(ctypes.byref(desc), fptr(dy_rows), fptr(w_perm), fptr(dx5), dptr(ws), ctypes.c_longlong(ws.numel()),
                                      stream_ptr()))
        dx = dx5 if ctx.nd == 5 else dx5.squeeze(2)
    with _WgradScope(dy5, dy_rows, col):
        tgt, acc, gw = _grad_slot(weight, ctx.needs_input_grad[1])
        if tgt is not None:
            ws2 = _ws_bytes(L.maavss_conv_tc3_workspace_bytes(ctypes.byref(desc), 2), dev)
            check(L.maavss_conv_tc3_wgrad(ctypes.byref(desc), fptr(col), fptr(dy_rows), fptr(tgt), acc, dptr(ws2),
                                          ctypes.c_longlong(ws2.numel()), stream_ptr()))
        if bias is not None:
            tgt, acc, gb = _grad_slot(bias, ctx.needs_input_grad[2])
            if tgt is not None:
                channel_sum(dy5, tgt, acc)
    return dx, gw, gb, None, None


ConvFn._backward_tc3 = staticmethod(_conv_backward_tc3)


class ConvTransposeFn(torch.autograd.Function):
    """nn.ConvTranspose2d forward = the input-gradient kernel of the mirrored convolution."""

    @staticmethod
    def forward(ctx, x, weight, bias, stride, padding, output_padding):
        nd = x.dim()
        x5, w5 = _as5(x.contiguous()), _as5(weight)          # weight: (C_in_t, C_out_t, k..) == conv (Cout, Cin, k..)
        s, p, op = _triple(stride, 1), _triple(padding, 0), _triple(output_padding, 0)
        out3 = tuple((x5.shape[2 + i] - 1) * s[i] - 2 * p[i] + w5.shape[2 + i] + op[i] for i in range(3))
        desc = make_conv_desc((x5.shape[0], w5.shape[1]) + out3, w5.shape, s, p, tuple(x5.shape[2:]))
        y = torch.empty((x5.shape[0], w5.shape[1]) + out3, device=x.device, dtype=torch.float32)
        check(lib().maavss_conv_dgrad(ctypes.byref(desc), fptr(x5), fptr(w5.contiguous()), _opt(bias), fptr(y), stream_ptr()))
        ctx.desc, ctx.nd = desc, nd
        ctx.save_for_backward(x5, weight, bias)
        return y if nd == 5 else y.squeeze(2)

    @staticmethod
    def backward(ctx, dy):
        x5, weight, bias = ctx.saved_tensors
        dy5 = _as5(dy.contiguous())
        desc = ctx.desc
        dx = gw = gb = None
        if ctx.needs_input_grad[0]:
            dx5 = torch.empty_like(x5)
            check(lib().maavss_conv_fwd(ctypes.byref(desc), fptr(dy5), fptr(weight), _null(), fptr(dx5), stream_ptr()))
            dx = dx5 if ctx.nd == 5 else dx5.squeeze(2)
        tgt, acc, gw = _grad_slot(weight, ctx.needs_input_grad[1])
        if tgt is not None:
            check(lib().maavss_conv_wgrad(ctypes.byref(desc), fptr(dy5), fptr(x5), fptr(tgt), acc, stream_ptr()))
        if bias is not None:
            tgt, acc, gb = _grad_slot(bias, ctx.needs_input_grad[2])
            if tgt is not None:
                channel_sum(dy5, tgt, acc)
        return dx, gw, gb, None, None, None


def _allreduce_partial(partial):
    import torch.distributed as dist
    from . import parallel
    if parallel.peer_exchange() is not None and partial.is_cuda and partial.numel() <= parallel.peer_capacity():
        # One single-CTA kernel over NVLink peer memory per exchange.  All exchanges of a rank run on ONE dedicated stream in
        # host issue order -- the same order on every rank -- whichever branch stream they belong to: two spinning exchange
        # kernels of different branches that the hardware happened to order differently on two ranks would wait for each
        # other forever (concurrent execution of graph branches is not guaranteed).
        cur = torch.cuda.current_stream(partial.device)
        ex = parallel.exchange_stream(partial.device)
        ex.wait_stream(cur)
        with torch.cuda.stream(ex):
            parallel.peer_allreduce_sum_(partial, 0)
        partial.record_stream(ex)
        cur.wait_stream(ex)
    elif parallel.native_comm() is not None and partial.is_cuda:
        parallel.allreduce_sum_(partial)
    else:
        dist.all_reduce(partial, group=SYNC_BN_GROUP)
    return dist.get_world_size(SYNC_BN_GROUP)


class BNActFn(torch.autograd.Function):
    """act(BatchNorm(x)) with batch statistics in training (running stats in eval), any channels-first rank.

    Training runs the two-launch grouped kernels: `BN_GROUPS` consecutive sub-batches of the batch dimension are
    normalised independently (the accumulation windows of one optimizer step batched by train_step); with a
    data-parallel group set, the (groups, C, S, 2) partial sums are all-reduced between the two launches."""

    @staticmethod
    def forward(ctx, x, gamma, beta, running_mean, running_var, nbt, training, momentum, eps, act, slope):
        x = x.contiguous()
        N, C = x.shape[0], x.shape[1]
        inner = x[0, 0].numel() if N else 0
        L = lib()
        use_batch = training or running_mean is None
        y = torch.empty_like(x)
        if use_batch:
            G = BN_GROUPS if (training and N % max(BN_GROUPS, 1) == 0) else 1
            S = L.maavss_bn_splits(N // G, C, inner)
            partial = torch.empty(G, C, S, 2, device=x.device, dtype=torch.float64)
            stats = torch.empty(G, C, 2, device=x.device, dtype=torch.float32)
            check(L.maavss_bn_reduce_grouped(fptr(x), N, C, inner, G, dptr(partial), stream_ptr()))
            world = 1
            if SYNC_BN_GROUP is not None and training:
                world = _allreduce_partial(partial)
            count = float((N // G) * inner * world)
            check(L.maavss_bn_apply_grouped(fptr(x), dptr(partial), N, C, inner, G, count, _opt(gamma), _opt(beta),
                                            _opt(running_mean) if training else _null(), _opt(running_var) if training else _null(),
                                            dptr(nbt) if (nbt is not None and training) else _null(),
                                            momentum if momentum is not None else 0.1, eps, act, slope, fptr(stats), fptr(y),
                                            stream_ptr()))
        else:
            G, world = 1, 1
            stats = torch.empty(1, C, 2, device=x.device, dtype=torch.float32)
            check(L.maavss_bn_finalize(_null(), 0, C, 1.0, eps, 0.0, fptr(running_mean), fptr(running_var), _null(), 1,
                                       fptr(stats), stream_ptr()))
            check(L.maavss_bn_apply_act(fptr(x), fptr(stats), _opt(gamma), _opt(beta), N, C, inner, act, slope, fptr(y), stream_ptr()))
        ctx.cfg = (N, C, inner, act, slope, use_batch, world, G)
        ctx.save_for_backward(x, y, stats, gamma, beta)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, y, stats, gamma, beta = ctx.saved_tensors
        N, C, inner, act, slope, use_batch, world, G = ctx.cfg
        dy = dy.contiguous()
        L = lib()
        gg_t, gg_acc, gg = _grad_slot(gamma, gamma is not None and ctx.needs_input_grad[1])
        gb_t, gb_acc, gb = _grad_slot(beta, beta is not None and ctx.needs_input_grad[2])
        if gg_t is not None and gb_t is not None and gg_acc != gb_acc:      # keep one accumulate flag for the fused kernel
            gb_t, gb_acc = torch.empty_like(beta), 0
            gb = gb_t
            if gg_acc:
                gg_t, gg_acc = torch.empty_like(gamma), 0
                gg = gg_t
        acc = gg_acc if gg_t is not None else gb_acc
        dx = torch.empty_like(x) if ctx.needs_input_grad[0] else None
        if use_batch:
            S = L.maavss_bn_splits(N // G, C, inner)
            partial = torch.empty(G, C, S, 2, device=x.device, dtype=torch.float64)
            check(L.maavss_bn_bwd_reduce_grouped(fptr(dy), fptr(y), fptr(x), fptr(stats), N, C, inner, G, act, slope, dptr(partial),
                                                 stream_ptr()))
            local = partial
            if world > 1:
                # parameter gradients come from the LOCAL sums (the gradient all-reduce averages them); the input
                # gradient needs the GLOBAL sums.
                local = partial.clone()
                _allreduce_partial(partial)
            check(L.maavss_bn_bwd_apply_grouped(fptr(dy), fptr(y), fptr(x), fptr(stats), _opt(gamma), dptr(partial), dptr(local), N, C,
                                                inner, G, float((N // G) * inner * world), act, slope, _opt(gg_t), _opt(gb_t), acc,
                                                fptr(dx) if dx is not None else _null(), stream_ptr()))
        else:
            S = L.maavss_bn_splits(N, C, inner)
            partial = torch.empty(C, S, 2, device=x.device, dtype=torch.float64)
            check(L.maavss_bn_bwd_reduce(fptr(dy), fptr(y), fptr(x), fptr(stats), N, C, inner, act, slope, dptr(partial), stream_ptr()))
            coef = torch.empty(C, 2, device=x.device, dtype=torch.float32)
            check(L.maavss_bn_bwd_finalize(dptr(partial), S, C, float(N * inner), 1, fptr(coef), _opt(gg_t), _opt(gb_t), acc,
                                           stream_ptr()))
            if dx is not None:
                check(L.maavss_bn_bwd_apply(fptr(dy), fptr(y), fptr(x), fptr(stats), _opt(gamma), fptr(coef), N, C, inner, act, slope,
                                            fptr(dx), stream_ptr()))
        return dx, gg, gb, None, None, None, None, None, None, None, None


class ActFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, act, slope):
        x = x.contiguous()
        y = torch.empty_like(x)
        check(lib().maavss_act_fwd(fptr(x), fptr(y), x.numel(), act, slope, stream_ptr()))
        ctx.cfg = (act, slope)
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        dy = dy.contiguous()
        dz = torch.empty_like(dy)
        check(lib().maavss_act_bwd(fptr(dy), fptr(y), fptr(dz), y.numel(), ctx.cfg[0], ctx.cfg[1], stream_ptr()))
        return dz, None, None


class MaxPoolFn(torch.autograd.Function):
    """nn.MaxPool3d((1, ph, pw)) / nn.MaxPool2d((ph, pw)) with stride = kernel (floor mode) over the last two dims."""

    @staticmethod
    def forward(ctx, x, ph, pw):
        x = x.contiguous()
        H, W = x.shape[-2], x.shape[-1]
        planes = x.numel() // (H * W) if H * W else 0
        y = torch.empty(x.shape[:-2] + (H // ph, W // pw), device=x.device, dtype=torch.float32)
        idx = torch.empty(y.shape, device=x.device, dtype=torch.uint8)
        check(lib().maavss_maxpool_fwd(fptr(x), fptr(y), dptr(idx), planes, H, W, ph, pw, stream_ptr()))
        ctx.cfg = (tuple(x.shape), planes, H, W, ph, pw)
        ctx.save_for_backward(idx)
        return y

    @staticmethod
    def backward(ctx, dy):
        (idx,) = ctx.saved_tensors
        shape, planes, H, W, ph, pw = ctx.cfg
        dx = torch.empty(shape, device=dy.device, dtype=torch.float32)
        check(lib().maavss_maxpool_bwd(fptr(dy.contiguous()), dptr(idx), fptr(dx), planes, H, W, ph, pw, stream_ptr()))
        return dx, None, None


# fp32-accurate tensor-core path (3xTF32 on tcgen05, csrc/tc_gemm3.cu) for the Linear / LSTM-projection products with at
# least this many rows (the accumulation windows stacked by train_step give M = 64); fewer rows stay on the weight-streaming
# SIMT kernels (linear_skinny.cu).  Accuracy is fp32 grade, so it is on by default.
TENSOR_CORE_LINEAR = True
TC_LINEAR_MIN_ROWS = 32


def gemm_tf32x3(A, a_mn, lda, B, b_mn, ldb, P, Q, K, out, ldc, transposed, bias=None, bias_on_q=0, act=ACT_NONE, slope=0.0,
                accumulate=0):
    """D[P, Q] = A . B^T through maavss_gemm_tf32x3 (see include/maavss_b200.h); returns False when the shape is outside
    the kernel's envelope (the caller then uses the SIMT kernels)."""
    L = lib()
    if not L.maavss_gemm_tf32x3_supported(P, Q, K, a_mn, lda, b_mn, ldb, ldc, transposed):
        return False
    if (A.data_ptr() | B.data_ptr() | out.data_ptr() | (bias.data_ptr() if bias is not None else 0)) & 15:
        return False
    nbytes = L.maavss_gemm_tf32x3_workspace_bytes(P, Q, K)
    ws = torch.empty(max(nbytes, 16) // 4, device=out.device, dtype=torch.float32)
    check(L.maavss_gemm_tf32x3(fptr(A), a_mn, lda, fptr(B), b_mn, ldb, P, Q, K, fptr(out), ldc, transposed, _opt(bias), bias_on_q, act,
                               slope, accumulate, fptr(ws), ctypes.c_longlong(nbytes), stream_ptr()))
    return True


def _tc_linear(M):
    return TENSOR_CORE_LINEAR and M >= TC_LINEAR_MIN_ROWS


def linear_fwd_into(x2, weight, bias, y, act=ACT_NONE, slope=0.0):
    """y[M, N] = act(x2[M, K] weight[N, K]^T + bias)."""
    M, K = x2.shape
    N = weight.shape[0]
    if _tc_linear(M) and gemm_tf32x3(weight, 0, K, x2, 0, K, N, M, K, y, N, 1, bias, 0, act, slope, 0):
        return
    check(lib().maavss_linear_fwd(fptr(x2), fptr(weight), _opt(bias), fptr(y), M, N, K, act, slope, stream_ptr()))


def linear_dgrad_into(dz, weight, dx2):
    """dx2[M, K] = dz[M, N] weight[N, K]."""
    M, N = dz.shape
    K = weight.shape[1]
    if _tc_linear(M) and gemm_tf32x3(weight, 1, K, dz, 0, N, K, M, N, dx2, K, 1):
        return
    check(lib().maavss_linear_dgrad(fptr(dz), fptr(weight), fptr(dx2), M, N, K, stream_ptr()))


def linear_wgrad_into(dz, x2, dw, dbias, accumulate):
    """dw[N, K] (+)= dz[M, N]^T x2[M, K];  dbias[N] (+)= column sums of dz (dbias may be None)."""
    M, N = dz.shape
    K = x2.shape[1]
    if _tc_linear(M) and gemm_tf32x3(dz, 1, N, x2, 1, K, N, K, M, dw, K, 0, None, 0, ACT_NONE, 0.0, accumulate):
        if dbias is not None:
            check(lib().maavss_colsum(fptr(dz), M, N, fptr(dbias), accumulate, stream_ptr()))
        return
    check(lib().maavss_linear_wgrad(fptr(dz), fptr(x2), fptr(dw), _opt(dbias), M, N, K, accumulate, stream_ptr()))


class LinearActFn(torch.autograd.Function):
    """act(x w^T + b) over the last dimension of x."""

    @staticmethod
    def forward(ctx, x, weight, bias, act, slope):
        shp = x.shape
        x2 = x.contiguous().reshape(-1, shp[-1])
        M, K = x2.shape
        N = weight.shape[0]
        y = torch.empty(M, N, device=x.device, dtype=torch.float32)
        linear_fwd_into(x2, weight, bias, y, act, slope)
        ctx.cfg = (shp, M, N, K, act, slope)
        ctx.save_for_backward(x2, weight, bias, y)
        return y.reshape(*shp[:-1], N)

    @staticmethod
    def backward(ctx, dy):
        x2, weight, bias, y = ctx.saved_tensors
        shp, M, N, K, act, slope = ctx.cfg
        dz = dy.contiguous().reshape(M, N)
        L = lib()
        if act != ACT_NONE:
            t = torch.empty_like(dz)
            check(L.maavss_act_bwd(fptr(dz), fptr(y), fptr(t), dz.numel(), act, slope, stream_ptr()))
            dz = t
        dx = gw = gb = None
        if ctx.needs_input_grad[0]:
            dx2 = torch.empty(M, K, device=dz.device, dtype=torch.float32)
            linear_dgrad_into(dz, weight, dx2)
            dx = dx2.reshape(shp)
        w_t, w_acc, gw = _grad_slot(weight, ctx.needs_input_grad[1])
        b_t, b_acc, gb = _grad_slot(bias, bias is not None and ctx.needs_input_grad[2])
        with _WgradScope(dz, x2):
            if w_t is not None:
                if b_t is not None and b_acc != w_acc:          # mixed: give the bias its own pass below
                    linear_wgrad_into(dz, x2, w_t, None, w_acc)
                    check(L.maavss_colsum(fptr(dz), M, N, fptr(b_t), b_acc, stream_ptr()))
                else:
                    linear_wgrad_into(dz, x2, w_t, b_t, w_acc)
            elif b_t is not None:
                check(L.maavss_colsum(fptr(dz), M, N, fptr(b_t), b_acc, stream_ptr()))
        return dx, gw, gb, None, None


class LSTMFn(torch.autograd.Function):
    """nn.LSTM(bias=False, batch_first=True, bidirectional=True, num_layers=1): x (B, T, I) -> y (B, T, 2H).
    One C-ABI call per direction of autograd (maavss_bilstm_fwd / maavss_bilstm_bwd): projections, recurrent steps, the
    h_{t-1} gather and every gradient product are enqueued by the library."""

    @staticmethod
    def forward(ctx, x, w_ih_f, w_hh_f, w_ih_r, w_hh_r):
        x = x.contiguous()
        B, T, I = x.shape
        H = w_hh_f.shape[1]
        L = lib()
        dev = x.device
        gin = torch.empty(2, B, T, 4 * H, device=dev, dtype=torch.float32)
        gates = torch.empty_like(gin)
        cells = torch.empty(2, B, T, H, device=dev, dtype=torch.float32)
        y = torch.empty(B, T, 2 * H, device=dev, dtype=torch.float32)
        nbytes = L.maavss_bilstm_workspace_bytes(B, T, I, H, 0)
        ws = torch.empty(nbytes, device=dev, dtype=torch.uint8)
        check(L.maavss_bilstm_fwd(fptr(x), fptr(w_ih_f), fptr(w_hh_f), fptr(w_ih_r), fptr(w_hh_r), fptr(gin), fptr(gates), fptr(cells),
                                  fptr(y), B, T, I, H, dptr(ws), ctypes.c_longlong(nbytes), stream_ptr()))
        ctx.cfg = (B, T, I, H)
        ctx.save_for_backward(x, w_ih_f, w_hh_f, w_ih_r, w_hh_r, gates, cells, y)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w_ih_f, w_hh_f, w_ih_r, w_hh_r, gates, cells, y = ctx.saved_tensors
        B, T, I, H = ctx.cfg
        L = lib()
        dev = dy.device
        dy = dy.contiguous()
        dgates = torch.empty_like(gates)
        slots = [_grad_slot(w, ctx.needs_input_grad[i + 1]) for i, w in enumerate((w_ih_f, w_hh_f, w_ih_r, w_hh_r))]
        accs = {a for t, a, _ in slots if t is not None}
        if len(accs) > 1:            # mixed accumulate flags (a caller dropped some .grad): fresh tensors for all, added below
            slots = [(torch.empty_like(w), 0, "add") if t is not None else (None, 0, None)
                     for (t, a, r), w in zip(slots, (w_ih_f, w_hh_f, w_ih_r, w_hh_r))]
        acc = next(iter({a for t, a, _ in slots if t is not None}), 0)
        dx = torch.empty(B, T, I, device=dev, dtype=torch.float32) if ctx.needs_input_grad[0] else None
        nbytes = L.maavss_bilstm_workspace_bytes(B, T, I, H, 1)
        ws = torch.empty(nbytes, device=dev, dtype=torch.uint8)
        # recurrent steps + input gradient on this stream; the four weight gradients on the companion stream
        check(L.maavss_bilstm_bwd(fptr(dy), fptr(x), fptr(w_ih_f), fptr(w_hh_f), fptr(w_ih_r), fptr(w_hh_r), fptr(gates), fptr(cells),
                                  fptr(y), fptr(dgates), _opt(dx), _null(), _null(), _null(), _null(), acc, B, T, I, H, dptr(ws),
                                  ctypes.c_longlong(nbytes), stream_ptr()))
        if any(t is not None for t, _, _ in slots):
            with _WgradScope(dgates, x, y):
                ws2 = torch.empty(nbytes, device=dev, dtype=torch.uint8)
                check(L.maavss_bilstm_wgrad(fptr(x), fptr(y), fptr(dgates), *[_opt(t) for t, _, _ in slots], acc, B, T, I, H, dptr(ws2),
                                            ctypes.c_longlong(nbytes), stream_ptr()))
        grads = [dx]
        for (t, a, ret), w in zip(slots, (w_ih_f, w_hh_f, w_ih_r, w_hh_r)):
            grads.append(t if isinstance(ret, str) or ret is not None else None)
        return tuple(grads)


class MSEFn(torch.autograd.Function):
    """nn.MSELoss() (mean reduction)."""

    @staticmethod
    def forward(ctx, a, b):
        a, b = a.contiguous(), b.contiguous()
        if a.shape != b.shape:
            raise RuntimeError(f"mse_loss: shapes differ {tuple(a.shape)} vs {tuple(b.shape)}")
        loss = torch.empty((), device=a.device, dtype=torch.float32)
        ws = _ws(lib().maavss_reduce_workspace_bytes(8), a.device)
        check(lib().maavss_mse_fwd(fptr(a), fptr(b), a.numel(), fptr(loss), dptr(ws), stream_ptr()))
        ctx.save_for_backward(a, b)
        return loss

    @staticmethod
    def backward(ctx, gout):
        a, b = ctx.saved_tensors
        gout = gout.contiguous().to(torch.float32)
        da = db = None
        if ctx.needs_input_grad[0] or ctx.needs_input_grad[1]:
            da = torch.empty_like(a)
            check(lib().maavss_mse_bwd(fptr(a), fptr(b), fptr(gout), a.numel(), fptr(da), 0, stream_ptr()))
            if ctx.needs_input_grad[1]:
                db = -da
            if not ctx.needs_input_grad[0]:
                da = None
        return da, db


def mse_loss(a, b):
    return MSEFn.apply(a, b)


_consts = {}


def const_scalar(value, device):
    """A cached fp32 device scalar (gradient seeds of the train steps: no fill kernel per step)."""
    key = (float(value), device.type, device.index)
    t = _consts.get(key)
    if t is None:
        t = _consts[key] = torch.full((), float(value), device=device, dtype=torch.float32)
    return t


def mse_grad(a, b, seed):
    """d(seed * mean((a - b)^2)) / da without the forward reduction: what MSEFn.backward returns for `gout = seed` (a device
    scalar).  The train steps seed the backward pass with it, so the loss VALUE (reporting only) leaves the critical path."""
    a, b = a.detach().contiguous(), b.detach().contiguous()
    if a.shape != b.shape:
        raise RuntimeError(f"mse_grad: shapes differ {tuple(a.shape)} vs {tuple(b.shape)}")
    da = torch.empty_like(a)
    check(lib().maavss_mse_bwd(fptr(a), fptr(b), fptr(seed), a.numel(), fptr(da), 0, stream_ptr()))
    return da


def stack_windows(x, windows, hop, span):
    """x (B, C, T, F) -> (windows*B, C, span, F): window j = x[:, :, j*hop : j*hop + span] in rows [j*B, (j+1)*B).  Inputs only (no autograd)."""
    x = x.detach().contiguous()
    B, C, T, F = x.shape
    out = torch.empty(windows * B, C, span, F, device=x.device, dtype=torch.float32)
    check(lib().maavss_stack_windows(fptr(x), fptr(out), B, C, T, F, windows, hop, span, stream_ptr()))
    return out


def conv(x, weight, bias, stride, padding):
    return ConvFn.apply(x, weight, bias, stride, padding)


def conv_transpose(x, weight, bias, stride, padding, output_padding):
    return ConvTransposeFn.apply(x, weight, bias, stride, padding, output_padding)


def bn_act(x, bn, act=ACT_NONE, slope=0.0):
    """bn: an nn.BatchNorm{2,3}d-like module (weight, bias, running_mean, running_var, num_batches_tracked, momentum, eps)."""
    return BNActFn.apply(x, bn.weight, bn.bias, bn.running_mean, bn.running_var, bn.num_batches_tracked, bn.training,
                         bn.momentum, bn.eps, act, slope)


def maxpool(x, ph, pw):
    return MaxPoolFn.apply(x, ph, pw)


def linear_act(x, weight, bias, act=ACT_NONE, slope=0.0):
    return LinearActFn.apply(x, weight, bias, act, slope)


def activation(x, act, slope=0.0):
    return ActFn.apply(x, act, slope)


def lstm(x, w_ih_f, w_hh_f, w_ih_r, w_hh_r):
    return LSTMFn.apply(x, w_ih_f, w_hh_f, w_ih_r, w_hh_r)

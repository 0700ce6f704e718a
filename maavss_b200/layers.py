"""nn.Module layers whose forward/backward run in libmaavss_b200.so.

Each class subclasses the torch module the reference uses (so parameter names, shapes, default initialisation and
state_dict keys are identical -- SURVEY 8b) and only replaces ``forward``.  ``FusedSequential`` keeps the
reference's flat ``nn.Sequential`` indexing (``stft_encoder.0.weight`` ...) but runs Conv -> BatchNorm -> activation
groups through the fused BN+activation kernels.
"""
import torch
import torch.nn as nn

from . import nn_ops as ops


def _need_cuda(x):
    if not x.is_cuda:
        raise RuntimeError("maavss_b200 layers need CUDA tensors on a B200 (there is no CPU fallback)")


class Conv2d(nn.Conv2d):
    def forward(self, x):
        _need_cuda(x)
        assert self.dilation == (1, 1) and self.groups == 1 and self.padding_mode == "zeros"
        return ops.conv(x, self.weight, self.bias, self.stride, self.padding)


class Conv3d(nn.Conv3d):
    def forward(self, x):
        _need_cuda(x)
        assert self.dilation == (1, 1, 1) and self.groups == 1 and self.padding_mode == "zeros"
        return ops.conv(x, self.weight, self.bias, self.stride, self.padding)


class ConvTranspose2d(nn.ConvTranspose2d):
    def forward(self, x, output_size=None):
        _need_cuda(x)
        assert output_size is None and self.dilation == (1, 1) and self.groups == 1
        return ops.conv_transpose(x, self.weight, self.bias, self.stride, self.padding, self.output_padding)


class _BNMixin:
    def fused_forward(self, x, act=ops.ACT_NONE, slope=0.0):
        _need_cuda(x)
        return ops.bn_act(x, self, act, slope)

    def forward(self, x):
        return self.fused_forward(x)


class BatchNorm2d(_BNMixin, nn.BatchNorm2d):
    pass


class BatchNorm3d(_BNMixin, nn.BatchNorm3d):
    pass


class MaxPool3d(nn.MaxPool3d):
    """Only the reference's use: kernel (1, ph, pw), stride = kernel, no padding / dilation (avse_model_final.py:36-57)."""

    def forward(self, x):
        _need_cuda(x)
        k = self.kernel_size
        if not (isinstance(k, tuple) and len(k) == 3 and k[0] == 1 and self.stride in (k, None) and self.padding in (0, (0, 0, 0))
                and self.dilation in (1, (1, 1, 1)) and not self.ceil_mode and not self.return_indices):
            raise NotImplementedError("maavss_b200.MaxPool3d supports kernel (1, ph, pw) with stride = kernel only")
        return ops.maxpool(x, k[1], k[2])


class Tanh(nn.Tanh):
    ACT = (ops.ACT_TANH, 0.0)

    def forward(self, x):
        return ops.activation(x, ops.ACT_TANH)


class Sigmoid(nn.Sigmoid):
    ACT = (ops.ACT_SIGMOID, 0.0)

    def forward(self, x):
        return ops.activation(x, ops.ACT_SIGMOID)


class LeakyReLU(nn.LeakyReLU):
    @property
    def ACT(self):
        return (ops.ACT_LEAKY, float(self.negative_slope))

    def forward(self, x):
        return ops.activation(x, ops.ACT_LEAKY, float(self.negative_slope))


class Linear(nn.Linear):
    def fused_forward(self, x, act=ops.ACT_NONE, slope=0.0):
        _need_cuda(x)
        return ops.linear_act(x, self.weight, self.bias, act, slope)

    def forward(self, x):
        return self.fused_forward(x)


class LSTM(nn.LSTM):
    """Only the configuration the reference uses: 1 layer, bias=False, batch_first, bidirectional."""

    def __init__(self, *a, **k):
        super().__init__(*a, **k)
        if not (self.num_layers == 1 and not self.bias and self.batch_first and self.bidirectional and self.proj_size == 0):
            raise NotImplementedError("maavss_b200.LSTM supports num_layers=1, bias=False, batch_first=True, bidirectional=True")

    def forward(self, x, hx=None):
        _need_cuda(x)
        assert hx is None
        y = ops.lstm(x, self.weight_ih_l0, self.weight_hh_l0, self.weight_ih_l0_reverse, self.weight_hh_l0_reverse)
        return y, None

    def flatten_parameters(self):      # cuDNN-specific in torch; a no-op here (and must not touch the flat arena views)
        return None

    def _apply(self, fn, recurse=True):   # skip nn.RNNBase's cuDNN weight re-flattening
        return nn.Module._apply(self, fn, recurse)


_ACTS = (Tanh, Sigmoid, LeakyReLU)


class FusedSequential(nn.Sequential):
    """nn.Sequential that runs (BatchNorm, activation) and (Linear, activation) pairs as one fused op."""

    def steps(self, x):
        """Generator over the fused ops: yields the running activation after each one (run_pair interleaves two of these)."""
        mods = list(self)
        i = 0
        while i < len(mods):
            m = mods[i]
            nxt = mods[i + 1] if i + 1 < len(mods) else None
            if isinstance(m, (_BNMixin, Linear)) and isinstance(nxt, _ACTS):
                act, slope = nxt.ACT
                x = m.fused_forward(x, act, slope)
                i += 2
            else:
                x = m(x)
                i += 1
            yield x

    def forward(self, x):
        for x in self.steps(x):
            pass
        return x


# The audio and the video encoder of the fusion models are independent until the LSTM (avse_model.py:680-690) and their
# layers are small (128-512 CTAs): run_pair() issues them on two streams, one layer of each in turn, so that the kernels
# of one branch fill the SMs the other leaves idle and (data parallel) the sync-BatchNorm collectives of one branch hide
# behind the other's convolutions.  Autograd replays each op's backward on its forward stream, so the backward overlaps
# the same way.  The fork / join are event waits: they are captured into the train step's CUDA graph like any launch.
PARALLEL_BRANCHES = True
_side_streams = {}


def _side_stream(device):
    s = _side_streams.get(device)
    if s is None:
        s = _side_streams[device] = torch.cuda.Stream(device=device)
    return s


_report_streams = {}


def _report_stream(device):
    """Stream of the loss values a train step only reports (train_step.FusionTrainStep): off the forward -> backward chain."""
    s = _report_streams.get(device)
    if s is None:
        s = _report_streams[device] = torch.cuda.Stream(device=device)
    return s


def run_pair(seq_main, x_main, seq_side, x_side):
    """(seq_main(x_main), seq_side(x_side)) with the two FusedSequentials interleaved on the current and a side stream."""
    if not (PARALLEL_BRANCHES and x_main.is_cuda and x_side.is_cuda):
        return seq_main(x_main), seq_side(x_side)
    cur = torch.cuda.current_stream(x_main.device)
    side = _side_stream(x_main.device)
    side.wait_stream(cur)
    ga, gb = seq_main.steps(x_main), seq_side.steps(x_side)
    a, b, done_a, done_b = x_main, x_side, False, False
    while not (done_a and done_b):
        if not done_b:
            with torch.cuda.stream(side):
                try:
                    b = next(gb)
                except StopIteration:
                    done_b = True
        if not done_a:
            try:
                a = next(ga)
            except StopIteration:
                done_a = True
    cur.wait_stream(side)
    b.record_stream(cur)
    return a, b

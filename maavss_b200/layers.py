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

    def forward(self, x):
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
        return x

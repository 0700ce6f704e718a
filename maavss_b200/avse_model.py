"""Drop-in for the reference `avse_model.AV_Fusion_Model` (avse_model.py:410-711): same constructor signature,
sub-module names, state_dict keys and methods; forward and backward run in libmaavss_b200.so.

The legacy generations in the same reference file (`AVSE_Model`, `AV_Model_STFT`, avse_model.py:11-405) are out of
scope."""
import torch
import torch.nn as nn

from . import _specs
from . import layers as L
from . import nn_ops as ops


def build_encoder(specs, conv_cls=L.Conv2d):
    mods = []
    for s in specs:
        mods += [conv_cls(s["cin"], s["cout"], kernel_size=s["k"], stride=s["s"], padding=s["p"], bias=s["bias"]),
                 L.BatchNorm2d(s["cout"]), L.Tanh()]
    return L.FusedSequential(*mods)


def build_decoder(specs):
    mods = []
    for s in specs:
        mods.append(L.ConvTranspose2d(s["cin"], s["cout"], kernel_size=s["k"], stride=s["s"], padding=s["p"],
                                      output_padding=s["op"], bias=s["bias"]))
        if not s["last"]:
            mods += [L.BatchNorm2d(s["cout"]), L.Tanh()]
    return L.FusedSequential(*mods)


def _default_device():
    # the reference hard-codes .to("cuda") in its constructor (avse_model.py:446 ff.)
    return torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")


class AV_Fusion_Model(nn.Module):
    VARIANT = "fc"

    def __init__(self, stft_shape, pgram_shape, alpha, latent_channels=64, fc_size=4096):
        super().__init__()
        self.stft_shape, self.pgram_shape, self.latent_channels = stft_shape, pgram_shape, latent_channels
        sp = _specs.fusion_specs(stft_shape, pgram_shape, latent_channels, fc_size, self.VARIANT)
        self.enc_hw = sp["enc_hw"]
        bias = sp["bias"]
        self.phasegram_encoder = build_encoder(sp["pg_enc"])
        self.phasegram_decoder = build_decoder(sp["pg_dec"])
        self.stft_encoder = build_encoder(sp["st_enc"])
        self.lstm = L.LSTM(input_size=sp["lstm_in"], hidden_size=256, num_layers=1, bias=False, batch_first=True,
                           dropout=0, bidirectional=True)
        self.fc1 = L.Linear(fc_size, fc_size // 2, bias=bias)
        self.fc2 = L.Linear(fc_size // 2, sp["fc2_out"], bias=bias)
        self.stft_decoder = build_decoder(sp["st_dec"])
        # the same module objects registered a second time: state_dict carries the aliases (avse_model.py:596-611)
        self.stft_autoencoder = L.FusedSequential(*self.stft_encoder, *self.stft_decoder)
        self.phasegram_autoencoder = L.FusedSequential(*self.phasegram_encoder, *self.phasegram_decoder)
        self._build_heads(sp["fc2_out"], bias)
        self.to(_default_device())

    def _build_heads(self, fc2_out, bias):
        s, p = self.stft_shape, self.pgram_shape
        self.a_fc1 = L.FusedSequential(L.Linear(fc2_out, 2 * s[-2] * s[-1]), L.LeakyReLU(negative_slope=0.3))
        self.v_fc1 = L.FusedSequential(L.Linear(fc2_out, p[-2] * p[-1]), L.LeakyReLU(negative_slope=0.3))

    # ---- avse_model.py:630-656 ----
    def toggle_fusion_grads(self, toggle):
        for m in (self.lstm, self.fc1, self.fc2, self.a_fc1, self.v_fc1):
            m.requires_grad_(toggle)

    def toggle_stft_ae_grads(self, toggle):
        for m in self.stft_autoencoder:
            m.requires_grad_(toggle)

    def toggle_phasegram_ae_grads(self, toggle):
        for m in self.phasegram_autoencoder:
            m.requires_grad_(toggle)

    def toggle_enc_grads(self, toggle):
        for m in list(self.stft_encoder) + list(self.phasegram_encoder):
            m.requires_grad_(toggle)

    def toggle_dec_grads(self, toggle):
        for m in list(self.stft_decoder) + list(self.phasegram_decoder):
            m.requires_grad_(toggle)

    # ---- avse_model.py:658-670 ----
    def av_fusion_forward(self, x_a, x_v):
        x = torch.cat((x_v.permute(0, 2, 1, 3), x_a.permute(0, 2, 1, 3)), dim=2).flatten(-2, -1)
        av = self.lstm(x)[0].flatten(1)
        av = self.fc1.fused_forward(av, ops.ACT_LEAKY, 0.3)
        return self.fc2.fused_forward(av, ops.ACT_LEAKY, 0.3)

    def visual_ae_forward(self, x_v):
        return self.phasegram_autoencoder(x_v)

    def audio_ae_forward(self, x_a):
        return self.stft_autoencoder(x_a)

    # ---- avse_model.py:680-711 ----
    def forward(self, x_a, x_v):
        x_a_enc, x_v_enc = L.run_pair(self.stft_encoder, x_a, self.phasegram_encoder, x_v)
        fused = self.av_fusion_forward(x_a_enc, x_v_enc)
        return self.a_fc1(fused).view(x_a.shape), self.v_fc1(fused).view(x_v.shape), fused

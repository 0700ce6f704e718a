"""Drop-in for the reference `avse_model_final.AV_Fusion_Model_Frames` (avse_model_final.py:14-274): same constructor
signature, sub-module names (state_dict keys) and methods; forward and backward run in libmaavss_b200.so.

Only constructible with latent_channels == 16, like the reference (cat of x_v and x_a along dim 2 needs equal channel
counts, avse_model_final.py:124)."""
import torch
import torch.nn as nn

from . import _specs
from . import layers as L
from . import nn_ops as ops
from .avse_model import _default_device, build_decoder, build_encoder


class AV_Fusion_Model_Frames(nn.Module):

    def __init__(self, stft_shape, frame_shape, hops_per_frame, latent_channels=16, fc_size=4096):
        super().__init__()
        self.stft_shape, self.frame_shape = stft_shape, frame_shape
        self.frame_channels = frame_shape[1]
        self.latent_channels = latent_channels
        self.output_stft_frames = hops_per_frame
        sp = _specs.frames_specs(stft_shape, frame_shape, latent_channels)
        if sp["vis"][-1]["cout"] != sp["st_enc"][-1]["cout"]:
            # the reference fails in torch.cat at construction for this case (avse_model_final.py:124)
            raise RuntimeError(f"Sizes of tensors must match except in dimension 2: visual encoder ends with "
                               f"{sp['vis'][-1]['cout']} channels, stft encoder with {sp['st_enc'][-1]['cout']}")
        mods = []
        for v in sp["vis"]:                                       # avse_model_final.py:33-57
            mods += [L.Conv3d(v["cin"], v["cout"], kernel_size=v["k"], stride=1, padding=v["p"], bias=False),
                     L.BatchNorm3d(v["cout"]), L.MaxPool3d((1, v["pool"], v["pool"])), L.LeakyReLU()]
        mods.append(nn.Flatten(start_dim=-2, end_dim=-1))
        self.visual_encoder = L.FusedSequential(*mods)
        self.stft_encoder = build_encoder(sp["st_enc"])
        tgt = sp["tgt"]
        self.lstm = L.LSTM(input_size=2 * tgt[0] * tgt[1], hidden_size=256, num_layers=1, bias=False, batch_first=True,
                           dropout=0, bidirectional=True)
        fc_in = 512 * latent_channels                             # fc_size is overwritten by the real size (:140)
        self.fc1 = L.Linear(fc_in, fc_in // 2, bias=False)
        self.fc2 = L.Linear(fc_in // 2, 512, bias=False)
        self.stft_decoder = build_decoder(sp["st_dec"])
        self.stft_autoencoder = L.FusedSequential(*self.stft_encoder, *self.stft_decoder)
        self.a_fc1 = L.FusedSequential(L.Linear(512, 2 * hops_per_frame * stft_shape[-1], bias=False))
        self.v_fc1 = L.FusedSequential(L.Linear(512, self.frame_channels * frame_shape[-2] * frame_shape[-1], bias=False))
        self.to(_default_device())

    # ---- avse_model_final.py:216-232 ----
    def toggle_fusion_grads(self, toggle):
        for m in (self.lstm, self.fc1, self.fc2, self.a_fc1, self.v_fc1):
            m.requires_grad_(toggle)

    def toggle_stft_ae_grads(self, toggle):
        for m in self.stft_autoencoder:
            m.requires_grad_(toggle)

    def toggle_enc_grads(self, toggle):
        for m in list(self.stft_encoder) + list(self.visual_encoder):
            m.requires_grad_(toggle)

    # ---- avse_model_final.py:235-251: the LSTM runs over the CHANNEL axis (no permute) ----
    def av_fusion_forward(self, x_a, x_v):
        x = torch.cat((x_v, x_a), dim=2).flatten(-2, -1)
        av = self.lstm(x)[0].flatten(1)
        av = self.fc1.fused_forward(av, ops.ACT_TANH)
        return self.fc2.fused_forward(av, ops.ACT_TANH)

    def audio_ae_forward(self, x_a):
        return self.stft_autoencoder(x_a)

    # ---- avse_model_final.py:258-274 ----
    def forward(self, x_a, x_v):
        x_a_enc, x_v_enc = L.run_pair(self.stft_encoder, x_a, self.visual_encoder, x_v)
        fused = self.av_fusion_forward(x_a_enc, x_v_enc)
        a = self.a_fc1[0].fused_forward(fused, ops.ACT_TANH)
        v = self.v_fc1[0].fused_forward(fused, ops.ACT_SIGMOID)
        v = v.view(x_v.shape[0], self.frame_channels, self.frame_shape[-2], self.frame_shape[-1])
        a = a.view(x_a.shape[0], 2, self.output_stft_frames, x_a.shape[-1])
        return a, v, fused

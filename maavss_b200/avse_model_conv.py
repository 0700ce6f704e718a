"""Drop-in for the reference `avse_model_conv.AV_Fusion_Model` (avse_model_conv.py:410-719): the bias-free variant
with (3, 9) STFT kernels whose forward runs the ConvTranspose2d decoders instead of the FC heads."""
from . import layers as L
from .avse_model import AV_Fusion_Model as _Base


class AV_Fusion_Model(_Base):
    VARIANT = "conv"

    def _build_heads(self, fc2_out, bias):
        s, p = self.stft_shape, self.pgram_shape
        self.a_fc1 = L.FusedSequential(L.Linear(fc2_out, 2 * s[-2] * s[-1], bias=False))
        self.v_fc1 = L.FusedSequential(L.Linear(fc2_out, p[-2] * p[-1], bias=False))

    # ---- avse_model_conv.py:700-719 ----
    def forward(self, x_a, x_v):
        x_a_enc, x_v_enc = L.run_pair(self.stft_encoder, x_a, self.phasegram_encoder, x_v)
        fused = self.av_fusion_forward(x_a_enc, x_v_enc)
        x_a_out = self.stft_decoder(fused.view(x_a_enc.shape))
        x_v_out = self.phasegram_decoder(fused.view(x_v_enc.shape))
        return x_a_out, x_v_out, fused

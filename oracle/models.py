"""Oracle (test infrastructure): the three MAAVSS fusion models restated on torch-CPU.

Follows /root/reference:
  avse_model.py:410-711        AV_Fusion_Model            (config 1, 2)
  avse_model_conv.py:410-719   AV_Fusion_Model (conv/bias-free variant; forward runs the deconv decoders)
  avse_model_final.py:14-274   AV_Fusion_Model_Frames     (config 4)
  train.py:136-162 / train_avse_frames.py:150-181  micro-step, loss, Adam

The reference builds its layers by pushing dummy tensors through them on "cuda"; here the
layer lists are derived by closed-form shape arithmetic (same loops, no tensors), giving
the same sub-module names and state_dict keys (including the ``stft_autoencoder`` /
``phasegram_autoencoder`` aliases that re-register the encoder/decoder modules,
avse_model.py:596-611).  Checked against the reference itself by tests/golden/make_golden.py.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F


def _conv_out(n, k, s, p):
    return (n + 2 * p - k) // s + 1


def _seq_enc(specs, act=nn.Tanh):
    mods = []
    for (cin, cout, k, s, p, bias) in specs:
        mods += [nn.Conv2d(cin, cout, k, s, p, bias=bias), nn.BatchNorm2d(cout), act()]
    return nn.Sequential(*mods)


def _seq_dec(specs):
    mods = []
    for i, (cin, cout, k, s, p, op, bias) in enumerate(specs):
        mods.append(nn.ConvTranspose2d(cin, cout, k, s, p, output_padding=op, bias=bias))
        if i != len(specs) - 1:
            mods += [nn.BatchNorm2d(cout), nn.Tanh()]
    return nn.Sequential(*mods)


def fusion_specs(stft_shape, pgram_shape, latent, fc_size, variant):
    """variant 'fc' = avse_model.py, 'conv' = avse_model_conv.py."""
    bias = variant == "fc"
    ka, pa = ((5, 5), (2, 2)) if variant == "fc" else ((3, 9), (1, 4))
    H, W = pgram_shape[-2], pgram_shape[-1]
    pg_enc, cin = [], 1
    while W * H * latent > fc_size // 2:                      # avse_model.py:432
        cout = min(cin * 2, latent)
        pg_enc.append((cin, cout, (1, 9), (1, 2), (0, 4), bias))
        W = _conv_out(W, 9, 2, 4)
        cin = cout
    pg_dec, cin, w = [], latent, W
    while w < pgram_shape[-1]:                                 # avse_model.py:452
        cout = max(cin // 2, 1)
        pg_dec.append((cin, cout, (1, 9), (1, 2), (0, 4), (0, 1), bias))
        w = (w - 1) * 2 - 8 + 9 + 1
        cin = cout
    tgt = [H, W]
    T, S = stft_shape[-2], stft_shape[-1]
    out_shape, st_enc, cin = [T, S], [], 2
    while out_shape != tgt:                                    # avse_model.py:480
        cout = min(cin * 4, latent)
        s = [1, 1]
        for d in (0, 1):
            if out_shape[d] > tgt[d]:
                s[d] = 2
                out_shape[d] //= 2
        st_enc.append((cin, cout, ka, tuple(s), pa, bias))
        cin = cout
    st_dec, cin, enc = [], latent, list(tgt)
    while enc != [T, S]:                                       # avse_model.py:569
        cout = max(cin // 4, stft_shape[1])
        s, op = [1, 1], [0, 0]
        for d, full in ((0, T), (1, S)):
            if enc[d] < full:
                s[d], op[d] = 2, 1
                enc[d] *= 2
        st_dec.append((cin, cout, ka, tuple(s), pa, tuple(op), bias))
        cin = cout
    lstm_in = latent * tgt[1] * 2
    fc2_out = 512 if variant == "fc" else latent * tgt[0] * tgt[1]
    return dict(pg_enc=pg_enc, pg_dec=pg_dec, st_enc=st_enc, st_dec=st_dec, lstm_in=lstm_in,
                fc2_out=fc2_out, enc_hw=tuple(tgt), bias=bias)


class AVFusionOracle(nn.Module):
    def __init__(self, stft_shape, pgram_shape, alpha=8, latent_channels=64, fc_size=4096,
                 variant="fc"):
        super().__init__()
        sp = fusion_specs(stft_shape, pgram_shape, latent_channels, fc_size, variant)
        self.variant, self.latent = variant, latent_channels
        self.phasegram_encoder = _seq_enc(sp["pg_enc"])
        self.phasegram_decoder = _seq_dec(sp["pg_dec"])
        self.stft_encoder = _seq_enc(sp["st_enc"])
        self.lstm = nn.LSTM(sp["lstm_in"], 256, 1, bias=False, batch_first=True, bidirectional=True)
        self.fc1 = nn.Linear(fc_size, fc_size // 2, bias=sp["bias"])
        self.fc2 = nn.Linear(fc_size // 2, sp["fc2_out"], bias=sp["bias"])
        self.stft_decoder = _seq_dec(sp["st_dec"])
        self.stft_autoencoder = nn.Sequential(*self.stft_encoder, *self.stft_decoder)
        self.phasegram_autoencoder = nn.Sequential(*self.phasegram_encoder, *self.phasegram_decoder)
        if variant == "fc":
            self.a_fc1 = nn.Sequential(nn.Linear(sp["fc2_out"], 2 * stft_shape[-2] * stft_shape[-1]),
                                       nn.LeakyReLU(0.3))
            self.v_fc1 = nn.Sequential(nn.Linear(sp["fc2_out"], pgram_shape[-2] * pgram_shape[-1]),
                                       nn.LeakyReLU(0.3))
        else:
            self.a_fc1 = nn.Sequential(nn.Linear(sp["fc2_out"], 2 * stft_shape[-2] * stft_shape[-1], bias=False))
            self.v_fc1 = nn.Sequential(nn.Linear(sp["fc2_out"], pgram_shape[-2] * pgram_shape[-1], bias=False))

    def av_fusion_forward(self, x_a, x_v):                    # avse_model.py:658-670
        x = torch.cat((x_v.permute(0, 2, 1, 3), x_a.permute(0, 2, 1, 3)), dim=2).flatten(-2, -1)
        av = self.lstm(x)[0].flatten(1)
        av = F.leaky_relu(self.fc1(av), 0.3)
        return F.leaky_relu(self.fc2(av), 0.3)

    def audio_ae_forward(self, x_a):
        return self.stft_autoencoder(x_a)

    def visual_ae_forward(self, x_v):
        return self.phasegram_autoencoder(x_v)

    def forward(self, x_a, x_v):                               # avse_model.py:680-711
        a_enc, v_enc = self.stft_encoder(x_a), self.phasegram_encoder(x_v)
        fused = self.av_fusion_forward(a_enc, v_enc)
        if self.variant == "fc":
            return self.a_fc1(fused).view(x_a.shape), self.v_fc1(fused).view(x_v.shape), fused
        # avse_model_conv.py:711-717
        return (self.stft_decoder(fused.view(a_enc.shape)),
                self.phasegram_decoder(fused.view(v_enc.shape)), fused)


class AVFusionFramesOracle(nn.Module):
    """avse_model_final.AV_Fusion_Model_Frames."""

    def __init__(self, stft_shape, frame_shape, hops_per_frame, latent_channels=16, fc_size=4096):
        super().__init__()
        L = latent_channels
        self.frame_shape, self.frame_channels, self.hpf = frame_shape, frame_shape[1], hops_per_frame
        chans = [(1, 16), (16, 32), (32, 64), (64, 64), (64, L)]
        pools = [2, 2, 2, 3, 3]
        mods, h, w = [], frame_shape[-2], frame_shape[-1]
        for i, ((ci, co), pk) in enumerate(zip(chans, pools)):
            pad = (1, 3, 3) if i == 4 else (1, 2, 2)           # avse_model_final.py:54
            mods += [nn.Conv3d(ci, co, (3, 5, 5), 1, pad, bias=False), nn.BatchNorm3d(co),
                     nn.MaxPool3d((1, pk, pk)), nn.LeakyReLU()]
            h, w = (h + 2 * pad[1] - 4) // pk, (w + 2 * pad[2] - 4) // pk
        mods.append(nn.Flatten(-2, -1))
        self.visual_encoder = nn.Sequential(*mods)
        tgt = [frame_shape[2], h * w]
        T, S = stft_shape[-2], stft_shape[-1]
        out_shape, specs, cin, first = [T, S], [], 2, True
        while out_shape != tgt:                                # avse_model_final.py:81
            cout = min(cin * 2, L)
            s = [1, 1]
            for d in (0, 1):
                if out_shape[d] > tgt[d]:
                    s[d] = 2
                    out_shape[d] //= 2
            specs.append((cin, cout, (3, 9), tuple(s), (1, 3 if first else 4), False))
            first, cin = False, cout
        self.stft_encoder = _seq_enc(specs)
        # cat((x_v, x_a), dim=2): (B, L, 2*T', HW) -> flatten -> LSTM input 2*T'*HW
        self.lstm = nn.LSTM(2 * tgt[0] * tgt[1], 256, 1, bias=False, batch_first=True,
                            bidirectional=True)
        fc_in = 512 * L
        self.fc1 = nn.Linear(fc_in, fc_in // 2, bias=False)
        self.fc2 = nn.Linear(fc_in // 2, 512, bias=False)
        dspecs, cin, enc, cur, k = [], L, list(tgt), list(tgt), (3, 9)
        while cur != [T, S]:                                   # avse_model_final.py:160
            cout = max(cin // 2, stft_shape[1])
            s, op = [1, 1], [0, 0]
            for d, full in ((0, T), (1, S)):
                if enc[d] < full:
                    s[d], op[d] = 2, 1
                    enc[d] *= 2
            dspecs.append((cin, cout, k, tuple(s), (1, 4), tuple(op), False))
            cur = [(cur[0] - 1) * s[0] - 2 + k[0] + op[0], (cur[1] - 1) * s[1] - 8 + k[1] + op[1]]
            k = (3, 10) if cur[1] == (S - 1) // 2 else (3, 9)
            cin = cout
        self.stft_decoder = _seq_dec(dspecs)
        self.stft_autoencoder = nn.Sequential(*self.stft_encoder, *self.stft_decoder)
        self.a_fc1 = nn.Sequential(nn.Linear(512, 2 * hops_per_frame * S, bias=False))
        self.v_fc1 = nn.Sequential(nn.Linear(512, self.frame_channels * frame_shape[-2] * frame_shape[-1], bias=False))

    def av_fusion_forward(self, x_a, x_v):                    # avse_model_final.py:235-251
        x = torch.cat((x_v, x_a), dim=2).flatten(-2, -1)       # LSTM runs over the CHANNEL axis
        av = self.lstm(x)[0].flatten(1)
        return torch.tanh(self.fc2(torch.tanh(self.fc1(av))))

    def audio_ae_forward(self, x_a):
        return self.stft_autoencoder(x_a)

    def forward(self, x_a, x_v):                               # avse_model_final.py:258-274
        fused = self.av_fusion_forward(self.stft_encoder(x_a), self.visual_encoder(x_v))
        a = torch.tanh(self.a_fc1(fused)).view(x_a.shape[0], 2, self.hpf, x_a.shape[-1])
        v = torch.sigmoid(self.v_fc1(fused)).view(x_v.shape[0], self.frame_channels,
                                                  self.frame_shape[-2], self.frame_shape[-1])
        return a, v, fused


def micro_step_loss(model, x_a, x_v, y_a, y_v, loss_coeff=0.001, num_seq=4):
    """train.py:152-158: MSE(a) + loss_coeff*MSE(v), divided by num_seq."""
    yh_a, yh_v, _ = model(x_a, x_v)
    a_loss = F.mse_loss(yh_a, y_a)
    v_loss = F.mse_loss(yh_v, y_v)
    return (a_loss + loss_coeff * v_loss) / num_seq, a_loss, v_loss

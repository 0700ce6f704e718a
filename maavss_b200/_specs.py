"""Closed-form layer geometry of the three fusion models.

The reference discovers its layer lists by pushing dummy tensors through freshly built layers on "cuda"
(avse_model.py:427-592, avse_model_conv.py:427-594, avse_model_final.py:33-193); the loops only depend on
shapes, so they are restated here as integer arithmetic (no tensors, no device)."""


def conv_out(n, k, s, p):
    return (n + 2 * p - k) // s + 1


def fusion_specs(stft_shape, pgram_shape, latent, fc_size, variant):
    """variant 'fc' = avse_model.py (5x5 kernels, biases, FC heads); 'conv' = avse_model_conv.py ((3,9), bias-free)."""
    bias = variant == "fc"
    ka, pa = ((5, 5), (2, 2)) if variant == "fc" else ((3, 9), (1, 4))
    H, W = pgram_shape[-2], pgram_shape[-1]
    pg_enc, cin = [], 1
    while W * H * latent > fc_size // 2:                       # avse_model.py:432
        cout = min(cin * 2, latent)
        pg_enc.append(dict(cin=cin, cout=cout, k=(1, 9), s=(1, 2), p=(0, 4), bias=bias))
        W = conv_out(W, 9, 2, 4)
        cin = cout
    pg_dec, cin, w = [], latent, W
    while w < pgram_shape[-1]:                                  # avse_model.py:452
        cout = max(cin // 2, 1)
        w = (w - 1) * 2 - 8 + 9 + 1
        pg_dec.append(dict(cin=cin, cout=cout, k=(1, 9), s=(1, 2), p=(0, 4), op=(0, 1), bias=bias,
                           last=(w == pgram_shape[-1])))
        cin = cout
    tgt = [H, W]
    T, S = stft_shape[-2], stft_shape[-1]
    cur, st_enc, cin = [T, S], [], 2
    while cur != tgt:                                           # avse_model.py:480
        cout = min(cin * 4, latent)
        s = [1, 1]
        for d in (0, 1):
            if cur[d] > tgt[d]:
                s[d] = 2
                cur[d] //= 2
        st_enc.append(dict(cin=cin, cout=cout, k=ka, s=tuple(s), p=pa, bias=bias))
        cin = cout
        if len(st_enc) > 64:
            raise ValueError("stft_shape cannot be reduced to the phasegram encoder's output shape")
    st_dec, cin, enc = [], latent, list(tgt)
    while enc != [T, S]:                                        # avse_model.py:569
        cout = max(cin // 4, stft_shape[1])
        s, op = [1, 1], [0, 0]
        for d, full in ((0, T), (1, S)):
            if enc[d] < full:
                s[d], op[d] = 2, 1
                enc[d] *= 2
        st_dec.append(dict(cin=cin, cout=cout, k=ka, s=tuple(s), p=pa, op=tuple(op), bias=bias, last=(enc == [T, S])))
        cin = cout
    return dict(pg_enc=pg_enc, pg_dec=pg_dec, st_enc=st_enc, st_dec=st_dec, lstm_in=latent * tgt[1] * 2,
                fc2_out=512 if variant == "fc" else latent * tgt[0] * tgt[1], enc_hw=tuple(tgt), bias=bias)


def frames_specs(stft_shape, frame_shape, latent):
    """avse_model_final.AV_Fusion_Model_Frames."""
    chans = [(1, 16), (16, 32), (32, 64), (64, 64), (64, latent)]
    pools = [2, 2, 2, 3, 3]
    vis, h, w = [], frame_shape[-2], frame_shape[-1]
    for i, ((ci, co), pk) in enumerate(zip(chans, pools)):
        pad = (1, 3, 3) if i == 4 else (1, 2, 2)                # avse_model_final.py:54
        vis.append(dict(cin=ci, cout=co, k=(3, 5, 5), p=pad, pool=pk))
        h, w = (h + 2 * pad[1] - 4) // pk, (w + 2 * pad[2] - 4) // pk
    tgt = [frame_shape[2], h * w]
    T, S = stft_shape[-2], stft_shape[-1]
    cur, st_enc, cin, first = [T, S], [], 2, True
    while cur != tgt:                                           # avse_model_final.py:81
        cout = min(cin * 2, latent)
        s = [1, 1]
        for d in (0, 1):
            if cur[d] > tgt[d]:
                s[d] = 2
                cur[d] //= 2
        st_enc.append(dict(cin=cin, cout=cout, k=(3, 9), s=tuple(s), p=(1, 3 if first else 4), bias=False))
        first, cin = False, cout
        if len(st_enc) > 64:
            raise ValueError("stft_shape cannot be reduced to the visual encoder's output shape")
    st_dec, cin, enc, cur, k = [], latent, list(tgt), list(tgt), (3, 9)
    while cur != [T, S]:                                        # avse_model_final.py:160
        cout = max(cin // 2, stft_shape[1])
        s, op = [1, 1], [0, 0]
        for d, full in ((0, T), (1, S)):
            if enc[d] < full:
                s[d], op[d] = 2, 1
                enc[d] *= 2
        cur = [(cur[0] - 1) * s[0] - 2 + k[0] + op[0], (cur[1] - 1) * s[1] - 8 + k[1] + op[1]]
        st_dec.append(dict(cin=cin, cout=cout, k=k, s=tuple(s), p=(1, 4), op=tuple(op), bias=False, last=(cur == [T, S])))
        k = (3, 10) if cur[1] == (S - 1) // 2 else (3, 9)
        cin = cout
        if len(st_dec) > 64:
            raise ValueError("stft decoder does not reach stft_shape")
    return dict(vis=vis, st_enc=st_enc, st_dec=st_dec, tgt=tuple(tgt))

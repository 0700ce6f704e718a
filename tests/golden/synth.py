"""Seeded synthetic inputs / weights shared by make_golden.py (reference side, authoring
container only) and the tests (oracle + CUDA side).  Everything derives from
torch.Generator on CPU so both sides see identical tensors without shipping weights."""
import numpy as np
import torch


def gen(seed):
    g = torch.Generator()
    g.manual_seed(seed)
    return g


def fill_state_dict(sd, seed, scale=None):
    """Deterministically overwrite every floating tensor of a state_dict, walking keys in
    sorted order and de-duplicating aliases (stft_autoencoder.* shares storage with
    stft_encoder.* / stft_decoder.*).  BN running_var is kept positive."""
    g = gen(seed)
    seen = {}
    for k in sorted(sd.keys()):
        v = sd[k]
        if not v.dtype.is_floating_point:
            continue
        if v.data_ptr() in seen:
            continue
        seen[v.data_ptr()] = k
        r = torch.randn(v.shape, generator=g)
        if k.endswith("running_var"):
            v.copy_(0.5 + r.abs())
        elif k.endswith("running_mean"):
            v.copy_(0.1 * r)
        elif v.dim() == 1 and ("norm" in k or k.split(".")[-2].isdigit()) and k.endswith("weight"):
            v.copy_(1.0 + 0.1 * r)            # BN / LN gains
        elif v.dim() == 1:
            v.copy_(0.05 * r)                  # biases
        else:
            fan_in = v[0].numel() if v.dim() > 1 else v.numel()
            s = scale if scale is not None else (1.0 / max(fan_in, 1)) ** 0.5
            v.copy_(s * r)
    return sd


def vit_state_dict(depth=12, dim=384, patch=8, pos_grid=28, seed=7):
    """State dict with dino VisionTransformer key names (dino/vision_transformer.py:135-160)."""
    g = gen(seed)
    n = lambda *s: torch.randn(*s, generator=g)
    sd = {
        "cls_token": 0.02 * n(1, 1, dim),
        "pos_embed": 0.02 * n(1, pos_grid * pos_grid + 1, dim),
        "patch_embed.proj.weight": n(dim, 3, patch, patch) * (3 * patch * patch) ** -0.5,
        "patch_embed.proj.bias": 0.02 * n(dim),
        "norm.weight": torch.ones(dim), "norm.bias": torch.zeros(dim),
    }
    for i in range(depth):
        p = f"blocks.{i}."
        sd[p + "norm1.weight"] = 1.0 + 0.1 * n(dim)
        sd[p + "norm1.bias"] = 0.05 * n(dim)
        sd[p + "attn.qkv.weight"] = n(3 * dim, dim) * (2.0 / dim) ** 0.5   # peaky attention
        sd[p + "attn.qkv.bias"] = 0.02 * n(3 * dim)
        sd[p + "attn.proj.weight"] = n(dim, dim) * dim ** -0.5 * 0.5
        sd[p + "attn.proj.bias"] = 0.02 * n(dim)
        sd[p + "norm2.weight"] = 1.0 + 0.1 * n(dim)
        sd[p + "norm2.bias"] = 0.05 * n(dim)
        sd[p + "mlp.fc1.weight"] = n(4 * dim, dim) * dim ** -0.5
        sd[p + "mlp.fc1.bias"] = 0.02 * n(4 * dim)
        sd[p + "mlp.fc2.weight"] = n(dim, 4 * dim) * (4 * dim) ** -0.5 * 0.5
        sd[p + "mlp.fc2.bias"] = 0.02 * n(dim)
    return sd


def audio_clip(n, seed=11, scale=0.1):
    return torch.randn(n, generator=gen(seed)) * scale


def frames(T, H, W, seed=13):
    return torch.randn(T, 3, H, W, generator=gen(seed))


def attn_clip(B, T, H, W, seed=17):
    return torch.rand(B, 1, T, H, W, generator=gen(seed))


def stft_batch(B, T, Fq, seed=19, scale=0.5):
    g = gen(seed)
    y = torch.randn(B, 2, T, Fq, generator=g) * scale
    x = y + 0.1 * torch.randn(B, 2, T, Fq, generator=g)
    return x, y


def sample_idx(n):
    return (torch.arange(8, dtype=torch.int64) * (n - 1)) // 7


def grad_digest(model):
    """Per-parameter (l2 norm, strided samples) of .grad -- small but sensitive."""
    norms, samples, names = [], [], []
    seen = set()
    for k, p in model.named_parameters():
        if p.data_ptr() in seen:
            continue
        seen.add(p.data_ptr())
        names.append(k)
        if p.grad is None:
            norms.append(-1.0)
            samples.append(np.zeros(8, np.float32))
            continue
        g = p.grad.flatten()
        norms.append(float(g.double().norm()))
        idx = sample_idx(g.numel())
        samples.append(g[idx].numpy().astype(np.float32))
    return names, np.array(norms), np.stack(samples)


def param_digest(model):
    out, seen = [], set()
    for k, p in model.named_parameters():
        if p.data_ptr() in seen:
            continue
        seen.add(p.data_ptr())
        f = p.detach().flatten()
        idx = sample_idx(f.numel())
        out.append(np.concatenate(([float(f.double().sum())], f[idx].double().numpy())))
    return np.stack(out)




def flow_frames(seed, H=256, W=256, shift=(1.7, -2.3)):
    """A textured gray frame with a few rectangles (strong corners) and the same scene translated by `shift` = (dx, dy) pixels
    (bilinear resampling) plus a slow brightness ramp: uint8 (H, W) x 2, as numpy arrays."""
    g = gen(seed)
    base = torch.rand(1, 1, H // 8 + 4, W // 8 + 4, generator=g)
    big = torch.nn.functional.interpolate(base, size=(H + 32, W + 32), mode="bicubic", align_corners=False)[0, 0]
    big = (big - big.min()) / (big.max() - big.min()) * 200.0 + 20.0
    rs = np.random.RandomState(seed)
    for _ in range(14):
        y, x = rs.randint(20, H - 20), rs.randint(20, W - 20)
        h, w = rs.randint(8, 30), rs.randint(8, 30)
        big[y:y + h, x:x + w] = float(rs.randint(0, 256))
    yy, xx = torch.meshgrid(torch.arange(H, dtype=torch.float32), torch.arange(W, dtype=torch.float32), indexing="ij")

    def sample(dx, dy):
        gx = (xx + 16 - dx) / (W + 32 - 1) * 2 - 1
        gy = (yy + 16 - dy) / (H + 32 - 1) * 2 - 1
        grid = torch.stack((gx, gy), -1)[None]
        return torch.nn.functional.grid_sample(big[None, None], grid, mode="bilinear", align_corners=True)[0, 0]
    a = sample(0.0, 0.0).round().clamp(0, 255).to(torch.uint8).numpy()
    b = (sample(*shift) * 1.02 + 1.0).round().clamp(0, 255).to(torch.uint8).numpy()
    return a, b

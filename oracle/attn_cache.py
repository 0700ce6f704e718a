"""Oracle (test infrastructure): the attention-frame cache arithmetic on torch-CPU.

Reference call sites: the writer `torchvision.utils.save_image(frame, path)` (av_dataset.py:272-276,
save_attn_videos.py:10-19) and the reader `imgLoadTransforms(Image.open(path))` = ToTensor (av_dataset.py:262-266),
followed by the temporal diff / clip normalise of the un-cached path (av_dataset.py:323-328).  The quantiser and the
de-quantiser live in torchvision (not vendored; the reference pins no version, requirements.txt:1-5): restated here
from the installed torchvision 0.26 --
  save_image:  ndarr = grid.mul(255).add_(0.5).clamp_(0, 255).permute(1, 2, 0).to("cpu", torch.uint8)
  ToTensor:    img.to(float32).div(255)
The JPEG coding between the two (lossy, 3 identical channels) is what the raw cache removes and is NOT restated; nor is
`torch.diff(attn, )` of av_dataset.py:268-269, which differences along W and changes the shape (reference
bug) -- the cached reader applies the same diff as the un-cached path.  Pinned by tests/test_oracle_golden.py against
torchvision's own save_image -> PNG (lossless) -> ToTensor round trip run in this container (tests/golden/attn_cache.npz).
"""
import torch


def pack_u8(maps):
    return maps.clone().mul(255).add_(0.5).clamp_(0, 255).to(torch.uint8)


def pack_f16(maps):
    return maps.to(torch.float16)


def unpack(stored):
    if stored.dtype == torch.uint8:
        return stored.to(torch.float32).div(255)
    return stored.to(torch.float32)


def load_clips(stored, attn_diff):
    """stored (clips, T, H, W) uint8 / float16 -> (clips, T, H, W) fp32, av_dataset.py:323-328 per clip."""
    out = []
    for clip in unpack(stored):
        attn = clip
        if attn_diff:
            pad = torch.zeros_like(attn[0:1])
            attn = torch.cat((pad, torch.diff(attn, dim=0)), 0)
        attn = attn * (1 / torch.max(attn))
        out.append(attn)
    return torch.stack(out)

"""Drop-in for the reference `video_attention.VideoAttention` (video_attention.py:24-160): DINO ViT last-layer
CLS->patch self-attention maps, summed over heads, nearest-upsampled by the patch size and max-normalised per frame.

The reference runs one frame at a time at batch 1 with a host round trip per frame (video_attention.py:41-96); here
all frames of a call go through the tcgen05 ViT of `vit_ops.ViTPlan` as one batch and only the finished maps are
copied back.  `_inference` keeps the reference contract (float32 CPU tensor (T, 1, H, W)); `inference_device` is the
device-resident fast path used by the dataset / clip pipeline."""
import math
import os

import torch

from . import attn_ops, vit_ops

_ARCH = {"vit_tiny": (192, 12), "vit_small": (384, 12), "vit_base": (768, 12)}   # dino/vision_transformer.py:229-254


def random_state_dict(architecture="vit_small", patch_size=8, img_size=224, seed=None):
    """Weights as dino's VisionTransformer initialises them (trunc_normal std .02, LayerNorm 1/0, dino/vision_transformer.py:162-172)."""
    dim, depth = _ARCH[architecture]
    g = torch.Generator().manual_seed(seed) if seed is not None else None
    tn = lambda *s: torch.nn.init.trunc_normal_(torch.empty(*s), std=0.02, generator=g)
    n0 = (img_size // patch_size) ** 2
    fan_in = 3 * patch_size * patch_size
    bound = 1.0 / math.sqrt(fan_in)
    sd = {"cls_token": tn(1, 1, dim), "pos_embed": tn(1, n0 + 1, dim),
          "patch_embed.proj.weight": (torch.rand(dim, 3, patch_size, patch_size, generator=g) * 2 - 1) * bound,
          "patch_embed.proj.bias": (torch.rand(dim, generator=g) * 2 - 1) * bound,
          "norm.weight": torch.ones(dim), "norm.bias": torch.zeros(dim)}
    for i in range(depth):
        p = f"blocks.{i}."
        sd.update({p + "norm1.weight": torch.ones(dim), p + "norm1.bias": torch.zeros(dim),
                   p + "attn.qkv.weight": tn(3 * dim, dim), p + "attn.qkv.bias": torch.zeros(3 * dim),
                   p + "attn.proj.weight": tn(dim, dim), p + "attn.proj.bias": torch.zeros(dim),
                   p + "norm2.weight": torch.ones(dim), p + "norm2.bias": torch.zeros(dim),
                   p + "mlp.fc1.weight": tn(4 * dim, dim), p + "mlp.fc1.bias": torch.zeros(4 * dim),
                   p + "mlp.fc2.weight": tn(dim, 4 * dim), p + "mlp.fc2.bias": torch.zeros(dim)})
    return sd


class VideoAttention:
    def __init__(self, patch_size=8, threshold=0.6, path_to_weights="dino_deitsmall8_pretrain.pth", architecture="vit_small",
                 resize=None, state_dict=None, device=None):
        self.resize = resize
        self.threshold = threshold          # has no effect on the output in the reference either (video_attention.py:59-78 is dead code)
        self.patch_size = patch_size
        self.checkpoint_key = "teacher"
        if not torch.cuda.is_available():
            raise RuntimeError("maavss_b200.VideoAttention needs a CUDA device (there is no CPU path)")
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.model = self.__load_model(architecture, path_to_weights, state_dict)

    def __load_model(self, arch, pretrained_weights, state_dict):
        if state_dict is None and pretrained_weights and os.path.isfile(pretrained_weights):
            state_dict = torch.load(pretrained_weights, map_location="cpu")
            if self.checkpoint_key is not None and self.checkpoint_key in state_dict:
                state_dict = state_dict[self.checkpoint_key]                      # video_attention.py:118-126
            state_dict = {k.replace("module.", "").replace("backbone.", ""): v for k, v in state_dict.items()}
        if state_dict is None:
            # the reference downloads the DINO weights here (video_attention.py:139-155); there is no network on the box
            print("No pretrained weights found => random weights (dino initialisation).")
            state_dict = random_state_dict(arch, self.patch_size)
        plan = vit_ops.ViTPlan(state_dict, self.device)
        if plan.patch != self.patch_size:
            raise ValueError(f"weights have patch size {plan.patch}, VideoAttention was built with {self.patch_size}")
        return plan

    @torch.no_grad()
    def inference_device(self, frames):
        """frames (T, 3, H, W) float -> (T, 1, H', W') on the device (H', W' cropped to multiples of the patch size)."""
        return self.model.attention_maps(frames.to(self.device, torch.float32, non_blocking=True))

    @torch.no_grad()
    def _inference(self, frames):
        return self.inference_device(frames).cpu()

    @torch.no_grad()
    def clip_attention(self, frames, attn_diff=False, clips=1):
        """The dataset's per-clip post-processing fused on the device (av_dataset.py:320-328): attention maps, optional
        temporal frame difference with a zero first frame, division by the clip's (signed) maximum."""
        f = frames.to(self.device, torch.float32, non_blocking=True)
        return self.model.attention_maps(f, post=2 if attn_diff else 1, clips=clips)

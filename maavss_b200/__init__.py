"""maavss_b200 -- B200-native (sm_100a) implementation of the MAAVSS data-parallel training hot path.

The arithmetic lives in libmaavss_b200.so (hand-written CUDA behind the C ABI of
include/maavss_b200.h); this package is the host-side mirror of the reference's Python interface
for that path (module / class / method names of av_dataset, generator, video_attention,
avse_model, avse_model_final, avse_model_conv, utilities, run_config).
There is no CPU or PyTorch fallback: the ops raise if the library or a CUDA device is missing.
"""
from . import _lib  # noqa: F401

__all__ = ["_lib"]

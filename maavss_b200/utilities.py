"""Host-side mirror of the hot-path helpers of the reference `utilities.py`.

Only the functions on the training hot path are provided: `calc_hop_size`
(utilities.py:24-28), checkpoint save/load with the reference's dict layout (utilities.py:165-204)
and `video_phasegram` (utilities.py:206-228).  File/ffmpeg/matplotlib helpers are out of scope.
"""
import glob
import os

import torch


def calc_hop_size(num_frames, hops_per_frame, fps, sr):
    """utilities.py:24-28 -- integer arithmetic must be bit-exact (float division, int() truncation)."""
    hop = int((sr / fps) / hops_per_frame)
    audio_sample_len = int(hops_per_frame * hop * num_frames)
    num_fft_frames = audio_sample_len // hop
    return hop, audio_sample_len, num_fft_frames


def save_model(path, model, overwrite=False):
    if not overwrite and os.path.exists(path):
        raise FileExistsError(path)
    torch.save(model.state_dict(), path)


def save_checkpoint(model_state, opt_state, epoch, loss, cp_dir, name="checkpoint"):
    """Same dict keys as utilities.py:171-178."""
    os.makedirs(cp_dir, exist_ok=True)
    path = os.path.join(cp_dir, f"{name}.pt")
    torch.save({"epoch": epoch, "model_state_dict": model_state, "optimizer_state_dict": opt_state, "loss": loss}, path)
    return path


def latest_file(directory, pattern="*.pt"):
    files = glob.glob(os.path.join(directory, pattern))
    return max(files, key=os.path.getctime) if files else None


def load_checkpoint(model, optimizer=None, cp_dir="checkpoints/", auto=True, path=None, load_opt=False):
    """utilities.py:180-204: strict=False model load, optional optimizer state."""
    if path is None and auto:
        path = latest_file(cp_dir)
    if path is None:
        return None
    cp = torch.load(path, map_location="cpu")
    model.load_state_dict(cp["model_state_dict"], strict=False)
    if load_opt and optimizer is not None:
        optimizer.load_state_dict(cp["optimizer_state_dict"])
    return cp


def video_phasegram(frames, resize=None, diff=True, cumulative=True, normalize=True):
    """utilities.py:206-228 on the device (fused CUDA kernels, see phasegram_ops)."""
    from . import phasegram_ops
    return phasegram_ops.video_phasegram(frames, resize=resize, diff=diff, cumulative=cumulative, normalize=normalize)

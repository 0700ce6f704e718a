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
    """utilities.py:162-166: always overwrites (the reference's `overwrite` handling is commented out, and
    train_autoencoder.py:131 saves to the same path every epoch)."""
    torch.save(model.state_dict(), path)


def save_checkpoint(model_dict, opt_dict, epoch, loss, name, dir):
    """utilities.py:168-175, same positional order (train_audio_net.py:130 calls it positionally) and dict keys."""
    os.makedirs(dir, exist_ok=True)
    path = f"{dir}/{name}.pt"
    torch.save({"epoch": epoch, "model_state_dict": model_dict, "optimizer_state_dict": opt_dict, "loss": loss}, path)
    return path


def latest_file(dir, ext="pt"):
    """utilities.py:199-204."""
    files = glob.glob(f"{dir}/*.{ext}", recursive=True)
    return max(files, key=os.path.getctime) if files else None


def load_checkpoint(model, optimizer, dir, auto=True, path=None, load_opt=False):
    """utilities.py:177-197: the newest *.pt of `dir` (auto) or `path`; strict=False model load; the optimizer state is
    loaded only on request and a failure there is reported, not raised."""
    if auto:
        path = latest_file(dir, "pt")
        if path is None:
            print("checkpoint not found, aborting cp load")
            return None
    elif path is None:
        return None
    cp = torch.load(path, map_location="cpu")
    model.load_state_dict(cp["model_state_dict"], strict=False)
    if load_opt:
        try:
            optimizer.load_state_dict(cp["optimizer_state_dict"])
        except Exception as e:
            print(f"Error loading optimizer: {e}")
    return cp


def video_phasegram(frames, resize=None, diff=True, cumulative=True, normalize=True):
    """utilities.py:206-228 on the device (fused CUDA kernels, see phasegram_ops)."""
    from . import phasegram_ops
    return phasegram_ops.video_phasegram(frames, resize=resize, diff=diff, cumulative=cumulative, normalize=normalize)

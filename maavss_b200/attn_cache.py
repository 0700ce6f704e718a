"""Raw attention-frame cache.

Replaces the reference's per-frame JPEG cache -- `AV_Dataset.get_frames_cached` (av_dataset.py:251-278), written by
`save_attn_videos.save_frames` (save_attn_videos.py:10-19) as `<attn_frames_path>/<video>/img_%05d.jpg` -- and finishes the
memmap idea of `preprocess_frames.py:11-24`: ONE memory-mapped file of raw uint8 (save_image's quantiser, no JPEG loss) or
fp16 frames, a validity byte per frame (the `utilities.verify_files` check) and a JSON index video name -> first frame slot.
Quantising, de-quantising and the diff / clip-normalise of av_dataset.py:323-328 run on the B200 through the C ABI
(`maavss_attn_cache_pack` / `maavss_attn_cache_load`); the host side only moves bytes between the memmap and pinned staging
buffers.  There is no CPU arithmetic path.

What is stored is the output of `VideoAttention._inference` (per-frame max-normalised, in [0, 1]) BEFORE the temporal
diff: overlapping clips of one video (frame_hop < num_frames) then share frames, as the reference's `img_{i+true_idx}`
naming does, and `attn_diff` stays a read-time switch.
"""
import json
import os

import numpy as np
import torch

from ._lib import check, dptr, fptr, lib, stream_ptr

FORMATS = {"uint8": (0, np.uint8, torch.uint8), "float16": (1, np.float16, torch.float16)}


def pack(maps, fmt="uint8"):
    """maps: cuda fp32, any shape -> cuda tensor of the same shape in the stored format."""
    code, _, tdt = FORMATS[fmt]
    m = maps.contiguous()
    out = torch.empty(m.shape, device=m.device, dtype=tdt)
    check(lib().maavss_attn_cache_pack(fptr(m), dptr(out), m.numel(), code, stream_ptr()))
    return out


def load(stored, attn_diff, fmt=None):
    """stored (clips, T, H, W) cuda uint8 / float16 -> (clips, T, H, W) fp32: dequantise + av_dataset.py:323-328 per clip."""
    if stored.dim() != 4:
        raise RuntimeError(f"attn_cache.load expects (clips, T, H, W), got {tuple(stored.shape)}")
    fmt = fmt or {torch.uint8: "uint8", torch.float16: "float16"}.get(stored.dtype)
    if fmt is None or FORMATS[fmt][2] != stored.dtype:
        raise RuntimeError(f"attn_cache.load: dtype {stored.dtype} is not a cache format")
    s = stored.contiguous()
    clips, T, H, W = s.shape
    out = torch.empty(s.shape, device=s.device, dtype=torch.float32)
    ws = torch.empty(max(clips, 1), device=s.device, dtype=torch.int32)
    check(lib().maavss_attn_cache_load(dptr(s), FORMATS[fmt][0], fptr(out), clips, T, H * W, int(bool(attn_diff)), dptr(ws),
                                       stream_ptr()))
    return out


class AttnFrameCache:
    """One memmap of `capacity` frames of `frame_hw` pixels.

    Files: `<path>` (frames), `<path>.valid` (one byte per frame), `<path>.json` (format, geometry, video index).
    A video is given a contiguous run of slots the first time it is seen (`reserve`); frame i of the video lives at slot
    `first + i`, so a clip starting at `true_idx = frame_hop * clip_idx` (av_dataset.py:257) is the slice
    `[first + true_idx, first + true_idx + num_frames)`.
    """

    def __init__(self, path, capacity=0, frame_hw=(256, 256), fmt="uint8", mode="r+"):
        self.path = path
        meta_path = path + ".json"
        if os.path.exists(meta_path) and mode != "w+":
            meta = json.load(open(meta_path))
            self.fmt, self.capacity = meta["format"], int(meta["capacity"])
            self.H, self.W = int(meta["height"]), int(meta["width"])
            self.index, self.used = {k: tuple(v) for k, v in meta["videos"].items()}, int(meta["used"])
            mm_mode = mode
        else:
            if capacity <= 0:
                raise ValueError("AttnFrameCache: a new cache needs capacity > 0 frames")
            if fmt not in FORMATS:
                raise ValueError(f"AttnFrameCache: unknown format {fmt!r} (uint8 | float16)")
            self.fmt, self.capacity = fmt, int(capacity)
            self.H, self.W = int(frame_hw[0]), int(frame_hw[1])
            self.index, self.used = {}, 0
            mm_mode = "w+"
        npdt = FORMATS[self.fmt][1]
        self._readonly = mm_mode == "r"
        self.frames = np.memmap(path, dtype=npdt, mode=mm_mode, shape=(self.capacity, self.H, self.W))
        self.valid = np.memmap(path + ".valid", dtype=np.uint8, mode=mm_mode, shape=(self.capacity,))
        self._stage_in = self._stage_out = self._h2d_done = None
        if mm_mode == "w+":
            self.flush()

    # ---- index ----
    def reserve(self, video, n_frames):
        """Slots for a video of n_frames frames (idempotent).  Returns the first slot."""
        if video in self.index:
            first, n = self.index[video]
            if n_frames > n:
                raise ValueError(f"AttnFrameCache: {video!r} was reserved with {n} frames, asked for {n_frames}")
            return first
        if self.used + n_frames > self.capacity:
            raise ValueError(f"AttnFrameCache: full ({self.used} + {n_frames} > {self.capacity} frames)")
        first = self.used
        # slots are handed out once per cache lifetime, but make a stale validity byte impossible anyway (a crash between a
        # write and the index save of an earlier run would otherwise turn into a hit on another video's frames)
        self.valid[first:first + n_frames] = 0
        self.index[video] = (first, int(n_frames))
        self.used += int(n_frames)
        self._save_index()              # the index is on disk before any frame of the video can be marked valid
        return first

    def _slots(self, video, start, n):
        if video not in self.index:
            return None
        first, total = self.index[video]
        if start < 0 or start + n > total:
            raise IndexError(f"AttnFrameCache: frames [{start}, {start + n}) outside {video!r} ({total} frames)")
        return first + start

    def has(self, video, start, n):
        """utilities.verify_files over img_{start..start+n-1} (av_dataset.py:258-260)."""
        s = self._slots(video, start, n)
        return s is not None and bool(np.all(self.valid[s:s + n]))

    def _save_index(self):
        if self._readonly:
            return
        meta = {"format": self.fmt, "capacity": self.capacity, "height": self.H, "width": self.W, "used": self.used,
                "videos": {k: list(v) for k, v in self.index.items()}}
        tmp = self.path + ".json.tmp"
        with open(tmp, "w") as f:
            json.dump(meta, f)
        os.replace(tmp, self.path + ".json")          # atomic: a reader sees the old or the new index, never a torn one

    def flush(self):
        if self._readonly:
            return
        self.frames.flush()
        self.valid.flush()
        self._save_index()

    def close(self):
        """Flush the maps and the index.  Also run from __del__; a cache that is simply dropped loses nothing either, since
        `reserve` saves the index itself and frame / validity bytes live in shared file mappings."""
        if getattr(self, "frames", None) is not None:
            self.flush()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- device <-> memmap ----
    def _pinned(self, which, n):
        cur = getattr(self, which)
        if cur is None or cur.shape[0] < n:
            cur = torch.empty((n, self.H, self.W), dtype=FORMATS[self.fmt][2]).pin_memory()
            setattr(self, which, cur)
        return cur[:n]

    def write(self, video, start, maps):
        """maps (n, 1, H, W) or (n, H, W) cuda fp32 = `_inference` output -> slots of `video` from frame `start`."""
        m = maps.reshape(-1, maps.shape[-2], maps.shape[-1])
        if tuple(m.shape[1:]) != (self.H, self.W):
            raise ValueError(f"AttnFrameCache: frame {tuple(m.shape[1:])} != cache geometry {(self.H, self.W)}")
        n = m.shape[0]
        s = self._slots(video, start, n)
        if s is None:
            raise KeyError(f"AttnFrameCache: reserve({video!r}, n_frames) first")
        host = self._pinned("_stage_out", n)
        host.copy_(pack(m, self.fmt), non_blocking=True)
        torch.cuda.current_stream().synchronize()
        self.frames[s:s + n] = host.numpy()
        self.valid[s:s + n] = 1

    def read_raw(self, requests):
        """requests [(video, start, n)] with one n -> (clips, n, H, W) cuda tensor in the stored format (H2D from pinned)."""
        n = requests[0][2]
        if self._h2d_done is not None:
            self._h2d_done.synchronize()          # the previous batch's copy out of the staging buffer has finished
        host = self._pinned("_stage_in", len(requests) * n)
        hv = host.numpy().reshape(len(requests), n, self.H, self.W)
        for c, (video, start, k) in enumerate(requests):
            if k != n:
                raise ValueError("AttnFrameCache.read_raw: all clips of a batch have the same length")
            s = self._slots(video, start, n)
            if s is None or not np.all(self.valid[s:s + n]):
                raise KeyError(f"AttnFrameCache: frames [{start}, {start + n}) of {video!r} are not cached")
            hv[c] = self.frames[s:s + n]
        dev = host.reshape(len(requests), n, self.H, self.W).cuda(non_blocking=True)
        self._h2d_done = torch.cuda.Event()
        self._h2d_done.record()
        return dev

    def read(self, requests, attn_diff):
        """-> (clips, 1, n, H, W) cuda fp32: the `attn` a batch of `gen_av_example` calls returns (av_dataset.py:319-332)."""
        return load(self.read_raw(requests), attn_diff, self.fmt).unsqueeze(1)

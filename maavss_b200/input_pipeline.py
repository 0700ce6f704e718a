"""Device-resident, double-buffered input pipeline for the train loops.

The reference builds every batch inside `Dataset.__getitem__` with `num_workers=0` (train.py:57-64): per clip it slices
the audio memmap (av_dataset.py:285-300, utilities.py:56-66, save_memmap.py:33-45), runs the STFT on the CPU, decodes
the frames, normalises them and pushes them through the ViT one frame at a time with a host round trip per frame
(av_dataset.py:302-333, video_attention.py:41-96) -- all in the training process, serialised with the optimizer step.

Here a whole batch goes through the device front end at once, on its own streams, while the previous step trains:

    host    AudioMemmap.slice_batch: B windows of the flat float32 memmap -> one pinned (B, n) buffer   (index-exact)
            frame_source(idx) -> decoded uint8 frames (T, H, W, 3) -> one pinned (B, T, H, W, 3) buffer
    copy    H2D on a copy stream (uint8 frames: 1/4 of the fp32 bytes the reference moves)
    front   maavss_frames_u8_normalize -> maavss_vit_attn_maps (all B*T frames, fused diff / clip normalise)
            maavss_stft_fwd (+ absmax normalise, + noise)                            on a front-end stream
    train   `get()` makes the training stream wait for that batch only

Video decoding itself (torchvision VideoClips / ffmpeg) is out of scope: `frame_source` supplies decoded
frames (synthetic in tests and benchmarks).  There is no CPU arithmetic path: the host only moves bytes.
"""
import ctypes
import os
import pickle

import numpy as np
import torch

from ._lib import check, dptr, fptr, lib

IMAGENET_MEAN, IMAGENET_STD = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)     # av_dataset.py:110-111


def save_audio_memmap(base_path, clips, names=None):
    """Write the reference's on-disk format (save_memmap.py:33-45): `audio_memmap.memmap` = all clips back to back as flat
    float32, `audio_index_map.obj` = pickle of [file names, (n_files, 2) int array of [start, end) offsets]."""
    os.makedirs(base_path, exist_ok=True)
    lengths = [int(np.asarray(c).shape[-1]) for c in clips]
    ends = np.cumsum(lengths)
    indexes = np.stack([ends - lengths, ends], 1)
    mm = np.memmap(os.path.join(base_path, "audio_memmap.memmap"), dtype="float32", mode="w+", shape=(int(ends[-1]),))
    for c, (a, b) in zip(clips, indexes):
        mm[a:b] = np.asarray(c, dtype=np.float32).reshape(-1)
    mm.flush()
    names = list(names) if names is not None else [f"clip_{i}.wav" for i in range(len(clips))]
    with open(os.path.join(base_path, "audio_index_map.obj"), "wb") as f:
        pickle.dump([names, indexes], f)
    return indexes


class AudioMemmap:
    """utilities.load_audio_map (utilities.py:56-66) + the memmap branch of AV_Dataset.get_audio (av_dataset.py:285-294)."""

    def __init__(self, base_path):
        with open(os.path.join(base_path, "audio_index_map.obj"), "rb") as f:
            self.files, self.indexes = pickle.load(f)
        self.indexes = np.asarray(self.indexes)
        self.length = int(self.indexes[-1, 1])
        self.map = np.memmap(os.path.join(base_path, "audio_memmap.memmap"), dtype="float32", mode="r", shape=(self.length,))

    @staticmethod
    def samples_start(clip_idx, frame_hop, fps, samplerate):
        """av_dataset.py:286-288 -- Python float arithmetic and round(), bit-exact."""
        seconds_start = (clip_idx * frame_hop) / fps
        return round(seconds_start * samplerate)

    def slice_batch(self, locations, frame_hop, fps, samplerate, n_samples, out=None):
        """locations: [(file_idx, clip_idx)] -> (B, n_samples) float32 (pinned when CUDA is present).  A window that runs past
        the end of the memmap is zero-filled at the tail (numpy slicing would return a short array the reference then fails on)."""
        B = len(locations)
        if out is None:
            out = torch.empty(B, n_samples, dtype=torch.float32, pin_memory=torch.cuda.is_available())
        o = out.numpy()
        for b, (file_idx, clip_idx) in enumerate(locations):
            a = int(self.indexes[file_idx][0]) + self.samples_start(clip_idx, frame_hop, fps, samplerate)
            e = min(a + n_samples, self.length)
            o[b, :e - a] = self.map[a:e]
            o[b, e - a:] = 0.0
        return out


def frames_u8_to_input(frames_u8, mean=IMAGENET_MEAN, std=IMAGENET_STD):
    """(.., H, W, 3) uint8 cuda -> (.., 3, H, W) fp32: permute, / 255, Normalize (av_dataset.py:312-317) in one kernel."""
    from ._lib import stream_ptr
    f = frames_u8.contiguous()
    if f.dtype != torch.uint8 or not f.is_cuda or f.shape[-1] != 3:
        raise RuntimeError("frames_u8_to_input expects a CUDA uint8 tensor (.., H, W, 3)")
    H, W = f.shape[-3], f.shape[-2]
    n = f.numel() // (H * W * 3)
    out = torch.empty(f.shape[:-3] + (3, H, W), device=f.device, dtype=torch.float32)
    m, s = (ctypes.c_float * 3)(*mean), (ctypes.c_float * 3)(*std)
    check(lib().maavss_frames_u8_normalize(dptr(f), fptr(out), n, H, W, ctypes.cast(m, ctypes.c_void_p), ctypes.cast(s, ctypes.c_void_p),
                                           stream_ptr()))
    return out


class ClipBatchPipeline:
    """submit(locations) stages one batch (host gather -> H2D -> device front end) on side streams; get() hands the oldest
    staged batch (x_stft, y_stft, attn, audio) to the current stream.  `slots` batches can be in flight."""

    def __init__(self, dataset, audio_map, frame_source, extractor, device=None, slots=2, noise_generator=None):
        self.ds, self.audio_map, self.frame_source, self.extractor = dataset, audio_map, frame_source, extractor
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.copy_stream = torch.cuda.Stream(device=self.device)
        self.front_stream = torch.cuda.Stream(device=self.device)
        self.slots = slots
        self.host = [None] * slots            # pinned staging (audio, frames) per slot
        self.h2d_done = [None] * slots
        self.pending, self.head = [], 0
        self.gen = noise_generator

    def _staging(self, slot, B, n, T, H, W):
        h = self.host[slot]
        if h is None or h[0].shape != (B, n) or h[1].shape != (B, T, H, W, 3):
            h = (torch.empty(B, n, dtype=torch.float32).pin_memory(), torch.empty(B, T, H, W, 3, dtype=torch.uint8).pin_memory())
            self.host[slot] = h
        return h

    def submit(self, locations, clip_indices=None):
        """locations: [(file_idx, clip_idx)] for the audio memmap; clip_indices: what frame_source is called with (default:
        position in `locations`)."""
        ds = self.ds
        B = len(locations)
        slot = self.head
        self.head = (self.head + 1) % self.slots
        if len(self.pending) >= self.slots:
            raise RuntimeError("ClipBatchPipeline: all slots in flight; call get() first")
        first = np.asarray(self.frame_source(clip_indices[0] if clip_indices is not None else 0))
        T, H, W = first.shape[:3]
        if self.h2d_done[slot] is not None:
            self.h2d_done[slot].synchronize()          # the staging buffers of this slot are free again
        audio_h, frames_h = self._staging(slot, B, ds.audio_sample_len, T, H, W)
        self.audio_map.slice_batch(locations, ds.frame_hop, ds.framerate, ds.samplerate, ds.audio_sample_len, out=audio_h)
        fh = frames_h.numpy()
        for b in range(B):
            fh[b] = first if b == 0 else np.asarray(self.frame_source(clip_indices[b] if clip_indices is not None else b))
        with torch.cuda.stream(self.copy_stream):
            audio = audio_h.to(self.device, non_blocking=True)
            frames = frames_h.to(self.device, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self.copy_stream)
        self.h2d_done[slot] = ev
        self.front_stream.wait_event(ev)
        with torch.cuda.stream(self.front_stream):
            video = frames_u8_to_input(frames.view(B * T, H, W, 3))
            attn = self.extractor.clip_attention(video, attn_diff=ds.attn_diff, clips=B)           # (B*T, 1, H', W')
            attn = attn.view(B, T, attn.shape[-2], attn.shape[-1]).unsqueeze(1)                     # (B, 1, T, H', W')
            noise = torch.randn(B, 2, ds.audio_sample_len // ds.hop, ds.fft_len // 2 + (0 if ds.trim_stft_end else 1),
                                device=self.device, generator=self.gen)
            x, y = ds.gen_stft_batch(audio, noise)
            done = torch.cuda.Event()
            done.record(self.front_stream)
        for t in (audio, frames, video, attn, noise, x, y):
            t.record_stream(self.front_stream)
        self.pending.append((done, (x, y, attn, audio)))

    def get(self):
        done, batch = self.pending.pop(0)
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(done)
        for t in batch:
            t.record_stream(cur)
        return batch

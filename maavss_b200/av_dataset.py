"""Host-side mirror of the reference `av_dataset.AV_Dataset` audio/attention front end.

Same constructor signature, attributes (`window`, `hop`, `fft_len`, `audio_sample_len`, ...) and
method names as /root/reference/av_dataset.py:15-377, with the arithmetic on the B200 through
libmaavss_b200.so.  Video-file decoding / clip indexing (torchvision VideoClips, ffmpeg) is out of
scope: clips come from a `clip_source` callable (idx -> (frames[T,3,H,W] float,
audio[n])) -- synthetic in tests and benches.

Device policy: a method returns its result on the device of its input (CPU in -> CPU out, like
the reference; CUDA in -> CUDA out with no host round trip).
"""
import torch

from . import audio_ops, utilities
from . import attn_ops


def _dev():
    if not torch.cuda.is_available():
        raise RuntimeError("maavss_b200 needs a CUDA device (B200); there is no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


class AV_Dataset:
    def __init__(self, num_frames=4, frame_hop=2, framerate=30, framesize=256, samplerate=16000, fft_len=512,
                 hops_per_frame=8, noise_std=0.01, use_polar=False, attn_diff=False, normalize_input_fft=True,
                 normalize_output_fft=True, autocontrast=False, compress_audio=False, shuffle_files=True,
                 data_path="./data/raw", max_clip_len=None, gen_stft=True, gen_video=True, wandb_run=None,
                 trim_stft_end=True, return_video_path=False, attn_frames_path=None,
                 clip_source=None, num_clips=0, attention_extractor=None, attn_cache=None, clip_location=None):
        self.gen_stft, self.gen_video = gen_stft, gen_video
        self.hop, self.audio_sample_len, _ = utilities.calc_hop_size(num_frames, hops_per_frame, framerate, samplerate)
        self.attention_extractor = attention_extractor        # a video_attention.VideoAttention (lazy: needs weights)
        self.num_frames, self.frame_hop = num_frames, frame_hop
        self.samplerate, self.framerate, self.framesize = samplerate, framerate, framesize
        self.fft_len = fft_len
        self.noise_std = noise_std
        self.normalize_input_fft, self.normalize_output_fft = normalize_input_fft, normalize_output_fft
        self.use_polar, self.attn_diff = use_polar, attn_diff
        self.autocontrast, self.compress_audio = autocontrast, compress_audio
        self.trim_stft_end = trim_stft_end
        self.return_video_path, self.attn_frames_path = return_video_path, attn_frames_path
        self.window = torch.hamming_window(self.fft_len)       # av_dataset.py:106
        self.clip_source, self._num_clips = clip_source, num_clips
        self.save_output_examples = False
        # raw frame cache in place of <attn_frames_path>/<video>/img_%05d.jpg (av_dataset.py:251-278): an
        # attn_cache.AttnFrameCache plus clip_location(idx) -> (video_name, clip_idx, frames_in_video), the
        # `video_clips.get_clip_location` / `video_paths` lookup of av_dataset.py:253-255
        self.attn_cache, self.clip_location = attn_cache, clip_location
        self.cache_ratio = [0, 0]                              # [hits, misses], av_dataset.py:96,261,271

    def toggle_dataset_mode(self, a, v):
        self.gen_stft, self.gen_video = a, v

    # ---- av_dataset.py:157-179 ----
    def stft_planes(self, audio):
        """Batched device path: audio (B, n) or (n,) -> planes (B, 2, T, F) / (2, T, F) on the GPU."""
        squeeze = audio.dim() == 1
        a = audio.reshape(1, -1) if squeeze else audio
        a = a.to(_dev(), torch.float32).contiguous()
        planes = audio_ops.stft_planes(a, self.window, self.fft_len, self.hop, self.trim_stft_end,
                                       self.normalize_input_fft, self.use_polar)
        return planes[0] if squeeze else planes

    def stft(self, audio, normalize=True, polar=False):
        """Reference layout: (F', T', 2) real view (or (2, F', T') = [mag; phase] when use_polar); the
        `normalize` / `polar` arguments are ignored exactly as in the reference (av_dataset.py:157,167,176)."""
        planes = self.stft_planes(audio).to(audio.device)
        if self.use_polar:
            return planes.transpose(-1, -2)                    # (2, F', T')
        return planes.permute(*range(planes.dim() - 3), -1, -2, -3)   # (F', T', 2) view; .permute(2,1,0) is contiguous

    # ---- av_dataset.py:181-201 ----
    def istft(self, stft):
        """stft: (2, T, F') planes (the layout the training loop holds) or batched (B, 2, T, F')."""
        squeeze = stft.dim() == 3
        p = stft.detach()
        p = (p.unsqueeze(0) if squeeze else p).to(_dev(), torch.float32).contiguous()
        audio = audio_ops.istft_audio(p, self.window, self.fft_len, self.hop, self.trim_stft_end,
                                      self.normalize_input_fft, self.use_polar)
        audio = audio[0] if squeeze else audio
        return audio.cpu() if not stft.is_cuda else audio        # reference forces .cpu() (av_dataset.py:194)

    # ---- av_dataset.py:217-220 ----
    def add_noise(self, tensor, noise=None):
        if noise is None:
            noise = torch.randn(tensor.shape, device=tensor.device)   # the gaussian draw stays torch's
        if tensor.is_cuda:
            return audio_ops.add_noise(tensor.contiguous(), noise.contiguous(), self.noise_std)
        return tensor + noise * self.noise_std

    # ---- av_dataset.py:335-342, batched ----
    def gen_stft_batch(self, audio, noise=None):
        """audio (B, n) -> x_stft, y_stft (B, 2, T, F') on the GPU (stft + permute + optional max-normalise + noise)."""
        y = self.stft_planes(audio)
        if self.normalize_output_fft:
            audio_ops.absmax_normalise_(y, 1e-7)
        if noise is None:
            noise = torch.randn(y.shape, device=y.device)
        x = audio_ops.add_noise(y, noise.to(y.device).contiguous(), self.noise_std)
        return x, y

    # ---- av_dataset.py:319-332, batched over the frames of one clip ----
    def gen_attention(self, video):
        """video (T,3,H,W) float (already transformed) -> attn (1,T,H,W) on the GPU."""
        if self.attention_extractor is None:
            raise RuntimeError("AV_Dataset was built without an attention_extractor (VideoAttention)")
        attn = self.attention_extractor.inference_device(video)            # (T,1,H,W) cuda
        attn = attn_ops.diff_normalise(attn, self.attn_diff)               # av_dataset.py:323-328
        return attn.permute(1, 0, 2, 3)

    # ---- av_dataset.py:251-283 ----
    def _cache_key(self, idx):
        if self.clip_location is None:
            return f"clip_{idx}", 0, self.num_frames           # every synthetic clip is its own "video"
        video, clip_idx, total = self.clip_location(idx)
        return video, self.frame_hop * clip_idx, total         # true_idx, av_dataset.py:257

    def get_frames_cached(self, idx):
        """-> (attn (1,T,H,W) cuda, video).  A hit reads the stored frames (video is zeros, as av_dataset.py:263); a miss
        runs the extractor and stores its per-frame maps.  The diff / normalise is the un-cached path's (:323-328), not the
        shape-changing `torch.diff(attn, )` of :268-269."""
        if self.attn_cache is None:
            raise RuntimeError("AV_Dataset was built without attn_cache (attn_cache.AttnFrameCache)")
        video_name, true_idx, total = self._cache_key(idx)
        if self.attn_cache.has(video_name, true_idx, self.num_frames):
            self.cache_ratio[0] += 1
            attn = self.attn_cache.read([(video_name, true_idx, self.num_frames)], self.attn_diff)[0]
            return attn, torch.zeros(3, self.num_frames, attn.shape[-2], attn.shape[-1])
        self.cache_ratio[1] += 1
        if self.attention_extractor is None:
            raise RuntimeError("AV_Dataset was built without an attention_extractor (VideoAttention)")
        video, _ = self.clip_source(idx)
        maps = self.attention_extractor.inference_device(video)            # (T,1,H,W) cuda, before the diff
        self.attn_cache.reserve(video_name, total)
        self.attn_cache.write(video_name, true_idx, maps)
        attn = attn_ops.diff_normalise(maps, self.attn_diff).permute(1, 0, 2, 3)
        return attn, video.permute(1, 0, 2, 3)

    def gen_av_example_cached(self, idx):
        attn, video = self.get_frames_cached(idx)
        x_stft, y_stft, audio = self.gen_stft_example(idx)
        return x_stft, y_stft, attn, audio, video

    def gen_stft_example(self, idx):
        _, audio = self.clip_source(idx)
        x, y = self.gen_stft_batch(audio.reshape(1, -1))
        return x[0], y[0], audio

    def gen_av_example(self, idx):
        video, audio = self.clip_source(idx)
        x, y = self.gen_stft_batch(audio.reshape(1, -1))
        attn = self.gen_attention(video)
        return x[0], y[0], attn, audio, video.permute(1, 0, 2, 3)

    def gen_video_example(self, idx):
        video, _ = self.clip_source(idx)
        return self.gen_attention(video), video.permute(1, 0, 2, 3)

    def __len__(self):
        return self._num_clips

    def __getitem__(self, idx):
        if self.clip_source is None:
            raise NotImplementedError("video-file clip indexing is out of scope; pass clip_source=")
        if self.gen_stft and self.gen_video:
            return self.gen_av_example(idx)
        if self.gen_stft:
            return self.gen_stft_example(idx)
        return self.gen_video_example(idx)


class STFT_Dataset(AV_Dataset):
    """av_dataset.py:381-491: the audio-only dataset shares stft/istft with AV_Dataset (:431-467)."""

    def __init__(self, **kw):
        kw.setdefault("gen_video", False)
        super().__init__(**kw)

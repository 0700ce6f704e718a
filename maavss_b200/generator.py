"""Host-side mirror of the tensor helpers of the reference `generator.DataGenerator`
(generator.py:60-133).  The file-loading generator loop itself is bit-rotted in the reference
 and out of scope; quirks of the helpers are reproduced (9.11): `ifft` returns
2*K samples with norm="forward" on the inverse, `cartesian_to_polar` returns cat(|z|, angle)."""
import torch

from . import audio_ops


class DataGenerator:
    def __init__(self, batch_size=2, num_vid_frames=4, framerate=30, framesize=256, samplerate=16000,
                 max_vid_frames=100, noise_std=0.01, center_fft=True, use_polar=True, normalize_input_fft=True,
                 shuffle_files=True, data_path="./data/raw", attention_extractor=None):
        assert batch_size > 1
        self.attention_extractor = attention_extractor
        self.max_vid_frames, self.batch_size, self.num_vid_frames = max_vid_frames, batch_size, num_vid_frames
        self.framerate, self.framesize, self.samplerate = framerate, framesize, samplerate
        self.noise_std = noise_std
        self.normalize_input_fft = normalize_input_fft
        self.fft_len = int((self.num_vid_frames / self.framerate) * self.samplerate)   # generator.py:44
        self.center_fft, self.use_polar = center_fft, use_polar
        self.device = torch.device("cuda")

    @staticmethod
    def _cuda(t):
        return t.to("cuda").contiguous()

    def fft(self, audio):                                    # generator.py:60-69
        a = self._cuda(audio.to(torch.float32))
        return audio_ops.dft_fwd(a, a.shape[-1] // 2, self.normalize_input_fft).to(audio.device)

    def ifft(self, fft):                                     # generator.py:71-81
        return audio_ops.dft_half_inv(self._cuda(fft), self.normalize_input_fft).to(fft.device)

    def cartesian_to_polar(self, cart):                      # generator.py:84-86
        ri = self._cuda(torch.view_as_real(cart).permute(0, 2, 1))        # (B,2,K)
        mp = audio_ops.magphase_split(ri.unsqueeze(-2)).squeeze(-2)       # planes (B,2,1,K)
        return torch.cat((mp[:, 0], mp[:, 1]), -1).to(cart.device)

    def polar_to_cartesian(self, polar):
        """The reference raises TypeError here (generator.py:94, torch.cat(a, b, -1)); the formula it
        intends -- polar (B, K, 2) = [mag, angle] -> ri (B, K, 2) -- is provided."""
        p = self._cuda(polar.permute(0, 2, 1)).unsqueeze(-2)              # (B,2,1,K)
        return audio_ops.magphase_combine(p).squeeze(-2).permute(0, 2, 1).to(polar.device)

    def complex_to_ri(self, tensor):                         # generator.py:99-103
        return torch.view_as_real(tensor).permute(0, 2, 1)

    def ri_to_complex(self, tensor):                         # generator.py:107-110
        return torch.view_as_complex(tensor.permute(0, 2, 1).contiguous())

    def center_fft_bins(self, fft_tensor):                   # generator.py:117-122
        return audio_ops.bins_center(self._cuda(fft_tensor)).to(fft_tensor.device)

    def decenter_fft_bins(self, fft_tensor):                 # generator.py:126-133
        return audio_ops.bins_decenter(self._cuda(fft_tensor)).to(fft_tensor.device)

    def reverse_process_fft(self, fft_tensor):               # generator.py:135-144
        if self.use_polar:
            fft_tensor = self.polar_to_cartesian(fft_tensor)
        if self.center_fft:
            fft_tensor = self.decenter_fft_bins(fft_tensor)
        return self.ri_to_complex(fft_tensor)

    def inference_to_audio(self, tensor):                    # generator.py:146-149
        return self.ifft(self.reverse_process_fft(tensor))

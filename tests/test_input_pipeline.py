"""Input pipeline: the audio memmap format / slicing host logic (CPU, bit-exact against the reference
arithmetic and, in the authoring container, the reference's own loader), and the device front end (GPU)."""
import importlib.util
import os
import sys

import numpy as np
import pytest
import torch

import synth
from maavss_b200 import input_pipeline as ip


def _make_map(tmp_path, lengths=(40000, 25000, 61234)):
    clips = [np.asarray(synth.audio_clip(n, seed=100 + i)) for i, n in enumerate(lengths)]
    idx = ip.save_audio_memmap(str(tmp_path), clips)
    return clips, idx


def test_audio_memmap_format_and_slicing(tmp_path):
    clips, idx = _make_map(tmp_path)
    assert idx.tolist() == [[0, 40000], [40000, 65000], [65000, 126234]]
    am = ip.AudioMemmap(str(tmp_path))
    assert am.length == 126234 and am.map.dtype == np.float32
    # av_dataset.py:286-294: seconds_start = clip_idx * frame_hop / fps; samples_start = round(seconds_start * sr)
    locs = [(0, 0), (1, 3), (2, 7), (0, 11), (2, 200)]
    out = am.slice_batch(locs, frame_hop=2, fps=30, samplerate=16000, n_samples=6336)
    assert out.shape == (5, 6336) and out.dtype == torch.float32
    for row, (f, c) in zip(out, locs):
        start = round(((c * 2) / 30) * 16000)
        whole = np.concatenate(clips)                    # the reference indexes the FLAT memmap: a window may run into the next file
        a = int(idx[f][0]) + start
        ref = whole[a:a + 6336]
        assert np.array_equal(row.numpy()[:len(ref)], ref)
        assert not row.numpy()[len(ref):].any()
    assert ip.AudioMemmap.samples_start(3, 2, 30, 16000) == 3200 and ip.AudioMemmap.samples_start(1, 1, 30, 16000) == 533


@pytest.mark.skipif(not os.path.isdir("/root/reference"), reason="the reference only exists in the authoring container")
def test_reference_loader_reads_our_memmap(tmp_path):
    import types
    import importlib
    stubbed = []
    for name in ("matplotlib", "matplotlib.pyplot", "cv2", "wandb", "torchsummary"):
        try:
            importlib.import_module(name)
        except ImportError:
            sys.modules[name] = types.ModuleType(name)
            stubbed.append(name)
    if "matplotlib" in stubbed:
        sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    sys.path.insert(0, "/root/reference")
    try:
        spec = importlib.util.spec_from_file_location("ref_utilities", "/root/reference/utilities.py")
        ref_util = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(ref_util)
    finally:
        sys.path.remove("/root/reference")
        for name in stubbed:
            sys.modules.pop(name, None)
    clips, idx = _make_map(tmp_path)
    m, index_map = ref_util.load_audio_map(str(tmp_path))
    assert m is not None and m.shape == (126234,)
    assert np.array_equal(np.asarray(index_map[1]), idx)
    assert np.array_equal(np.asarray(m[40000:40010]), clips[1][:10])


@pytest.mark.gpu
def test_frames_u8_normalize_bit_exact():
    g = synth.gen(3)
    fr = torch.randint(0, 256, (5, 40, 36, 3), generator=g, dtype=torch.uint8)
    out = ip.frames_u8_to_input(fr.cuda()).cpu()
    v = fr.permute(0, 3, 1, 2).type(torch.float32) / 255.                      # av_dataset.py:312-313
    mean, std = torch.tensor(ip.IMAGENET_MEAN).view(1, 3, 1, 1), torch.tensor(ip.IMAGENET_STD).view(1, 3, 1, 1)
    ref = (v - mean) / std                                                     # torchvision Normalize
    assert torch.equal(out, ref)


@pytest.mark.gpu
def test_clip_batch_pipeline_equals_per_clip_path(tmp_path):
    """One pipelined batch == the per-clip dataset calls (gen_stft_example + gen_attention) on the same clips and noise."""
    from maavss_b200 import av_dataset, video_attention
    from oracle import audio as oa
    clips, idx = _make_map(tmp_path)
    am = ip.AudioMemmap(str(tmp_path))
    va = video_attention.VideoAttention(state_dict=synth.vit_state_dict(depth=2))
    ds = av_dataset.AV_Dataset(num_frames=12, frame_hop=2, fft_len=256, hops_per_frame=8, noise_std=0.1, attn_diff=True,
                               attention_extractor=va)
    g = synth.gen(5)
    vids = [torch.randint(0, 256, (12, 64, 72, 3), generator=g, dtype=torch.uint8).numpy() for _ in range(4)]
    locs = [(0, 1), (1, 0), (2, 9), (0, 4)]
    pipe = ip.ClipBatchPipeline(ds, am, lambda i: vids[i], va, noise_generator=torch.Generator(device="cuda").manual_seed(7))
    pipe.submit(locs, clip_indices=[0, 1, 2, 3])
    pipe.submit(locs[::-1], clip_indices=[3, 2, 1, 0])          # two batches in flight
    x, y, attn, audio = pipe.get()
    x2, y2, attn2, audio2 = pipe.get()
    torch.cuda.synchronize()
    assert attn.shape == (4, 1, 12, 64, 72) and x.shape == y.shape == (4, 2, 96, 128) and audio.shape == (4, 6336)
    w = oa.hamming_window(256)
    whole = np.concatenate(clips)
    for b, (f, c) in enumerate(locs):
        a = int(idx[f][0]) + round(((c * 2) / 30) * 16000)
        assert np.array_equal(audio[b].cpu().numpy(), whole[a:a + 6336])                       # index-exact
        ref_y = oa.stft_planes(torch.from_numpy(whole[a:a + 6336]), w, 256, 66, True, True, normalize_output=True)
        assert float((y[b].cpu() - ref_y).abs().max()) < 2e-5 * float(ref_y.abs().max())
        per_clip = ds.gen_attention(ip.frames_u8_to_input(torch.from_numpy(vids[b]).cuda()))   # (1, T, H, W)
        assert torch.equal(attn[b], per_clip)
    assert float((x - y).std()) == pytest.approx(0.1, rel=0.05)                                # noise_std * N(0, 1)
    assert torch.equal(attn2[0], attn[3]) and torch.equal(audio2[3], audio[0])

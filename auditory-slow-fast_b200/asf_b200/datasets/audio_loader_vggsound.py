"""VGG-Sound twin of the loader, on the GPU extraction kernels.

Mirror of /root/reference/audio_slowfast/datasets/audio_loader_vggsound.py:
  ``get_start_end_idx(audio_size, clip_size, clip_idx, num_clips)``  :8-36  (no start_sample argument here)
  ``_log_specgram(cfg, audio, window_size, step_size, eps)``         :57-69 -- NO hop quirk (hop = step, not
        win - step as in the EPIC loader, audio_loader_epic.py:74-75), n_fft 2048 and 128 mels hard-coded
  ``_extract_sound_feature(cfg, samples, start_idx, end_idx)``       :72-84 -- a file shorter than one clip is
        transformed whole and edge-padded to NUM_FRAMES rows; otherwise the slice, with no padding
  ``pack_audio(cfg, audio_record, temporal_sample_index)``           :39-54 -- reads the .wav with librosa in the
        reference; here ``samples`` are passed in (waveform I/O is outside the hot path)
plus the batched entry ``logmel_batch_vggsound`` for waveforms already on the device.
"""
from __future__ import annotations

import random

import numpy as np
import torch

from .audio_loader_epic import _stft_sizes, logmel_batch, logmel_cuda


def get_start_end_idx(audio_size, clip_size, clip_idx, num_clips):
    delta = max(audio_size - clip_size, 0)
    if clip_idx == -1:
        start_idx = random.uniform(0, delta)
    else:
        start_idx = np.linspace(0, delta, num=num_clips)[clip_idx]
    end_idx = start_idx + clip_size - 1
    return start_idx, end_idx


def _log_specgram(cfg, audio, window_size=10, step_size=5, eps=1e-6):
    """float32 numpy [N] -> float32 numpy [T, 128]; hop = step_size (the VGG-Sound loader has no EPIC hop quirk)."""
    win, hop = _stft_sizes(cfg, window_size, step_size, epic_quirk=False)
    x = torch.as_tensor(np.ascontiguousarray(audio, dtype=np.float32)).reshape(1, -1).cuda(non_blocking=True)
    out = logmel_cuda(x, cfg.AUDIO_DATA.SAMPLING_RATE, 2048, win, hop, 128, eps, None)
    return out[0, 0].cpu().numpy()


def _extract_sound_feature(cfg, samples, start_idx, end_idx):
    clip = int(round(cfg.AUDIO_DATA.SAMPLING_RATE * cfg.AUDIO_DATA.CLIP_SECS))
    if samples.shape[0] < clip:
        spec = _log_specgram(cfg, samples, window_size=cfg.AUDIO_DATA.WINDOW_LENGTH, step_size=cfg.AUDIO_DATA.HOP_LENGTH)
        pad = cfg.AUDIO_DATA.NUM_FRAMES - spec.shape[0]
        spec = np.pad(spec, ((0, pad), (0, 0)), "edge")  # (a negative pad raises, as in the reference)
    else:
        samples = samples[start_idx:end_idx]
        spec = _log_specgram(cfg, samples, window_size=cfg.AUDIO_DATA.WINDOW_LENGTH, step_size=cfg.AUDIO_DATA.HOP_LENGTH)
    return torch.tensor(spec).unsqueeze(0)


def pack_audio(cfg, audio_record, temporal_sample_index, samples=None, sr=None):
    """``samples``: the decoded waveform of ``audio_record`` (the reference loads
    ``cfg.VGGSOUND.AUDIO_DATA_DIR/<video>.wav`` with librosa here)."""
    if samples is None:
        raise RuntimeError("asf_b200: pass the decoded waveform (file decoding is outside the hot path)")
    if sr is not None:
        assert sr == cfg.AUDIO_DATA.SAMPLING_RATE, \
            "Audio sampling rate ({}) does not match target sampling rate ({})".format(sr, cfg.AUDIO_DATA.SAMPLING_RATE)
    start_idx, end_idx = get_start_end_idx(
        samples.shape[0], int(round(cfg.AUDIO_DATA.SAMPLING_RATE * cfg.AUDIO_DATA.CLIP_SECS)), temporal_sample_index,
        cfg.TEST.NUM_ENSEMBLE_VIEWS)
    return _extract_sound_feature(cfg, samples, int(start_idx), int(end_idx))


def logmel_batch_vggsound(cfg, waveforms, augment=None):
    """CUDA float32 [B, N] -> per-pathway model inputs with the VGG-Sound framing (hop = HOP_LENGTH)."""
    return logmel_batch(cfg, waveforms, epic_quirk=False, augment=augment)

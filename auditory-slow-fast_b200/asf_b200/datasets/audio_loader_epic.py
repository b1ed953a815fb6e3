"""Log-mel extraction on the GPU behind the reference's loader signatures.

Mirror of /root/reference/audio_slowfast/datasets/audio_loader_epic.py:
  ``_log_specgram(cfg, audio, window_size, step_size, eps)``                       :66-97
  ``_extract_sound_feature(cfg, samples, audio_record, start_idx, end_idx, ...)``  :100-156
  ``pack_audio(cfg, audio_dataset, audio_record, temporal_sample_index, ...)``     :16-63
plus the additive batched entry ``logmel_batch`` (waveforms already on the device -> per-pathway model
inputs), which is what the training / inference step uses: the reference runs librosa per sample inside
DataLoader workers; here the loader hands over raw waveforms and extraction is one kernel per batch.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np
import torch

from .. import _lib as L
from .utils import get_start_end_idx, slow_pathway_indices

_TABLES = {}


def _stft_sizes(cfg, window_size, step_size, epic_quirk=True):
    sr = cfg.AUDIO_DATA.SAMPLING_RATE
    win = int(round(window_size * sr / 1e3))
    hop = int(round(step_size * sr / 1e3))
    if epic_quirk and win - hop > 0:  # audio_loader_epic.py:74-75 (absent from the VGG-Sound twin)
        hop = win - hop
    return win, hop


def _tables(device, sr, n_fft, win, n_mels):
    key = (device.index, sr, n_fft, win, n_mels)
    t = _TABLES.get(key)
    if t is None:
        n_bins = n_fft // 2 + 1
        nbp = (n_bins + 7) // 8 * 8
        basis = torch.zeros((win, nbp, 2), dtype=torch.float32, device=device)
        melT = torch.zeros((nbp, n_mels), dtype=torch.float32, device=device)
        rng = torch.zeros((n_mels, 2), dtype=torch.int32, device=device)
        L.call("asf_logmel_tables", L.ptr(basis), L.ptr(melT), L.ptr(rng), sr, n_fft, win, n_mels, nbp, L.stream_ptr())
        t = (basis, melT, rng, nbp)
        _TABLES[key] = t
    return t


def _tc_tables(device, sr, n_fft, win, hop, n_mels):
    """Tables of the tensor-core kernel (pre-split TF32 basis image, per-bin mel pair); None when the mel layout
    cannot be expressed as adjacent filter pairs (then the CUDA-core kernel runs)."""
    key = ("tc", device.index, sr, n_fft, win, hop, n_mels)
    if key not in _TABLES:
        nbp = (n_fft // 2 + 1 + 103) // 104 * 104
        btab = torch.zeros(L.lib().asf_logmel_tc_btab_floats(n_fft), dtype=torch.float32, device=device)
        melw = torch.zeros((nbp, 2), dtype=torch.float32, device=device)
        melj = torch.zeros(nbp, dtype=torch.int32, device=device)
        rc = L.lib().asf_logmel_tc_tables(L.ptr(btab), L.ptr(melw), L.ptr(melj), sr, n_fft, win, hop, n_mels,
                                          L.stream_ptr())
        _TABLES[key] = (btab, melw, melj) if rc == 0 else None
    return _TABLES[key]


def logmel_cuda(waveforms, sr=24000, n_fft=2048, win=240, hop=120, n_mels=128, eps=1e-6, num_frames=None, impl=None):
    """waveforms: CUDA float32 [B, N] -> CUDA float32 [B, 1, T, n_mels] with T = max(1 + N//hop, num_frames)
    (rows past the last frame replicate it, as np.pad(..., "edge") does in the reference).
    ``impl``: None = the tcgen05 kernel when the configuration allows it (win = 240, hop = 120: every shipped
    config), else the CUDA-core kernel; "simt" / "tc" force one of them."""
    L.require_cuda_sm100()
    if waveforms.dim() != 2 or waveforms.dtype != torch.float32 or not waveforms.is_cuda:
        raise ValueError("logmel_cuda expects a CUDA float32 tensor of shape [B, N]")
    if waveforms.stride(1) != 1:
        waveforms = waveforms.contiguous()
    B, N = waveforms.shape
    T = 1 + N // hop
    T_out = max(T, num_frames or 0)
    out = torch.empty((B, 1, T_out, n_mels), dtype=torch.float32, device=waveforms.device)
    if impl is None:
        impl = os.environ.get("ASF_LOGMEL_IMPL") or None
    tc = None
    if impl != "simt" and L.lib().asf_logmel_tc_supported(n_fft, win, hop):
        tc = _tc_tables(waveforms.device, sr, n_fft, win, hop, n_mels)
    if impl == "tc" and tc is None:
        raise RuntimeError("logmel_cuda: the tensor-core kernel does not support n_fft=%d win=%d hop=%d n_mels=%d"
                           % (n_fft, win, hop, n_mels))
    if tc is not None:
        L.call("asf_logmel_fwd_tc", L.ptr(waveforms), waveforms.stride(0), B, N, n_fft, win, hop, n_mels,
               C.c_float(eps), L.ptr(tc[0]), L.ptr(tc[1]), L.ptr(tc[2]), L.ptr(out), T_out, L.stream_ptr())
        return out
    basis, melT, rng, nbp = _tables(waveforms.device, sr, n_fft, win, n_mels)
    L.call("asf_logmel_fwd", L.ptr(waveforms), waveforms.stride(0), B, N, n_fft, win, hop, n_mels, C.c_float(eps),
           L.ptr(basis), L.ptr(melT), L.ptr(rng), nbp, L.ptr(out), T_out, L.stream_ptr())
    return out


def logmel_batch(cfg, waveforms, epic_quirk=True, augment=None):
    """Batched waveform -> model inputs: CUDA float32 [B, N] -> ``[slow (B,1,T//alpha,F), fast (B,1,T,F)]``
    (or ``[x]`` for single-pathway archs), i.e. _extract_sound_feature + pack_pathway_output for a batch.
    ``augment``: a ``datasets.spec_augment.SpecAugmenter`` applied to the spectrograms BEFORE the pathway packing,
    which is where the reference's training dataset applies SpecAugment (datasets/epickitchens.py:160-170)."""
    win, hop = _stft_sizes(cfg, cfg.AUDIO_DATA.WINDOW_LENGTH, cfg.AUDIO_DATA.HOP_LENGTH, epic_quirk)
    spec = logmel_cuda(waveforms, cfg.AUDIO_DATA.SAMPLING_RATE, cfg.AUDIO_DATA.N_FFT, win, hop,
                       cfg.AUDIO_DATA.NUM_FREQUENCIES, 1e-6, cfg.AUDIO_DATA.NUM_FRAMES)
    if augment is not None:
        spec = augment(spec)
    if cfg.MODEL.ARCH in cfg.MODEL.SINGLE_PATHWAY_ARCH:
        return [spec]
    if cfg.MODEL.ARCH in cfg.MODEL.MULTI_PATHWAY_ARCH:
        B, _, T, F = spec.shape
        idx = slow_pathway_indices(T, cfg.SLOWFAST.ALPHA, device=spec.device).to(torch.int32)
        slow = torch.empty((B, 1, idx.numel(), F), dtype=torch.float32, device=spec.device)
        L.call("asf_gather_rows", L.ptr(spec), L.ptr(slow), L.ptr(idx), B, T, idx.numel(), F, L.stream_ptr())
        return [slow, spec]
    raise NotImplementedError(
        "Model arch {} is not in {}".format(cfg.MODEL.ARCH, cfg.MODEL.SINGLE_PATHWAY_ARCH + cfg.MODEL.MULTI_PATHWAY_ARCH))


def _log_specgram(cfg, audio, window_size=10, step_size=5, eps=1e-6):
    """Reference signature: float32 numpy [N] -> float32 numpy [T, n_mels] (log-mel, transposed).
    Host buffers in, host buffers out; the arithmetic runs on the GPU (one clip = a batch of one)."""
    win, hop = _stft_sizes(cfg, window_size, step_size, True)
    x = torch.as_tensor(np.ascontiguousarray(audio, dtype=np.float32)).reshape(1, -1).cuda(non_blocking=True)
    out = logmel_cuda(x, cfg.AUDIO_DATA.SAMPLING_RATE, cfg.AUDIO_DATA.N_FFT, win, hop,
                      cfg.AUDIO_DATA.NUM_FREQUENCIES, eps, None)
    return out[0, 0].cpu().numpy()


def _extract_sound_feature(cfg, samples, audio_record, start_idx, end_idx, transform=None):
    """audio_loader_epic.py:100-156: slice the clip (whole action when it is shorter than a clip), optional
    waveform transform, log-mel, edge-pad to NUM_FRAMES rows; returns a CPU FloatTensor [1, T, n_mels]."""
    clip = int(round(cfg.AUDIO_DATA.SAMPLING_RATE * cfg.AUDIO_DATA.CLIP_SECS))
    if audio_record.num_audio_samples < clip:
        samples = samples[audio_record.start_audio_sample:audio_record.end_audio_sample]
    else:
        samples = samples[start_idx:end_idx]
    if transform is not None:
        samples = transform(samples, sample_rate=cfg.AUDIO_DATA.SAMPLING_RATE)
    spec = _log_specgram(cfg=cfg, audio=samples, window_size=cfg.AUDIO_DATA.WINDOW_LENGTH,
                         step_size=cfg.AUDIO_DATA.HOP_LENGTH)
    pad = cfg.AUDIO_DATA.NUM_FRAMES - spec.shape[0]
    if pad > 0:
        spec = np.pad(spec, ((0, pad), (0, 0)), "edge")
    return torch.tensor(spec).unsqueeze(0)


def pack_audio(cfg, audio_dataset, audio_record, temporal_sample_index, transform=None):
    """audio_loader_epic.py:16-63.  ``audio_dataset`` is any mapping video name -> float32 sample array."""
    samples = audio_dataset[audio_record.untrimmed_video_name][()]
    start_idx, end_idx = get_start_end_idx(
        audio_record.num_audio_samples,
        int(round(cfg.AUDIO_DATA.SAMPLING_RATE * cfg.AUDIO_DATA.CLIP_SECS)),
        temporal_sample_index, cfg.TEST.NUM_ENSEMBLE_VIEWS, start_sample=audio_record.start_audio_sample)
    return _extract_sound_feature(cfg=cfg, samples=samples, audio_record=audio_record, start_idx=int(start_idx),
                                  end_idx=int(end_idx), transform=transform)

"""Clip sampling and pathway packing (mirror of /root/reference/audio_slowfast/datasets/utils.py:12-72)."""
import random

import numpy as np
import torch


def get_start_end_idx(audio_size, clip_size, clip_idx, num_clips, start_sample=0):
    """Clip window inside an audio of ``audio_size`` samples (datasets/utils.py:12-40): random jitter when
    ``clip_idx == -1``, otherwise the clip_idx-th of ``num_clips`` uniformly spaced windows.  Returns float
    (start, end) with end = start + clip_size - 1; the caller truncates with int()."""
    delta = max(audio_size - clip_size, 0)
    if clip_idx == -1:
        start_idx = random.uniform(0, delta)
    else:
        start_idx = np.linspace(0, delta, num=num_clips)[clip_idx]
    end_idx = start_idx + clip_size - 1
    return start_sample + start_idx, start_sample + end_idx


def slow_pathway_indices(num_frames, alpha, device=None):
    """The reference's temporal sub-sampling indices: ``torch.linspace(0, T-1, T//alpha).long()``
    (datasets/utils.py:59-63) -- NOT alpha*i (for T=400, alpha=4 there are three stride-5 jumps)."""
    return torch.linspace(0, num_frames - 1, num_frames // alpha, device=device).long()


def pack_pathway_output(cfg, spectrogram):
    """[C, T, F] spectrogram -> list of per-pathway tensors (datasets/utils.py:43-72).  Works on CPU or CUDA
    tensors; a pure index_select, bit-exact by construction."""
    if cfg.MODEL.ARCH in cfg.MODEL.SINGLE_PATHWAY_ARCH:
        return [spectrogram]
    if cfg.MODEL.ARCH in cfg.MODEL.MULTI_PATHWAY_ARCH:
        idx = slow_pathway_indices(spectrogram.shape[1], cfg.SLOWFAST.ALPHA).to(spectrogram.device)
        return [torch.index_select(spectrogram, 1, idx), spectrogram]
    raise NotImplementedError(
        "Model arch {} is not in {}".format(cfg.MODEL.ARCH, cfg.MODEL.SINGLE_PATHWAY_ARCH + cfg.MODEL.MULTI_PATHWAY_ARCH))

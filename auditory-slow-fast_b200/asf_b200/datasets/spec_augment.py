"""SpecAugment on the GPU for a whole batch, between the log-mel kernel and the stems.

Mirror of /root/reference/audio_slowfast/datasets/spec_augment.py (``spec_augment`` :144-191 = ``time_mask(freq_mask(
time_warp(x)))``, with ``time_warp`` going through datasets/sparse_image_warp.py) as the reference applies it to every
training spectrogram (datasets/epickitchens.py:160-168).  There it runs per sample on the CPU inside the DataLoader
workers; with the extraction on the GPU that would mean a device -> host -> device round trip per batch, so here it is
two launches on the batch that the extraction kernel just wrote (csrc/specaug.cu):

    aug = SpecAugmenter(seed=0)
    fast = aug(fast)                      # (B, 1, T, F) CUDA float32 -> same shape, augmented

The random quantities (warp column and displacement, the 3 x 3 normal block of the reference's spline fit, the mask
intervals) come from a counter-based generator on the device (Philox4x32-10): the same DISTRIBUTIONS as the reference's
``random.randrange`` / ``torch.randn`` calls, not the same bits, and a CUDA graph that captured the two launches draws
fresh numbers on every replay (the counter lives in device memory).  ``spec_augment(spec)`` keeps the reference's
per-sample signature ((1, F, T) tensor in, same out).
"""
from __future__ import annotations

import ctypes as C

import torch

from .. import _lib as L

PARAM_BYTES = C.sizeof(L.SpecAugParams)


class SpecAugmenter:
    """Batched SpecAugment with the reference's defaults (spec_augment.py:144-151): 2 frequency masks of width < 27,
    2 time masks of width < 25, time warp W = 5."""

    def __init__(self, seed=0, num_freq_masks=2, num_time_masks=2, F=27, T=25, W=5):
        if num_freq_masks > L.SPECAUG_MAX_MASKS or num_time_masks > L.SPECAUG_MAX_MASKS:
            raise ValueError("at most %d masks of a kind" % L.SPECAUG_MAX_MASKS)
        self.seed = int(seed) & 0xFFFFFFFFFFFFFFFF
        self.nf, self.nt, self.F, self.T, self.W = num_freq_masks, num_time_masks, F, T, W
        self._counter = {}
        self._bufs = {}

    def _state(self, spec):
        key = (spec.device.index, tuple(spec.shape))
        if key not in self._bufs:
            B = spec.shape[0]
            self._bufs[key] = (torch.zeros(B * PARAM_BYTES, dtype=torch.uint8, device=spec.device),
                               torch.empty_like(spec))
        if spec.device.index not in self._counter:
            self._counter[spec.device.index] = torch.zeros(1, dtype=torch.int64, device=spec.device)
        return self._bufs[key] + (self._counter[spec.device.index],)

    def draw(self, spec):
        """Draws one set of parameters per clip of ``spec`` (B, 1, T, F); returns the device buffer of
        asf_specaug_params_t records (kept by this object and overwritten by the next call)."""
        _check(spec)
        params, _, counter = self._state(spec)
        B, _, T, F = spec.shape
        L.call("asf_specaug_params", L.ptr(spec), B, T, F, C.c_ulonglong(self.seed), L.ptr(counter), self.W, self.F, self.T,
               self.nf, self.nt, L.ptr(params), L.stream_ptr())
        return params

    def __call__(self, spec):
        """(B, 1, T, F) CUDA float32 -> augmented tensor of the same shape (a buffer owned by this object: static
        addresses, as a captured training step needs)."""
        params = self.draw(spec)
        _, out, _ = self._state(spec)
        return apply_params(spec, params, out)


def _check(spec):
    L.require_cuda_sm100()
    if spec.dim() != 4 or spec.shape[1] != 1 or spec.dtype != torch.float32 or not spec.is_cuda or not spec.is_contiguous():
        raise ValueError("SpecAugment expects a contiguous CUDA float32 tensor of shape (B, 1, T, F)")


def apply_params(spec, params, out=None):
    """Applies explicit per-clip parameters (device uint8 buffer of asf_specaug_params_t records)."""
    _check(spec)
    B, _, T, F = spec.shape
    if params.numel() != B * PARAM_BYTES:
        raise ValueError("expected %d parameter records" % B)
    if out is None:
        out = torch.empty_like(spec)
    L.call("asf_specaug_apply", L.ptr(spec), L.ptr(out), L.ptr(params), B, T, F, L.stream_ptr())
    return out


def pack_params(records, device):
    """Host dicts {"coef": (a_f, a_t, a_0), "fmask": [(z, e), ...], "tmask": [(z, e), ...]} -> device parameter buffer."""
    arr = (L.SpecAugParams * len(records))()
    for r, d in zip(arr, records):
        for j in range(3):
            r.coef[j] = float(d["coef"][j])
        for name in ("fmask", "tmask"):
            ivs = list(d.get(name, ()))
            if len(ivs) > L.SPECAUG_MAX_MASKS:
                raise ValueError("at most %d masks of a kind" % L.SPECAUG_MAX_MASKS)
            for m, (z, e) in enumerate(ivs):
                getattr(r, name)[m][0], getattr(r, name)[m][1] = int(z), int(e)
    return torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8).to(device)


_DEFAULT = None


def spec_augment(spec, num_freq_masks=2, num_time_masks=2, F=27, T=25, W=5):
    """Reference signature (spec_augment.py:144): ``spec`` (1, F, T) -> (1, F, T).  A CUDA tensor is augmented on the
    device (one clip = a batch of one); host tensors are copied over and back."""
    global _DEFAULT
    key = (num_freq_masks, num_time_masks, F, T, W)
    if _DEFAULT is None or _DEFAULT[0] != key:
        _DEFAULT = (key, SpecAugmenter(torch.initial_seed(), num_freq_masks, num_time_masks, F, T, W))
    x = spec.permute(0, 2, 1).unsqueeze(0).contiguous().float()  # (1, 1, T, F)
    was_cuda = x.is_cuda
    y = _DEFAULT[1](x.cuda())
    y = y[0].permute(0, 2, 1)
    return y.clone() if was_cuda else y.cpu()

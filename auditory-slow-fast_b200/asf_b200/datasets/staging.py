"""Waveform staging: pinned-host ring -> device, and clip slicing on the device.

What the reference does per item on the host (/root/reference/audio_slowfast/datasets/audio_loader_epic.py:16-63:
read the WHOLE video array from HDF5 (:45), ``get_start_end_idx`` (datasets/utils.py:12-40), slice (:138)) becomes:

  * ``clip_starts``      the reference's start arithmetic (random crop in training, ``np.linspace(0, delta, 10)[i]`` for
                         the 10 test views, datasets/epickitchens.py:46-47) for a whole batch, float64 on the host like
                         the reference, truncated with ``int()`` like its caller (audio_loader_epic.py:53).
  * ``ResidentAudio``    raw waveforms kept in HBM (one flat fp32 buffer); a training crop, the 10 test views or a
                         sliding window are strided reads of it (``asf_gather_clips``), not copies made on the host.
  * ``WaveformRing``     a ring of pinned host batches with their device twins on a copy stream, so that the H2D copy
                         of batch i + 1 overlaps the step of batch i (the training step reads a static device buffer:
                         ``TrainStep.load`` becomes a device-to-device copy ordered by an event).
"""
from __future__ import annotations

import random

import numpy as np
import torch

from .. import _lib as L
from .utils import get_start_end_idx


def clip_starts(num_samples, clip_size, clip_idx, num_clips, start_sample=0):
    """Start sample of clip ``clip_idx`` of every record (lists / arrays broadcast): ``int(start_idx)`` of the
    reference's ``get_start_end_idx`` (same RNG use for clip_idx == -1).  Returns int64 numpy."""
    num_samples = np.atleast_1d(np.asarray(num_samples, dtype=np.int64))
    start_sample = np.broadcast_to(np.asarray(start_sample, dtype=np.int64), num_samples.shape)
    clip_idx = np.broadcast_to(np.asarray(clip_idx, dtype=np.int64), num_samples.shape)
    out = np.empty(num_samples.shape, dtype=np.int64)
    for i in range(num_samples.size):
        s, _ = get_start_end_idx(int(num_samples[i]), clip_size, int(clip_idx[i]), num_clips, start_sample=int(start_sample[i]))
        out[i] = int(s)
    return out


def multi_view_starts(num_samples, clip_size, num_views, start_sample=0):
    """(records, views) start samples of the test-time ensemble views (TEST.NUM_ENSEMBLE_VIEWS = 10)."""
    num_samples = np.atleast_1d(np.asarray(num_samples, dtype=np.int64))
    start_sample = np.broadcast_to(np.asarray(start_sample, dtype=np.int64), num_samples.shape)
    return np.stack([clip_starts(num_samples, clip_size, v, num_views, start_sample) for v in range(num_views)], 1)


class ResidentAudio:
    """Raw waveforms resident on the device as ONE flat fp32 buffer."""

    def __init__(self, waveforms, device=None):
        """``waveforms``: mapping name -> 1-D float32 array / tensor (or a list, named by index)."""
        L.require_cuda_sm100()
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        items = list(waveforms.items()) if hasattr(waveforms, "items") else list(enumerate(waveforms))
        self.names = [k for k, _ in items]
        lens = [int(len(v)) for _, v in items]
        self.offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
        host = torch.empty(int(self.offsets[-1]), dtype=torch.float32).pin_memory()
        for (k, v), o, n in zip(items, self.offsets, lens):
            host[o:o + n] = torch.as_tensor(np.asarray(v, dtype=np.float32))
        self.samples = host.to(dev, non_blocking=True)
        self._host = host  # keeps the pinned staging buffer alive until the copy is done
        self.index = {k: i for i, k in enumerate(self.names)}
        self.device = dev

    def crop(self, names, starts, n, out=None, valid=None):
        """out[b] = waveform(names[b])[starts[b] : starts[b] + n], zero-filled past the end of that waveform (Python's
        slice would stop there; ``valid[b]`` reports the number of real samples).  ``starts``: host ints."""
        vid = np.array([self.index[k] for k in names], dtype=np.int64)
        st = self.offsets[vid] + np.asarray(starts, dtype=np.int64)
        lim = self.offsets[vid + 1]
        B = len(vid)
        table = torch.from_numpy(np.stack([st, lim])).to(self.device, non_blocking=True)
        if out is None:
            out = torch.empty((B, n), dtype=torch.float32, device=self.device)
        L.call("asf_gather_clips", L.ptr(self.samples), L.ptr(table[0]), L.ptr(table[1]), B, n, L.ptr(out), out.stride(0),
               L.ptr(valid), L.stream_ptr())
        return out

    def views(self, name, clip_size, num_views, num_samples=None, start_sample=0, n=None):
        """All test views of one record: (num_views, n) device tensor, n = clip_size - 1 (the reference slices
        [start, start + clip_size - 1), audio_loader_epic.py:138 with end = start + clip_size - 1)."""
        total = int(self.offsets[self.index[name] + 1] - self.offsets[self.index[name]])
        ns = total if num_samples is None else int(num_samples)
        st = multi_view_starts([ns], clip_size, num_views, [start_sample])[0]
        return self.crop([name] * num_views, st, clip_size - 1 if n is None else n)


class WaveformRing:
    """``slots`` pinned host batches [B, n] (+ integer labels) with device twins; ``put`` fills a host slot and starts
    its H2D copy on the ring's copy stream, ``get`` makes the current stream wait for that copy and returns the device
    tensors.  With two or more slots the copy of the next batch runs under the current training step."""

    def __init__(self, batch, n_samples, slots=2, label_names=("verb", "noun"), device=None):
        L.require_cuda_sm100()
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.device, self.slots = dev, slots
        self.host_w = [torch.empty((batch, n_samples), dtype=torch.float32).pin_memory() for _ in range(slots)]
        self.dev_w = [torch.empty((batch, n_samples), dtype=torch.float32, device=dev) for _ in range(slots)]
        self.label_names = tuple(label_names)
        self.host_l = [{k: torch.empty(batch, dtype=torch.int64).pin_memory() for k in self.label_names} for _ in range(slots)]
        self.dev_l = [{k: torch.empty(batch, dtype=torch.int64, device=dev) for k in self.label_names} for _ in range(slots)]
        self.stream = torch.cuda.Stream(device=dev)
        self.ready = [None] * slots   # H2D copy of the slot finished
        self.free = [None] * slots    # the consumer is done with the slot's device tensors

    def put(self, slot, waveforms, labels):
        """``waveforms``: host array / tensor [B, n]; ``labels``: dict name -> host ints (or one array for a single
        head).  Returns immediately; the copy runs on the ring's stream."""
        if self.ready[slot] is not None:
            self.ready[slot].synchronize()  # the previous copy out of this pinned slot must have finished reading it
        self.host_w[slot].copy_(torch.as_tensor(waveforms))
        lab = labels if isinstance(labels, dict) else {self.label_names[0]: labels}
        for k in lab:
            self.host_l[slot][k].copy_(torch.as_tensor(lab[k]))
        with torch.cuda.stream(self.stream):
            if self.free[slot] is not None:
                self.stream.wait_event(self.free[slot])
            self.dev_w[slot].copy_(self.host_w[slot], non_blocking=True)
            for k in lab:
                self.dev_l[slot][k].copy_(self.host_l[slot][k], non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self.stream)
        self.ready[slot] = ev

    def get(self, slot):
        """Device tensors of ``slot`` once its copy is ordered before the current stream's next work."""
        cur = torch.cuda.current_stream(self.device)
        if self.ready[slot] is None:
            raise RuntimeError("WaveformRing.get(%d) before put" % slot)
        cur.wait_event(self.ready[slot])
        lab = self.dev_l[slot]
        return self.dev_w[slot], (lab if len(self.label_names) > 1 else lab[self.label_names[0]])

    def release(self, slot):
        """Marks the point on the current stream after which the slot's device tensors may be overwritten."""
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        self.free[slot] = ev

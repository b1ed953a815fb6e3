"""Data-parallel wrapper: one process per GPU, gradients all-reduced over NCCL (NVLink 5 / NVSwitch) in buckets
that are launched from INSIDE the fused backward pass, so communication overlaps the remaining dgrad / wgrad work.

Takes the place of ``torch.nn.parallel.DistributedDataParallel(module=model, device_ids=[dev], output_device=dev,
find_unused_parameters=True)`` in /root/reference/audio_slowfast/models/build.py:50-57.  Same surface for the
callers: ``model.module`` is the wrapped network (train_net.py:392, :710; checkpoint.py:111, :156), ``forward``
passes through, parameters are broadcast from rank 0 at construction, and after ``loss.backward()`` every
``p.grad`` holds the gradient averaged over ranks.

How it differs from DDP: the engine writes all parameter gradients into ONE flat fp32 buffer laid out in
reverse registration order (= the order in which the backward pass finishes them: head, s5, s4, ... stems), so
a "bucket" is just the next contiguous byte range of that buffer -- no gradient copies, no per-parameter
autograd hooks.  The engine calls ``grad_hook`` after each residual block's gradients are final; the bucketer
then all-reduces (mean) the finished prefix on a side stream whenever at least ``bucket_bytes`` are ready.
s5 + s4 hold ~92 % of the 107.5 MB of gradients and finish first, so almost all of the all-reduce hides under
the s3 / s2 / stem backward (SURVEY.md 5 "distributed communication backend").
"""
from __future__ import annotations

import torch
import torch.distributed as dist
import torch.nn as nn


class GradBucketer:
    """All-reduces (mean) a flat gradient buffer in contiguous buckets as prefixes of it become final.

    Device-agnostic on purpose (CPU tensors + gloo in the tests, CUDA tensors + NCCL in production): on CUDA the
    collectives run on ``comm_stream`` behind an event recorded on the producing stream."""

    def __init__(self, flat, group=None, bucket_bytes=16 << 20):
        self.flat = flat
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.bucket_elems = max(1, bucket_bytes // flat.element_size())
        self.launched = 0
        self.pending = []
        self.buckets = []  # (start, end) of every bucket of the last pass, for tests / logging
        self.comm_stream = torch.cuda.Stream(device=flat.device) if flat.is_cuda else None
        self.on_reduced = None  # callable(lo, hi), called on the communication stream right behind a bucket's all-reduce
        # gloo has no AVG: reduce with SUM and scale afterwards
        self.avg_op = dist.ReduceOp.AVG if flat.is_cuda else dist.ReduceOp.SUM

    def begin(self):
        self.launched = 0
        self.pending = []
        self.buckets = []

    def _launch(self, end):
        if end <= self.launched:
            return
        if self.world == 1:
            if self.on_reduced is not None:
                self.on_reduced(self.launched, end)
            self.launched = end
            return
        seg = self.flat[self.launched:end]
        self.buckets.append((self.launched, end))
        if self.comm_stream is not None:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(self.flat.device))
            with torch.cuda.stream(self.comm_stream):
                self.comm_stream.wait_event(ev)
                dist.all_reduce(seg, op=self.avg_op, group=self.group)
                if self.on_reduced is not None:
                    self.on_reduced(self.launched, end)
        else:
            w = dist.all_reduce(seg, op=self.avg_op, group=self.group, async_op=True)
            self.pending.append((w, seg))
        self.launched = end

    def ready(self, done):
        """``done``: number of leading elements of the flat buffer whose gradients are final."""
        if done - self.launched >= self.bucket_elems:
            self._launch(done)

    def finish(self):
        """Launches the tail bucket and makes the consumer (current stream / caller) wait for every bucket."""
        self._launch(self.flat.numel())
        if self.comm_stream is not None:
            torch.cuda.current_stream(self.flat.device).wait_stream(self.comm_stream)
        else:
            for w, seg in self.pending:
                w.wait()
                seg.div_(self.world)
            self.pending = []


class PeerExchange:
    """Sum-over-ranks of small double vectors through NVLink peer memory (``asf_peer_allreduce_f64``, csrc/peer.cu):
    the SyncBatchNorm statistics of every train-mode BatchNorm layer, forward and backward.  One symmetric buffer per
    rank (torch.distributed._symmetric_memory) holds, for every exchange key, 2 (epoch parity) x ``world`` slots of n
    doubles and ``world`` flags; the kernel pushes its values into its slot of every rank's buffer, signals the peers,
    waits for their signals and adds the slots of its own buffer in rank order.
    Construction is collective (every rank, same keys, same order)."""

    def __init__(self, group, device, layout):
        import torch.distributed._symmetric_memory as symm

        from . import _lib as L

        self.L = L
        self.group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)
        self.where = {}
        off = 0
        for key, n in layout:
            slot = off
            off += 2 * self.world * n * 8
            flag = off
            off += (self.world * 4 + 15) // 16 * 16
            self.where[key] = (len(self.where), n, slot, flag)
        self.buf = symm.empty(off, dtype=torch.uint8, device=device)
        self.buf.zero_()
        torch.cuda.synchronize(device)
        self.handle = symm.rendezvous(self.buf, self.group)
        self.ptrs = torch.tensor([int(p) for p in self.handle.buffer_ptrs], dtype=torch.int64, device=device)
        self.epochs = torch.zeros(len(self.where), dtype=torch.int32, device=device)
        torch.cuda.synchronize(device)
        dist.barrier(group=self.group)  # every rank's buffer is zeroed before anybody signals into it

    def allreduce(self, key, t):
        i, n, slot, flag = self.where[key]
        if t.numel() != n or t.dtype != torch.float64:
            raise ValueError("PeerExchange: %r expects %d doubles" % (key, n))
        L = self.L
        L.call("asf_peer_allreduce_f64", L.ptr(t), n, L.ptr(self.ptrs), self.rank, self.world, slot, flag,
               L.ptr(self.epochs[i:i + 1]), L.stream_ptr())


class DataParallel(nn.Module):
    """``DataParallel(model)``: data-parallel replica of an asf_b200 engine model (one process per GPU)."""

    def __init__(self, module, process_group=None, bucket_bytes=16 << 20, broadcast_parameters=True, sync_bn=True):
        """``sync_bn``: train-mode BatchNorm layers normalise with the statistics of the global batch (two small
        all-reduces per such layer and pass: sum(x), sum(x^2) forward; sum(dz), sum(dz * xhat) backward), which is
        what the reference gets from ``torch.nn.SyncBatchNorm.convert_sync_batchnorm`` (models/build.py:45-46).
        With the EPIC recipe (BN.FREEZE True) only the two stems and s1_fuse are in training mode."""
        super().__init__()
        if not dist.is_initialized():
            raise RuntimeError("asf_b200.parallel.DataParallel needs torch.distributed to be initialised "
                               "(one process per GPU, backend 'nccl')")
        self.module = module
        self.process_group = process_group
        self.bucket_bytes = bucket_bytes
        self._bucketer = None
        self._done = 0
        if broadcast_parameters:
            with torch.no_grad():
                for t in list(module.parameters()) + list(module.buffers()):
                    dist.broadcast(t, src=0, group=process_group)
        module.engine.grad_hook = self._on_grads
        world = dist.get_world_size(process_group)
        module.engine.sync_bn = (process_group, world) if (sync_bn and world > 1) else None

    def forward(self, x, bboxes=None):
        return self.module(x, bboxes)

    # called by the engine from inside backward: names = units whose parameter gradients are now final
    def _on_grads(self, engine, names):
        flat = engine._flat_grad
        if self._bucketer is None or self._bucketer.flat is not flat:
            self._bucketer = GradBucketer(flat, self.process_group, self.bucket_bytes)
        b = self._bucketer
        b.on_reduced = (lambda lo, hi: engine.update_hook(engine, lo, hi)) if engine.update_hook is not None else None
        if names and names[0] == "head":
            b.begin()
            self._done = 0
            for n, p in self.module.named_parameters():
                if p.requires_grad and p.grad is not None and p.grad.data_ptr() != engine._grad_views[n].data_ptr():
                    raise RuntimeError("DataParallel: %s.grad is not the engine's gradient buffer; use "
                                       "optimizer.zero_grad(set_to_none=True) or leave .grad untouched" % n)
        self._done = max(self._done, engine.grad_prefix_end(names))
        if names and names[-1] == "stems":
            b.finish()
        else:
            b.ready(self._done)

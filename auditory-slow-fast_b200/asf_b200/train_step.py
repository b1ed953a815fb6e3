"""One training iteration of the hot path as a replayable CUDA graph.

What ``train_epoch`` does per batch in the reference (/root/reference/audio_slowfast/tools/train_net.py:414-493:
host->device copy, ``model(x=inputs)``, ``compute_loss`` = mean of the verb and noun cross-entropies
(tools/train_utils.py:76-101), ``optimizer.zero_grad(); loss.backward(); optimizer.step()``) -- preceded here by the
batched log-mel extraction that the reference runs per sample in DataLoader workers -- is recorded ONCE into a CUDA
graph over static buffers and replayed every iteration: ~900 kernel launches become one ``cudaGraphLaunch``, which is
what keeps the GPU busy once the kernels themselves take only a few milliseconds.

    step = TrainStep(cfg, model, optimizer, batch_size, n_samples)
    loss = step(waveforms, labels)            # device or pinned-host tensors; returns the (device) loss scalar

Learning-rate schedules keep working: ``optimizer.param_groups[i]["lr"]`` is published to the device before each
replay (asf_b200.optim.FusedSGD).  ``graph=False`` runs the same calls eagerly.
"""
from __future__ import annotations

import os

import torch
import torch.nn.functional as F

from .datasets.audio_loader_epic import logmel_batch


_CE_TICKETS = {}


def _ce_ticket(device):
    """Zero-initialised, self-resetting arrival counter of asf_softmax_ce (one per device and stream: launches that
    share a ticket must be stream-ordered)."""
    key = (device.index, torch.cuda.current_stream(device).cuda_stream)
    t = _CE_TICKETS.get(key)
    if t is None:
        t = _CE_TICKETS[key] = torch.zeros(1, dtype=torch.int32, device=device)
    return t


class _FusedCE(torch.autograd.Function):
    """Both cross-entropies, their mean and d loss / d logits in ONE launch (asf_softmax_ce, csrc/head.cu) instead of
    the ~25 small launches of two F.cross_entropy calls and their autograd backward."""

    @staticmethod
    def forward(ctx, labels, *logits):
        from . import _lib as L
        heads, B = len(logits), logits[0].shape[0]
        logits = [lg if lg.stride(-1) == 1 else lg.contiguous() for lg in logits]
        loss = torch.empty((), dtype=torch.float32, device=logits[0].device)
        parts = torch.empty(2, dtype=torch.float32, device=logits[0].device)
        need = [ctx.needs_input_grad[1 + h] for h in range(heads)]
        dl = [torch.empty_like(lg, memory_format=torch.contiguous_format) if n else None for lg, n in zip(logits, need)]
        # (held in variables until the call has been issued: a temporary would be freed, and its block handed to the next
        # allocation, while its address is still on its way into the launch)
        scratch = torch.empty(heads * B, dtype=torch.float32, device=loss.device)
        ticket = _ce_ticket(loss.device)
        L.call("asf_softmax_ce", heads, L.ptr_array(logits), L.i64_array([lg.stride(0) for lg in logits]),
               L.int_array([lg.shape[1] for lg in logits]), L.ptr_array(list(labels)), B, L.ptr(loss), L.ptr(parts),
               L.ptr_array(dl),
               L.i64_array([d.stride(0) if d is not None else 0 for d in dl]),
               L.ptr(scratch), L.ptr(ticket), L.stream_ptr())
        ctx.dl = dl
        ctx.mark_non_differentiable(parts)
        return loss, parts

    @staticmethod
    def backward(ctx, go, _):
        return (None,) + tuple(None if d is None else d * go for d in ctx.dl)


def _fusable(preds, labels):
    ts = list(preds) if isinstance(preds, (tuple, list)) else [preds]
    ls = [labels["verb"], labels["noun"]] if isinstance(preds, (tuple, list)) else [labels]
    ok = all(t.is_cuda and t.dtype == torch.float32 and t.dim() == 2 for t in ts) and all(
        l.is_cuda and l.dtype == torch.int64 and l.is_contiguous() for l in ls) and len(ts) <= 2
    return ok, ts, ls


def compute_loss(preds, labels):
    """tools/train_utils.py:76-101: two heads -> mean of the verb and noun CE; one head -> CE."""
    ok, ts, ls = _fusable(preds, labels)
    if ok:
        loss, parts = _FusedCE.apply(tuple(ls), *ts)
        return loss, parts[0], parts[1 if len(ts) > 1 else 0]
    if isinstance(preds, (tuple, list)):
        lv = F.cross_entropy(preds[0], labels["verb"])
        ln = F.cross_entropy(preds[1], labels["noun"])
        return 0.5 * (lv + ln), lv, ln
    loss = F.cross_entropy(preds, labels)
    return loss, loss, loss


class TrainStep:
    def __init__(self, cfg, model, optimizer, batch_size, n_samples, graph=True, warmup=3, augment=None,
                 pipeline=False, segmented_update=False):
        """``augment``: optional ``datasets.spec_augment.SpecAugmenter`` (the reference's training-time SpecAugment,
        datasets/epickitchens.py:160-168) applied on the device between extraction and the stems; its random numbers
        come from a device-side counter, so every replay of the captured step augments differently.

        ``segmented_update`` (FusedSGD only): parameters are updated piece by piece from inside the backward pass, as
        soon as a piece of the flat gradient buffer is final (and, data parallel, reduced), instead of by one launch
        after it.  Same arithmetic per parameter, so the result is bit-identical.

        ``pipeline``: the log-mel extraction (+ SpecAugment) of batch k+1 runs inside step k, beside the end of its
        backward pass, the way the reference's DataLoader workers extract the next batches while the GPU trains
        (datasets/epickitchens.py:160-170 runs in the workers).  Needs the ``prefetch()`` / ``step()`` protocol: the
        batch after the one being trained must have been handed over before ``step()`` is called, otherwise that
        step extracts its own batch up front (the first and the last step of a run)."""
        self.cfg, self.model, self.opt = cfg, model, optimizer
        net = model.module if hasattr(model, "module") else model
        dev = next(net.parameters()).device
        self.device = dev
        self.two_heads = len(cfg.MODEL.NUM_CLASSES) > 1
        self.wave = torch.zeros((batch_size, n_samples), dtype=torch.float32, device=dev)
        if self.two_heads:
            self.labels = {"verb": torch.zeros(batch_size, dtype=torch.int64, device=dev),
                           "noun": torch.zeros(batch_size, dtype=torch.int64, device=dev)}
        else:
            self.labels = torch.zeros(batch_size, dtype=torch.int64, device=dev)
        self.loss = torch.zeros((), dtype=torch.float32, device=dev)
        self.augment = augment
        self.pipeline = bool(pipeline)
        self.segmented = bool(segmented_update) and hasattr(optimizer, "step_range")
        self._x_cur = self._x_next = None
        self._x_next_holds = None   # id of the staged batch whose spectrogram is in _x_next
        self._batch_ids = 0
        self.graph = None
        self._want_graph = graph
        self._warmup = warmup
        self._calls = 0

    def _extract(self):
        return logmel_batch(self.cfg, self.wave, augment=self.augment)

    def _extract_next(self):
        """pipeline mode: spectrograms of the waveforms in ``self.wave`` -> the ``_x_next`` buffers (current stream)."""
        x = self._extract()
        if self._x_next is None:
            self._x_next = [torch.empty_like(t) for t in x]
            self._x_cur = [torch.empty_like(t) for t in x]
        for dst, src in zip(self._x_next, x):
            dst.copy_(src)

    def _body(self):
        net = self.model.module if hasattr(self.model, "module") else self.model
        cur = torch.cuda.current_stream(self.device)
        eng = getattr(net, "engine", None)
        fork = torch.cuda.Event()
        fork.record(cur)
        late = None
        if self.pipeline:
            # this batch was extracted during the previous step (or up front by __call__)
            for dst, src in zip(self._x_cur, self._x_next):
                dst.copy_(src)
            x = self._x_cur
            taken = torch.cuda.Event()
            taken.record(cur)
            late = {"ev": None}
            if eng is not None:
                after = os.environ.get("ASF_EXTRACT_AFTER", "s2_fuse")

                def stage_hook(engine, name, late=late, after=after):
                    if name == after and late["ev"] is None:
                        late["ev"] = torch.cuda.Event()
                        late["ev"].record(torch.cuda.current_stream(self.device))
                eng.stage_hook = stage_hook
        else:
            x = self._extract()
        if eng is not None:
            # the optimiser step of the previous iteration left every packed weight copy stale: refresh them on a side
            # stream forked BEFORE the log-mel kernel but launched after it, so the HBM-bound re-layout runs beside the
            # tensor-core kernel (the forward joins that stream before its first launch)
            eng.refresh_weights(net.training, after=fork)
        self.opt.zero_grad(set_to_none=True)
        seg = self.segmented and eng is not None
        if seg:
            self.opt.begin_step(self.device)
            eng.update_hook = lambda engine, lo, hi: self.opt.step_range(engine._flat_grad, lo, hi)
        try:
            preds = self.model(x)
            loss, _, _ = compute_loss(preds, self.labels)
            loss.backward()
        finally:
            if eng is not None:
                eng.update_hook = None
                eng.stage_hook = None
        self.opt.step()
        self.loss.copy_(loss.detach())
        if self.pipeline:
            # the next batch's extraction: a branch of the step that starts once the backward pass has reached the
            # chosen stage boundary (and the spectrogram buffers have been handed to this step) and joins at the end
            if getattr(self, "_ext_stream", None) is None:
                self._ext_stream = torch.cuda.Stream(device=self.device)
            ext = self._ext_stream
            ext.wait_event(taken)
            if late["ev"] is not None:
                ext.wait_event(late["ev"])
            with torch.cuda.stream(ext):
                self._extract_next()
            cur.wait_stream(ext)

    def prefetch(self, waveforms, labels):
        """Starts the host->device copy of the NEXT batch (pinned host tensors) into a staging slot on a copy stream and
        returns at once: the copy runs under the step that is executing.  The next ``step()`` call without arguments
        consumes it (device-to-device into the static buffers the graph reads, 12 MB at HBM speed).  Two slots."""
        if not hasattr(self, "_stage"):
            self._copy_stream = torch.cuda.Stream(device=self.device)
            mk = lambda t: torch.empty_like(t)
            self._stage = [(mk(self.wave), ({k: mk(v) for k, v in self.labels.items()} if self.two_heads else mk(self.labels)))
                           for _ in range(2)]
            self._stage_ready = [None, None]
            self._stage_free = [None, None]
            self._stage_next = 0
            self._stage_pending = []
        if len(self._stage_pending) >= 2:
            raise RuntimeError("TrainStep.prefetch: both staging slots hold batches that have not been trained on yet")
        k = self._stage_next
        self._stage_next ^= 1
        self._batch_ids += 1
        w, lab = self._stage[k]
        with torch.cuda.stream(self._copy_stream):
            if self._stage_free[k] is not None:
                self._copy_stream.wait_event(self._stage_free[k])
            w.copy_(waveforms, non_blocking=True)
            if self.two_heads:
                lab["verb"].copy_(labels["verb"], non_blocking=True)
                lab["noun"].copy_(labels["noun"], non_blocking=True)
            else:
                lab.copy_(labels, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self._copy_stream)
        self._stage_ready[k] = ev
        self._stage_pending.append((k, self._batch_ids))

    def _consume_prefetched(self):
        k, _ = self._stage_pending.pop(0)
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(self._stage_ready[k])
        w, lab = self._stage[k]
        self.load(w, lab)
        ev = torch.cuda.Event()
        ev.record(cur)
        self._stage_free[k] = ev

    def _load_labels(self, labels):
        if self.two_heads:
            self.labels["verb"].copy_(labels["verb"], non_blocking=True)
            self.labels["noun"].copy_(labels["noun"], non_blocking=True)
        else:
            self.labels.copy_(labels, non_blocking=True)

    def _consume_pipelined(self):
        """pipeline mode, before the step runs: labels of the batch to train on -> static buffers; its spectrogram is
        in ``_x_next`` already unless nobody announced it in time (then: extract it now, on the current stream);
        waveforms of the batch AFTER it (if announced) -> ``self.wave`` for the extraction branch of this step."""
        if not getattr(self, "_stage_pending", None):
            raise RuntimeError("TrainStep(pipeline=True): call prefetch(waveforms, labels) before step()")
        cur = torch.cuda.current_stream(self.device)
        k, bid = self._stage_pending.pop(0)
        cur.wait_event(self._stage_ready[k])
        w, lab = self._stage[k]
        if self._x_next_holds != bid:
            self.wave.copy_(w, non_blocking=True)
            self._extract_next()
        self._load_labels(lab)
        if self._stage_pending:
            k2, bid2 = self._stage_pending[0]
            cur.wait_event(self._stage_ready[k2])
            self.wave.copy_(self._stage[k2][0], non_blocking=True)
            self._x_next_holds = bid2
        else:
            self._x_next_holds = None  # the extraction branch re-extracts stale waveforms; nobody reads the result
        ev = torch.cuda.Event()
        ev.record(cur)
        self._stage_free[k] = ev

    def load(self, waveforms, labels):
        """Copies one batch (device tensors, or pinned host tensors -> async H2D) into the static input buffers."""
        self.wave.copy_(waveforms, non_blocking=True)
        if self.two_heads:
            self.labels["verb"].copy_(labels["verb"], non_blocking=True)
            self.labels["noun"].copy_(labels["noun"], non_blocking=True)
        else:
            self.labels.copy_(labels, non_blocking=True)

    def _snapshot(self):
        """Everything a training iteration mutates: parameters, buffers (BN running statistics, num_batches_tracked),
        optimiser state and the CUDA RNG (dropout)."""
        net = self.model.module if hasattr(self.model, "module") else self.model
        snap = {"params": [p.detach().clone() for p in net.parameters()],
                "buffers": [b.detach().clone() for b in net.buffers()],
                "rng": torch.cuda.get_rng_state(self.device),
                "opt": {id(p): {k: (v.detach().clone() if torch.is_tensor(v) else v) for k, v in st.items()}
                        for p, st in self.opt.state.items()},
                "opt_steps": getattr(self.opt, "_steps", None)}
        return snap

    def _restore(self, snap):
        net = self.model.module if hasattr(self.model, "module") else self.model
        with torch.no_grad():
            for p, q in zip(net.parameters(), snap["params"]):
                p.copy_(q)
            for b, q in zip(net.buffers(), snap["buffers"]):
                b.copy_(q)
            for p, st in self.opt.state.items():
                old = snap["opt"].get(id(p))
                for k, v in st.items():
                    if torch.is_tensor(v):
                        # state created by the warm-up (momentum buffers) goes back to its initial value: zero
                        v.copy_(old[k]) if old is not None and k in old else v.zero_()
                    elif old is not None and k in old:
                        st[k] = old[k]
        if snap["opt_steps"] is not None:
            self.opt._steps = snap["opt_steps"]
        torch.cuda.set_rng_state(snap["rng"], self.device)
        if hasattr(net, "engine"):
            net.engine.invalidate_weights()

    def _capture(self):
        """Capture of the whole iteration.  torch's capture rules need a few eager warm-up runs on a side stream first
        (allocations, lazy plans, cudaFuncSetAttribute); those runs are REAL iterations, so everything they mutate --
        parameters, momentum buffers, BN running statistics / num_batches_tracked, the dropout RNG -- is restored
        afterwards: the first replay is the first update, exactly one per batch like the reference's train_epoch
        (tools/train_net.py:472-493)."""
        snap = self._snapshot()
        # pipeline mode: the warm-up runs overwrite the hand-over buffers with the NEXT batch's spectrograms
        x_next = [t.clone() for t in self._x_next] if self.pipeline else None
        s = torch.cuda.Stream(device=self.device)
        s.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(s):
            for _ in range(max(1, self._warmup)):
                self._body()
        torch.cuda.current_stream(self.device).wait_stream(s)
        torch.cuda.synchronize(self.device)
        self._restore(snap)
        if x_next is not None:
            for dst, src in zip(self._x_next, x_next):
                dst.copy_(src)
        net = self.model.module if hasattr(self.model, "module") else self.model
        if hasattr(net, "engine"):
            # constants (folded affine of the frozen BatchNorm layers) are derived once, here; only the re-pack of the
            # weights that the optimiser changes every step goes into the graph
            net.engine.prepare_weights(net.training)
            net.engine.invalidate_weights(folds=False)
        torch.cuda.synchronize(self.device)
        if hasattr(self.opt, "refresh_lr"):
            self.opt.refresh_lr()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self._body()
        self.graph = g

    def __call__(self, waveforms=None, labels=None):
        if self.pipeline:
            if waveforms is not None:
                self.prefetch(waveforms, labels)
            self._consume_pipelined()
        elif waveforms is not None:
            self.load(waveforms, labels)
        elif getattr(self, "_stage_pending", None):
            self._consume_prefetched()
        if not self._want_graph:
            self._body()
            return self.loss
        if self.graph is None:
            self._capture()
        if hasattr(self.opt, "refresh_lr"):
            self.opt.refresh_lr()
        self.graph.replay()
        self._calls += 1
        # the replayed optimiser step wrote the fp32 masters without running the Python that bumps autograd's version
        # counters: an eager forward between replays (validation) must not reuse packed copies from before the update
        net = self.model.module if hasattr(self.model, "module") else self.model
        if hasattr(net, "engine"):
            net.engine.invalidate_weights()
        return self.loss

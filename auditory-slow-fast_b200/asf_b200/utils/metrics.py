"""Accuracy metrics on the device, behind the reference's function names.

Mirror of /root/reference/audio_slowfast/utils/metrics.py:11-207 (``topks_correct``, ``multitask_topks_correct``,
``topk_errors``, ``topk_accuracies``, ``multitask_topk_accuracies``).  Same arguments, same meaning of the results;
the results stay ON THE DEVICE (0-dim tensors), so the training loop's nine ``.item()`` synchronisations per iteration
(tools/train_net.py:497-539) become none: accumulate with ``TrainMetrics`` and read once per logging period.
One launch of ``asf_topk_counts`` (one warp per sample, rank of the label's score) replaces topk + transpose + eq + sum.
"""
from __future__ import annotations

import torch

from .. import _lib as L


def _prep(preds, labels):
    if not preds.is_cuda:
        raise RuntimeError("asf_b200.utils.metrics works on CUDA tensors (there is no CPU path)")
    if preds.dim() != 2:
        raise ValueError("predictions must be (batch, classes)")
    assert preds.size(0) == labels.size(0), "Batch dim of predictions and labels must match"
    p = preds.detach()
    if p.dtype != torch.float32:
        p = p.float()
    if p.stride(1) != 1:
        p = p.contiguous()
    lab = labels.detach().to(device=p.device, dtype=torch.int64).contiguous()
    return p, lab


def topk_counts(preds, labels, ks=(1, 5), out=None):
    """Counts of top-k hits.  ``preds`` / ``labels``: one task (tensor, tensor) or two tasks (tuples of two).
    Returns a device int64 tensor: one task -> [hits@k1, hits@k2]; two tasks -> [a@k1, a@k2, b@k1, b@k2, both@k1, both@k2].
    ``out`` (int64[6], device): accumulate into it instead of a fresh zero buffer."""
    L.require_cuda_sm100()
    k1, k2 = (int(ks[0]), int(ks[-1]))
    two = isinstance(preds, (tuple, list))
    pa, la = _prep(preds[0], labels[0]) if two else _prep(preds, labels)
    if out is None:
        out = torch.zeros(6, dtype=torch.int64, device=pa.device)
    if two:
        if len(preds) != 2:
            raise NotImplementedError("asf_topk_counts fuses at most two tasks (verb, noun)")
        pb, lb = _prep(preds[1], labels[1])
        L.call("asf_topk_counts", L.ptr(pa), pa.size(1), pa.stride(0), L.ptr(la), L.ptr(pb), pb.size(1), pb.stride(0),
               L.ptr(lb), pa.size(0), k1, k2, L.ptr(out), L.stream_ptr())
    else:
        L.call("asf_topk_counts", L.ptr(pa), pa.size(1), pa.stride(0), L.ptr(la), None, 0, 0, None, pa.size(0), k1, k2,
               L.ptr(out), L.stream_ptr())
    return out


def topks_correct(preds, labels, ks):
    """metrics.py:11-41: number of correct predictions for each k (device float tensors, like the reference's)."""
    out = []
    for i in range(0, len(ks), 2):
        pair = ks[i:i + 2]
        c = topk_counts(preds, labels, (pair[0], pair[-1]))
        out.extend([c[j].float() for j in range(len(pair))])
    return out


def multitask_topks_correct(preds, labels, ks=(1,)):
    """metrics.py:80-108: samples whose labels are inside the top-k of EVERY task."""
    out = []
    for i in range(0, len(ks), 2):
        pair = ks[i:i + 2]
        c = topk_counts(tuple(preds), tuple(labels), (pair[0], pair[-1]))
        out.extend([c[4 + j].float() for j in range(len(pair))])
    return out


def topk_errors(preds, labels, ks):
    return [(1.0 - x / preds.size(0)) * 100.0 for x in topks_correct(preds, labels, ks)]


def topk_accuracies(preds, labels, ks):
    return [(x / preds.size(0)) * 100.0 for x in topks_correct(preds, labels, ks)]


def multitask_topk_accuracies(preds, labels, ks):
    return [(x / preds[0].size(0)) * 100.0 for x in multitask_topks_correct(preds, labels, ks)]


class TrainMetrics:
    """Running sums of one logging period, on the device: top-1 / top-5 hits of verb, noun and action, the three losses
    and the sample count.  ``update`` launches one kernel + one tiny add and never synchronises; ``summary`` is the one
    read-back (and, under data parallelism, ONE all-reduce of 10 numbers instead of the reference's three per
    iteration, tools/train_net.py:501-531)."""

    def __init__(self, device):
        self.counts = torch.zeros(6, dtype=torch.int64, device=device)
        self.sums = torch.zeros(4, dtype=torch.float64, device=device)  # loss, loss_verb, loss_noun, samples

    def reset(self):
        self.counts.zero_()
        self.sums.zero_()

    def update(self, preds, labels, loss, loss_verb=None, loss_noun=None):
        two = isinstance(preds, (tuple, list))
        if two:
            topk_counts(tuple(preds), (labels["verb"], labels["noun"]), (1, 5), out=self.counts)
            n = preds[0].size(0)
        else:
            topk_counts(preds, labels, (1, 5), out=self.counts)
            n = preds.size(0)
        vals = [loss, loss_verb if loss_verb is not None else loss, loss_noun if loss_noun is not None else loss]
        self.sums[:3] += torch.stack([v.detach().double() for v in vals]) * n
        self.sums[3] += n

    def summary(self, group=None):
        import torch.distributed as dist

        buf = torch.cat([self.counts.double(), self.sums])
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
            dist.all_reduce(buf, group=group)
        v = buf.cpu().tolist()  # the one synchronisation of the period
        n = max(v[9], 1.0)
        keys = ("verb_top1_acc", "verb_top5_acc", "noun_top1_acc", "noun_top5_acc", "top1_acc", "top5_acc")
        out = {k: 100.0 * c / n for k, c in zip(keys, v[:6])}
        out.update({"loss": v[6] / n, "verb_loss": v[7] / n, "noun_loss": v[8] / n, "samples": int(v[9])})
        return out

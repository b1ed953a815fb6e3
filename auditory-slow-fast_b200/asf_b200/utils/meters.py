"""Multi-view test ensemble on the device.

Mirror of ``EPICTestMeter`` (/root/reference/audio_slowfast/utils/meters.py:1560-1755: ``update_stats`` :1641-1681,
``finalize_metrics`` :1712-1755) as ``perform_test`` drives it (tools/test_net.py:194-239).  The reference copies every
batch of predictions to the host and walks it sample by sample in Python; here the scores of every (video, view) are
scattered into device buffers by one launch per head, the views are added (or maxed) in view order by one launch at the
end, and the verb / noun / action top-k accuracies come out of ``asf_topk_counts`` -- the only host synchronisation is
reading the six counters in ``finalize_metrics``.
"""
from __future__ import annotations

import numpy as np
import torch

from .. import _lib as L
from . import metrics


class EPICTestMeter:
    def __init__(self, num_audios, num_clips, num_cls, overall_iters=0, ensemble_method="sum", device=None):
        if ensemble_method not in ("sum", "max"):
            raise NotImplementedError("Ensemble Method {} is not supported".format(ensemble_method))
        L.require_cuda_sm100()
        self.num_audios, self.num_clips, self.num_cls = int(num_audios), int(num_clips), tuple(num_cls)
        self.overall_iters = overall_iters
        self.ensemble_method = ensemble_method
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.device = dev
        self.verb_audio_preds_clips = torch.zeros((num_audios, num_clips, num_cls[0]), device=dev)
        self.noun_audio_preds_clips = torch.zeros((num_audios, num_clips, num_cls[1]), device=dev)
        self.verb_audio_labels = torch.zeros(num_audios, dtype=torch.int64, device=dev)
        self.noun_audio_labels = torch.zeros(num_audios, dtype=torch.int64, device=dev)
        self._seen = torch.zeros((num_audios, num_clips), dtype=torch.int32, device=dev)
        self.metadata = np.zeros(num_audios, dtype=object)
        self._meta_pending = []
        self.stats = {}

    def reset(self):
        for t in (self.verb_audio_preds_clips, self.noun_audio_preds_clips, self.verb_audio_labels,
                  self.noun_audio_labels, self._seen):
            t.zero_()
        self.metadata.fill(0)
        self._meta_pending = []

    def update_stats(self, preds, labels, metadata, clip_ids):
        """preds: (verb (N, Cv), noun (N, Cn)) device tensors; labels: (verb (N,), noun (N,)); clip_ids (N,): video
        index * num_clips + view (meters.py:1655-1657).  No host synchronisation."""
        ids = clip_ids.detach().to(device=self.device, dtype=torch.int64).contiguous()
        st = L.stream_ptr()
        for p, lab, clips, vlab in ((preds[0], labels[0], self.verb_audio_preds_clips, self.verb_audio_labels),
                                    (preds[1], labels[1], self.noun_audio_preds_clips, self.noun_audio_labels)):
            p, lab = metrics._prep(p, lab)
            L.call("asf_ensemble_scatter", L.ptr(p), p.size(1), p.stride(0), L.ptr(lab), L.ptr(ids), p.size(0),
                   self.num_clips, self.num_audios, L.ptr(clips), L.ptr(vlab), L.ptr(self._seen), st)
        if metadata is not None:
            self._meta_pending.append((ids, list(metadata["narration_id"])))  # resolved at finalize (needs the ids on the host)

    @property
    def clip_count(self):
        return self._seen.sum(1)

    def _ensemble(self, clips):
        ens = torch.empty((self.num_audios, clips.shape[2]), device=self.device)
        L.call("asf_ensemble_reduce", L.ptr(clips), clips.shape[2], self.num_clips, self.num_audios,
               0 if self.ensemble_method == "sum" else 1, L.ptr(ens), L.stream_ptr())
        return ens

    def finalize_metrics(self, ks=(1, 5)):
        """meters.py:1712-1755: ensembled verb / noun / action top-k accuracies; returns the same 4-tuple (numpy)."""
        self.verb_audio_preds = self._ensemble(self.verb_audio_preds_clips)
        self.noun_audio_preds = self._ensemble(self.noun_audio_preds_clips)
        counts = metrics.topk_counts((self.verb_audio_preds, self.noun_audio_preds),
                                     (self.verb_audio_labels, self.noun_audio_labels), ks)
        c = counts.cpu().tolist()  # the one synchronisation of the test run
        n = float(self.num_audios)
        self.stats = {"split": "test_final"}
        for name, off in (("verb", 0), ("noun", 2), ("action", 4)):
            for j, k in enumerate(ks[:2]):
                self.stats["{}_top{}_acc".format(name, k)] = "{:.{prec}f}".format(100.0 * c[off + j] / n, prec=2)
        self.complete = bool((self.clip_count == self.num_clips).all())
        for ids, names in self._meta_pending:
            for i, nm in zip((ids // self.num_clips).cpu().tolist(), names):
                self.metadata[i] = nm
        self._meta_pending = []
        return ((self.verb_audio_preds.cpu().numpy().copy(), self.noun_audio_preds.cpu().numpy().copy()),
                (self.verb_audio_preds_clips.cpu().numpy().copy(), self.noun_audio_preds_clips.cpu().numpy().copy()),
                (self.verb_audio_labels.cpu().numpy().copy(), self.noun_audio_labels.cpu().numpy().copy()),
                self.metadata.copy())

"""Norm-layer factory (mirror of /root/reference/audio_slowfast/models/batchnorm_helper.py:16-31).

Only ``BN.NORM_TYPE == "batchnorm"`` is on the hot path (every shipped yaml uses the default);
``sub_batchnorm`` / ``sync_batchnorm`` are never enabled by a shipped config and raise instead of
silently running something else.
"""
import torch.nn as nn


def get_norm(cfg):
    kind = cfg.BN.NORM_TYPE
    if kind == "batchnorm":
        return nn.BatchNorm2d
    if kind in ("sub_batchnorm", "sync_batchnorm"):
        raise NotImplementedError(
            "Norm type {} is outside the B200 hot path (SURVEY.md 2.1 row 6); use 'batchnorm'".format(kind))
    raise NotImplementedError("Norm type {} is not supported".format(kind))

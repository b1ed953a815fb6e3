"""Precise BatchNorm statistics.

Mirror of /root/reference/audio_slowfast/utils/bn_helper.py:12-74 (``compute_and_update_bn_stats``) and its caller
``calculate_and_update_precise_bn`` (tools/train_net.py:659-680): with the weights fixed, run ``num_batches`` training-
mode forwards with BN momentum 1.0 (so each forward leaves THAT batch's statistics in running_mean / running_var --
written by the engine's train-mode BN kernels), average E[x] and E[x^2] over the batches, and store mean and
E[x^2] - mean^2 as the running statistics.  The per-layer Python arithmetic of the reference (2 x 110 tiny tensor ops per
batch) is done with multi-tensor (``torch._foreach``) calls: a handful of launches per batch, no host synchronisation.
Only layers that are in training mode are recomputed, exactly like the reference (frozen / eval layers do not update
their running statistics in a forward pass, so the reference's averaging leaves them where they were).
"""
from __future__ import annotations

import itertools

import torch


@torch.no_grad()
def compute_and_update_bn_stats(model, data_loader, num_batches=200):
    """``data_loader`` yields ``(inputs, labels, index)`` like the reference's loaders, or bare ``inputs``."""
    bn_layers = [m for m in model.modules()
                 if isinstance(m, (torch.nn.BatchNorm1d, torch.nn.BatchNorm2d, torch.nn.BatchNorm3d))]
    if not bn_layers:
        return 0
    momentum_actual = [bn.momentum for bn in bn_layers]
    for bn in bn_layers:
        bn.momentum = 1.0
    running_mean = [torch.zeros_like(bn.running_mean) for bn in bn_layers]
    running_square_mean = [torch.zeros_like(bn.running_var) for bn in bn_layers]
    rms = [bn.running_mean for bn in bn_layers]
    rvs = [bn.running_var for bn in bn_layers]
    done = 0
    try:
        for ind, batch in enumerate(itertools.islice(data_loader, num_batches)):
            inputs = batch[0] if isinstance(batch, tuple) else batch  # (inputs, labels, index) or bare inputs
            if isinstance(inputs, (list,)):
                inputs = [t.float().cuda(non_blocking=True) for t in inputs]
            else:
                inputs = inputs.cuda(non_blocking=True)
            model(inputs)
            # running_mean += (bn.running_mean - running_mean) / (ind + 1), for all layers at once
            d = torch._foreach_sub(rms, running_mean)
            torch._foreach_add_(running_mean, d, alpha=1.0 / (ind + 1))
            sq = torch._foreach_mul(rms, rms)
            torch._foreach_add_(sq, rvs)                      # E[x^2] = Var(x) + E[x]^2
            torch._foreach_sub_(sq, running_square_mean)
            torch._foreach_add_(running_square_mean, sq, alpha=1.0 / (ind + 1))
            done += 1
    finally:
        for bn, mom in zip(bn_layers, momentum_actual):
            bn.momentum = mom
    if done:
        m2 = torch._foreach_mul(running_mean, running_mean)
        var = torch._foreach_sub(running_square_mean, m2)
        for bn, m, v in zip(bn_layers, running_mean, var):
            bn.running_mean.copy_(m)
            bn.running_var.copy_(v)
    return done


def calculate_and_update_precise_bn(loader, model, num_iters=200, use_gpu=True):
    """tools/train_net.py:659-680."""
    if not use_gpu:
        raise RuntimeError("asf_b200 models run on the GPU only")

    def _gen_loader():
        for inputs, *_ in loader:
            yield inputs, None, None

    return compute_and_update_bn_stats(model, _gen_loader(), num_iters)

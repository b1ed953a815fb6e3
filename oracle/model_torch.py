"""TEST INFRASTRUCTURE ONLY -- plain-PyTorch fp32 restatement of the reference network.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference``
legs may import this module; the product package never does.

A *functional* restatement (state_dict in, tensors out; F.conv2d / F.batch_norm / F.max_pool2d /
F.avg_pool2d / F.linear on NCHW fp32) of

  AudioSlowFast            /root/reference/audio_slowfast/models/audio_model_builder.py:115-345
  ResNet (slow / fast)     audio_model_builder.py:624-796
  FuseFastToSlow           audio_model_builder.py:54-111
  ResNetBasicStem          /root/reference/audio_slowfast/models/stem_helper.py:98-168
  BottleneckTransform      /root/reference/audio_slowfast/models/resnet_helper.py:101-225
  ResBlock / ResStage      resnet_helper.py:228-341, :344-485
  ResNetBasicHead          /root/reference/audio_slowfast/models/head_helper.py:10-109
  init_weights             /root/reference/audio_slowfast/utils/weight_init_helper.py:10-39
  freeze_fn                audio_model_builder.py:327-344
  compute_loss             /root/reference/audio_slowfast/tools/train_utils.py:76-101

Pinned in this container against the *real* reference classes imported through
``oracle/ref_shim.py`` (tests/test_oracle_model.py: state_dict keys/shapes/values bit-exact for the
same seed, outputs and gradients equal to float rounding) and against the committed golden
vectors tests/golden/model_*.npz that ``tests/golden/make_golden.py`` produced by running the
reference itself.
"""
from __future__ import annotations

from collections import OrderedDict
from types import SimpleNamespace

import torch
import torch.nn as nn
import torch.nn.functional as F

STAGE_DEPTH = {50: (3, 4, 6, 3), 101: (3, 4, 23, 3)}
# audio_model_builder.py:23-45 (temporal kernel of conv1, res2..res5 per pathway)
TEMP_KERNEL = {
    "slow": [[1], [1], [1], [3], [3]],
    "fast": [[5], [3], [3], [3], [3]],
    "slowfast": [[1, 5], [1, 3], [1, 3], [3, 3], [3, 3]],
}
UNFROZEN_BN = ("s1.pathway0_stem.bn", "s1.pathway1_stem.bn", "s1_fuse.bn")
BN_EPS, BN_MOM = 1e-5, 0.1


def epic_cfg(arch="slowfast", num_classes=(97, 300), bn_freeze=True, dropout=0.5, zero_init_final_bn=True):
    """Plain-namespace equivalent of models/asf/config/asf-original-augment.yaml over defaults.py."""
    return SimpleNamespace(
        arch=arch, num_classes=list(num_classes), depth=50, width=64, alpha=4, beta_inv=8,
        fusion_ratio=2, fusion_kernel=7, num_frames=400, num_freq=128, dropout=dropout,
        bn_freeze=bn_freeze, zero_init_final_bn=zero_init_final_bn, fc_init_std=0.01,
        freq_strides=[1, 2, 2, 2], head_act="softmax",
    )


def network_spec(cfg):
    """List of (name, kind, meta) in module-registration order of the reference
    (``_construct_network``): the order that fixes both RNG consumption and state_dict order."""
    npath = 2 if cfg.arch == "slowfast" else 1
    tk = TEMP_KERNEL[cfg.arch]
    w = cfg.width
    if npath == 2:
        dims = [w, w // cfg.beta_inv]
    else:
        dims = [w] if cfg.arch == "slow" else [w // cfg.beta_inv]
    # NB: the reference ResNet builder uses width_per_group for the single pathway regardless of
    # arch (audio_model_builder.py:672-682) -- "fast" alone keeps 64 channels.
    if npath == 1:
        dims = [w]
    spec = []
    for p in range(npath):
        k = tk[0][p]
        spec.append(("s1.pathway%d_stem.conv" % p, "conv", dict(cin=1, cout=dims[p], k=(k, 7), s=(2, 2), p=(k // 2, 3))))
        spec.append(("s1.pathway%d_stem.bn" % p, "bn", dict(c=dims[p], final=False)))
    depths = STAGE_DEPTH[cfg.depth]
    dim_in = list(dims)
    out_ratio = cfg.beta_inv // cfg.fusion_ratio
    for si in range(4):
        stage = si + 2
        if npath == 2:
            cf = dim_in[1]
            spec.append(("s%d_fuse.conv_f2s" % (stage - 1), "conv",
                         dict(cin=cf, cout=cf * cfg.fusion_ratio, k=(cfg.fusion_kernel, 1), s=(cfg.alpha, 1),
                              p=(cfg.fusion_kernel // 2, 0))))
            spec.append(("s%d_fuse.bn" % (stage - 1), "bn", dict(c=cf * cfg.fusion_ratio, final=False)))
            dim_in[0] = dim_in[0] + cf * cfg.fusion_ratio
        mult = 2 ** si
        dim_out = [w * 4 * mult] + ([w * 4 * mult // cfg.beta_inv] if npath == 2 else [])
        inner = [w * mult] + ([w * mult // cfg.beta_inv] if npath == 2 else [])
        for p in range(npath):
            for i in range(depths[si]):
                cin = dim_in[p] if i == 0 else dim_out[p]
                st = cfg.freq_strides[si] if i == 0 else 1
                pre = "s%d.pathway%d_res%d." % (stage, p, i)
                if cin != dim_out[p] or st != 1:
                    spec.append((pre + "branch1", "conv", dict(cin=cin, cout=dim_out[p], k=(1, 1), s=(1, st), p=(0, 0))))
                    spec.append((pre + "branch1_bn", "bn", dict(c=dim_out[p], final=False)))
                t = tk[si + 1][p]
                spec.append((pre + "branch2.a", "conv", dict(cin=cin, cout=inner[p], k=(t, 1), s=(1, 1), p=(t // 2, 0))))
                spec.append((pre + "branch2.a_bn", "bn", dict(c=inner[p], final=False)))
                spec.append((pre + "branch2.b", "conv", dict(cin=inner[p], cout=inner[p], k=(1, 3), s=(1, st), p=(0, 1))))
                spec.append((pre + "branch2.b_bn", "bn", dict(c=inner[p], final=False)))
                spec.append((pre + "branch2.c", "conv", dict(cin=inner[p], cout=dim_out[p], k=(1, 1), s=(1, 1), p=(0, 0))))
                spec.append((pre + "branch2.c_bn", "bn", dict(c=dim_out[p], final=True)))
        dim_in = list(dim_out)
    feat = sum(dim_in)
    if len(cfg.num_classes) > 1:
        spec.append(("head.projection_verb", "fc", dict(cin=feat, cout=cfg.num_classes[0])))
        spec.append(("head.projection_noun", "fc", dict(cin=feat, cout=cfg.num_classes[1])))
    else:
        spec.append(("head.projection", "fc", dict(cin=feat, cout=cfg.num_classes[0])))
    return spec


def _registration_order(spec):
    """state_dict order differs from construction order in one place: inside a ResStage the
    reference registers all blocks of pathway0 then pathway1 -- which network_spec already does --
    and fuse modules are registered *after* the stage they precede in ``forward`` order only for
    s1 (s1, s1_fuse, s2, s2_fuse, ...).  network_spec emits fuse before the stage it feeds, which
    is the same sequence.  So registration order == spec order."""
    return spec


def init_state_dict(cfg, seed=0):
    """Reproduces ``torch.manual_seed(seed); build_model(cfg).state_dict()`` of the reference
    bit-exactly: first every Conv2d/Linear constructor draws its default init in construction
    order, then ``init_weights`` redraws conv (kaiming_normal fan_out) and fc (N(0, std)) weights
    in ``modules()`` order (weight_init_helper.py:17-39)."""
    torch.manual_seed(seed)
    spec = _registration_order(network_spec(cfg))
    mods = []
    for name, kind, m in spec:
        if kind == "conv":
            mods.append(nn.Conv2d(m["cin"], m["cout"], list(m["k"]), stride=list(m["s"]), padding=list(m["p"]), bias=False))
        elif kind == "fc":
            mods.append(nn.Linear(m["cin"], m["cout"], bias=True))
        else:
            mods.append(nn.BatchNorm2d(m["c"], eps=BN_EPS, momentum=BN_MOM))
    for (name, kind, m), mod in zip(spec, mods):
        if kind == "conv":
            nn.init.kaiming_normal_(mod.weight, mode="fan_out", nonlinearity="relu")
        elif kind == "bn":
            mod.weight.data.fill_(0.0 if (m["final"] and cfg.zero_init_final_bn) else 1.0)
            mod.bias.data.zero_()
        else:
            mod.weight.data.normal_(mean=0.0, std=cfg.fc_init_std)
            mod.bias.data.zero_()
    sd = OrderedDict()
    for (name, kind, m), mod in zip(spec, mods):
        for k, v in mod.state_dict().items():
            sd[name + "." + k] = v.detach().clone()
    return sd


def perturb_bn(sd, seed=7):
    """Deterministically randomises every BN's gamma/beta/running stats (and nothing else) so
    that ZERO_INIT_FINAL_BN no longer zeroes each residual branch.  Applied identically to the
    real reference when the golden vectors are made."""
    g = torch.Generator().manual_seed(seed)
    out = OrderedDict()
    for k, v in sd.items():
        if k.endswith("_bn.weight") or k.endswith(".bn.weight"):
            out[k] = 0.5 + torch.rand(v.shape, generator=g)
        elif k.endswith("_bn.bias") or k.endswith(".bn.bias"):
            out[k] = 0.2 * torch.randn(v.shape, generator=g)
        elif k.endswith("running_mean"):
            out[k] = 0.1 * torch.randn(v.shape, generator=g)
        elif k.endswith("running_var"):
            out[k] = 0.5 + torch.rand(v.shape, generator=g)
        else:
            out[k] = v.clone()
    return out


def synthetic_inputs(cfg, batch, seed=1234):
    """spectrogram-level synthetic input = oracle log-mel of seeded synthetic
    waveforms, packed per pathway.  Returns ([slow, fast] or [x]) float32 NCHW tensors."""
    import numpy as np
    from . import logmel_np as L

    n = int(round(24000 * (cfg.num_frames / 200.0 - 0.001))) - 1
    wav = L.synthetic_waveforms(batch, n, seed)
    spec = np.stack([L.extract_sound_feature(w, num_frames=cfg.num_frames) for w in wav])  # (B,1,T,F)
    spec = torch.from_numpy(spec)
    if cfg.arch == "slowfast":
        idx = torch.from_numpy(L.slow_indices(spec.shape[2], cfg.alpha))
        return [spec[:, :, idx, :].contiguous(), spec]
    return [spec]


def synthetic_labels(cfg, batch, seed=1235):
    g = torch.Generator().manual_seed(seed)
    if len(cfg.num_classes) > 1:
        return {"verb": torch.randint(0, cfg.num_classes[0], (batch,), generator=g),
                "noun": torch.randint(0, cfg.num_classes[1], (batch,), generator=g)}
    return torch.randint(0, cfg.num_classes[0], (batch,), generator=g)


def calibrated_state(cfg, seed=0, bn_seed=7, calib_batch=2):
    """A 'trained-looking' deterministic state for parity tests: reference random init (seed),
    random BN gamma/beta (so ZERO_INIT_FINAL_BN does not zero every residual branch), and running
    statistics set to the batch statistics of a seeded calibration batch (one training-mode
    forward with momentum 1), so that frozen/eval BN sees sanely normalised activations."""
    sd = init_state_dict(cfg, seed)
    g = torch.Generator().manual_seed(bn_seed)
    spec = {n: (k, m) for n, k, m in network_spec(cfg)}
    for name, (kind, m) in spec.items():
        if kind != "bn":
            continue
        lo, hi = (0.2, 0.5) if m["final"] else (0.7, 1.3)
        sd[name + ".weight"] = lo + (hi - lo) * torch.rand(m["c"], generator=g)
        sd[name + ".bias"] = 0.1 * torch.randn(m["c"], generator=g)
    import copy

    c2 = copy.copy(cfg)
    c2.bn_freeze = False
    c2.dropout = 0.0
    with torch.no_grad():
        Net(c2, sd, bn_momentum=1.0).forward(synthetic_inputs(cfg, calib_batch, seed=4321), training=True)
    for k in sd:
        if k.endswith("num_batches_tracked"):
            sd[k].zero_()
    return sd


def bn_is_train(name, training, bn_freeze):
    """model.train() + (BN.FREEZE -> freeze_fn('bn_statistics')): audio_model_builder.py:337-344."""
    if not training:
        return False
    if bn_freeze and not any(name.startswith(u) for u in UNFROZEN_BN):
        return False
    return True


class Net:
    """Functional network bound to a state_dict (tensors may require grad)."""

    def __init__(self, cfg, sd, bn_momentum=BN_MOM):
        self.cfg = cfg
        self.sd = sd
        self.bn_momentum = bn_momentum
        self.spec = {n: (k, m) for n, k, m in network_spec(cfg)}

    def conv(self, name, x):
        _, m = self.spec[name]
        return F.conv2d(x, self.sd[name + ".weight"], None, stride=m["s"], padding=m["p"])

    def bn(self, name, x, training):
        sd = self.sd
        tr = bn_is_train(name, training, self.cfg.bn_freeze)
        y = F.batch_norm(x, sd[name + ".running_mean"], sd[name + ".running_var"], sd[name + ".weight"],
                         sd[name + ".bias"], tr, self.bn_momentum, BN_EPS)
        if tr:
            sd[name + ".num_batches_tracked"] += 1
        return y

    # ``forced`` (test aid): the ReLU / max-pool DECISIONS of another execution of the same network.  A ReLU named
    # ``n`` becomes ``x * forced[n]`` ({0,1} mask, NCHW) and the stem max-pool gathers ``forced[n + ".pool"]`` (int64
    # flat (t, f) indices per output element) instead of taking its own arg-max.  With forced decisions the network is
    # a fixed piecewise-linear map, so two executions in different precisions differ by accumulated rounding only (a
    # few 1e-3 in bf16) instead of by the sqrt(eps) amplification of flipped ReLU masks -- which is what lets a
    # gradient parity test tell a wrong backward from rounding noise.  forced = None is the plain network.
    forced = None

    def relu(self, name, x):
        if self.forced is None:
            return F.relu(x)
        return x * self.forced[name].to(x.dtype)

    def stem(self, p, x, training):
        pre = "s1.pathway%d_stem." % p
        name = pre[:-1]
        x = self.relu(name, self.bn(pre + "bn", self.conv(pre + "conv", x), training))
        if self.forced is None:
            return F.max_pool2d(x, kernel_size=[3, 3], stride=[2, 2], padding=[1, 1])
        idx = self.forced[name + ".pool"]  # (B, C, Tp, Fp) flat indices into the (T, F) plane
        return torch.gather(x.flatten(2), 2, idx.flatten(2)).view(idx.shape)

    def fuse(self, stage, xs, xf, training):
        pre = "s%d_fuse." % stage
        f = self.relu(pre[:-1], self.bn(pre + "bn", self.conv(pre + "conv_f2s", xf), training))
        return torch.cat([xs, f], 1)

    def block(self, pre, x, training):
        if (pre + "branch1") in self.spec:
            sc = self.bn(pre + "branch1_bn", self.conv(pre + "branch1", x), training)
        else:
            sc = x
        h = self.relu(pre + "branch2.a", self.bn(pre + "branch2.a_bn", self.conv(pre + "branch2.a", x), training))
        h = self.relu(pre + "branch2.b", self.bn(pre + "branch2.b_bn", self.conv(pre + "branch2.b", h), training))
        h = self.bn(pre + "branch2.c_bn", self.conv(pre + "branch2.c", h), training)
        return self.relu(pre + "branch2.c", sc + h)

    def forward(self, inputs, training=False, dropout_mask=None, taps=None):
        """inputs: [slow (B,1,Ts,F), fast (B,1,Tf,F)] or [x].  Returns what the reference head
        returns: raw logits when training, softmax probabilities (spatially averaged) in eval.
        ``dropout_mask`` (B, feat) of {0,1}: applied as x*mask/(1-p) in training (the reference
        draws it from torch's RNG; passing it in makes the comparison deterministic).
        ``taps``: optional dict filled with intermediate activations."""
        cfg = self.cfg
        npath = 2 if cfg.arch == "slowfast" else 1
        assert len(inputs) == npath, "Input tensor does not contain {} pathway".format(npath)
        x = [self.stem(p, inputs[p], training) for p in range(npath)]
        depths = STAGE_DEPTH[cfg.depth]
        for si in range(4):
            stage = si + 2
            if npath == 2:
                x[0] = self.fuse(stage - 1, x[0], x[1], training)
            for p in range(npath):
                h = x[p]
                for i in range(depths[si]):
                    h = self.block("s%d.pathway%d_res%d." % (stage, p, i), h, training)
                x[p] = h
            if taps is not None:
                taps["s%d" % stage] = [t for t in x]
        pooled = []
        for p in range(npath):
            T = cfg.num_frames // (cfg.alpha if (npath == 2 and p == 0) else 1) // 4
            pooled.append(F.avg_pool2d(x[p], [T, cfg.num_freq // 32], stride=1))
        f = torch.cat(pooled, 1).permute(0, 2, 3, 1)
        if training and cfg.dropout > 0:
            if dropout_mask is not None:
                f = f * dropout_mask.view(f.shape[0], 1, 1, -1) / (1.0 - cfg.dropout)
            else:
                f = F.dropout(f, cfg.dropout, True)
        outs = []
        names = ["head.projection_verb", "head.projection_noun"] if len(cfg.num_classes) > 1 else ["head.projection"]
        for n in names:
            o = F.linear(f, self.sd[n + ".weight"], self.sd[n + ".bias"])
            if not training:
                o = F.softmax(o, dim=3).mean([1, 2])
            outs.append(o.reshape(o.shape[0], -1))
        return tuple(outs) if len(outs) > 1 else outs[0]


def record_decisions(cfg, sd, inputs, training=True):
    """The oracle's own ReLU masks and stem pooling indices in the ``forced`` format (used to pin the forced path:
    a network forced with its own decisions reproduces itself)."""
    forced = {}

    class Rec(Net):
        def relu(self, name, x):
            forced[name] = (x > 0)
            return F.relu(x)

        def stem(self, p, x, training):
            pre = "s1.pathway%d_stem." % p
            x = self.relu(pre[:-1], self.bn(pre + "bn", self.conv(pre + "conv", x), training))
            y, idx = F.max_pool2d(x, kernel_size=[3, 3], stride=[2, 2], padding=[1, 1], return_indices=True)
            forced[pre[:-1] + ".pool"] = idx
            return y

    with torch.no_grad():
        Rec(cfg, OrderedDict((k, v.clone()) for k, v in sd.items())).forward(inputs, training=training)
    return forced


def compute_loss(preds, labels):
    """train_utils.py:76-101 (ONLY_ACTION_RECOGNITION branch): mean of verb and noun CE."""
    if isinstance(preds, (tuple, list)):
        lv = F.cross_entropy(preds[0], labels["verb"])
        ln = F.cross_entropy(preds[1], labels["noun"])
        return 0.5 * (lv + ln), lv, ln
    loss = F.cross_entropy(preds, labels)
    return loss, loss, loss


def loss_and_grads(cfg, sd, inputs, labels, dropout_mask=None, forced=None, taps=None):
    """One training-mode forward + backward; returns (loss, preds, {param: grad}, updated sd).
    ``forced``: see Net.relu (ReLU masks / pooling indices of another execution)."""
    sd = OrderedDict((k, v.clone()) for k, v in sd.items())
    params = {}
    for k, v in sd.items():
        if v.dtype.is_floating_point and "running_" not in k:
            frozen_bn = cfg.bn_freeze and ("bn." in k) and not any(k.startswith(u) for u in UNFROZEN_BN)
            if not frozen_bn:
                v.requires_grad_(True)
                params[k] = v
    net = Net(cfg, sd)
    net.forced = forced
    preds = net.forward(inputs, training=True, dropout_mask=dropout_mask, taps=taps)
    loss, _, _ = compute_loss(preds, labels)
    grads = torch.autograd.grad(loss, list(params.values()))
    return loss.detach(), preds, dict(zip(params.keys(), grads)), sd

"""GPU aid: isolates the fast stem's backward (fp32 parity mode) against a float64 torch replica fed with the engine's
own upstream gradient, with the replica's own max-pool / ReLU decisions and with the engine's (forced)."""
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "auditory-slow-fast_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False

from asf_b200.config.defaults import epic_slowfast_cfg  # noqa: E402
from asf_b200.models import build_model  # noqa: E402
from gpu_util import engine_decisions, nchw  # noqa: E402
from oracle import model_torch as M  # noqa: E402


def rel(a, b):
    return float((a.double() - b.double()).norm() / (b.double().norm() + 1e-30))


ocfg = M.epic_cfg(dropout=0.0, bn_freeze=True)
sd = M.calibrated_state(ocfg)
cfg = epic_slowfast_cfg((97, 300), True)
cfg.MODEL.DROPOUT_RATE = 0.0
cfg.B200.COMPUTE_DTYPE = sys.argv[1] if len(sys.argv) > 1 else "float32"
cfg.NUM_GPUS = 1
model = build_model(cfg)
model.load_state_dict(sd)
model.train()
model.freeze_fn("bn_parameters")
model.freeze_fn("bn_statistics")
inputs, labels = M.synthetic_inputs(ocfg, 4, seed=4242), M.synthetic_labels(ocfg, 4, seed=4243)
preds = model([t.cuda() for t in inputs])
loss, _, _ = M.compute_loss(preds, {k: v.cuda() for k, v in labels.items()})
loss.backward()
torch.cuda.synchronize()
forced = engine_decisions(model)
for p in (0, 1):
    u = model.engine.plan.stems[p]
    name = u.name
    dpool = nchw(u.pooled.grad.view().double())
    raw = nchw(u.y.view().double())
    x = inputs[p].cuda().double()
    w = sd[name + ".conv.weight"].cuda().double().requires_grad_(True)
    gamma = sd[name + ".bn.weight"].cuda().double().requires_grad_(True)
    beta = sd[name + ".bn.bias"].cuda().double().requires_grad_(True)
    conv = F.conv2d(x, w, stride=u.conv.stride, padding=u.conv.padding)
    print("== %s: raw conv rel err %.3e" % (name, rel(raw, conv.detach())))
    bn = F.batch_norm(conv, None, None, gamma, beta, True, 0.1, 1e-5)
    own_mask = bn.detach() > 0
    act = F.relu(bn)
    pooled, idx = F.max_pool2d(act, 3, 2, 1, return_indices=True)
    print("   premask mismatches: %d of %d; pool index mismatches: %d of %d (of which pooled value > 0: %d)" % (
        int((own_mask != forced[name]).sum()), own_mask.numel(), int((idx != forced[name + ".pool"]).sum()), idx.numel(),
        int(((idx != forced[name + ".pool"]) & (pooled.detach() > 0)).sum())))
    ours = {"conv.weight": u.conv.weight.grad, "bn.weight": u.bn.weight.grad, "bn.bias": u.bn.bias.grad}
    ga = torch.autograd.grad(pooled, [w, gamma, beta], dpool, retain_graph=True)
    fa = bn * forced[name].double()
    pooled_f = torch.gather(fa.flatten(2), 2, forced[name + ".pool"].flatten(2)).view(idx.shape)
    gb = torch.autograd.grad(pooled_f, [w, gamma, beta], dpool)
    for k, a, b in zip(ours, ga, gb):
        print("   %-12s ours vs replica(own decisions) %.3e   ours vs replica(forced) %.3e   own vs forced %.3e" % (
            k, rel(ours[k], a), rel(ours[k], b), rel(a, b)))
    print("   pooled value: ours vs replica %.3e" % rel(nchw(u.pooled.view().double()), pooled.detach()))

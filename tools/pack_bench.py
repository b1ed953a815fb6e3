"""GPU aid: time of the weight re-pack (asf_pack_weights_multi, both launches) for the bench model."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "auditory-slow-fast_b200")):
    sys.path.insert(0, p)
import torch
from asf_b200.config.defaults import epic_slowfast_cfg
from asf_b200.models import build_model

cfg = epic_slowfast_cfg((97, 300), True)
cfg.NUM_GPUS = 1
m = build_model(cfg)
m.train()
m.freeze_fn("bn_parameters"); m.freeze_fn("bn_statistics")
x = [torch.randn(4, 1, 100, 128).cuda(), torch.randn(4, 1, 400, 128).cuda()]
with torch.no_grad():
    m(x)
eng = m.engine
params = [p for p in m.parameters()]
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
ts = []
for _ in range(6):
    torch.autograd.graph.increment_version(params)
    torch.cuda.synchronize()
    e0.record()
    eng._prepare_all(True)
    e1.record()
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1) * 1e3)
print("re-pack of %.1f M parameters: %s us" % (sum(p.numel() for p in params) / 1e6, ["%.0f" % t for t in ts]))

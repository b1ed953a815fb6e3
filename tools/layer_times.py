"""GPU aid: per-launch tap-GEMM timings (shape, tile, achieved TFLOP/s) for one step of bench's train workload."""
import ctypes as C
import collections
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "auditory-slow-fast_b200")):
    sys.path.insert(0, p)
import torch
from asf_b200 import _lib as L
from asf_b200.config.defaults import epic_slowfast_cfg
from asf_b200.models import build_model

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
freeze = int(sys.argv[2]) if len(sys.argv) > 2 else 0
cfg = epic_slowfast_cfg((97, 300), bool(freeze))
cfg.BN.FREEZE = bool(freeze)
cfg.NUM_GPUS = 1
m = build_model(cfg)
m.train()
if freeze:
    m.freeze_fn("bn_parameters"); m.freeze_fn("bn_statistics")
x = [torch.randn(B, 1, 100, 128).cuda(), torch.randn(B, 1, 400, 128).cuda()]
def step():
    v, n = m(x)
    (v.float().mean() + n.float().mean()).backward()
for _ in range(3):
    step()
torch.cuda.synchronize()
L.call("asf_prof_start", 8192)
step()
torch.cuda.synchronize()
ms, fl, n = (C.c_double * 2)(), (C.c_double * 2)(), (C.c_longlong * 2)()
L.call("asf_prof_stop", ms, fl, n)
tot = n[0] + n[1]
agg = collections.OrderedDict()
for i in range(tot):
    r = (C.c_double * 9)()
    L.call("asf_prof_record", i, r)
    key = (int(r[0]), int(r[3]), int(r[4]), int(r[5]), int(r[6]), int(r[7]), int(r[8]))
    a = agg.setdefault(key, [0, 0.0, 0.0])
    a[0] += 1; a[1] += r[1]; a[2] += r[2]
print("mode      M     N     K taps   bn grid | count  total_us  avg_us  TFLOP/s")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%4d %7d %5d %5d %4d %4d %4d | %5d %9.1f %7.1f %8.1f" % (k + (a[0], a[1] * 1e3, a[1] * 1e3 / a[0], a[2] / (a[1] * 1e-3) / 1e12)))
print("fwd/dgrad: %.2f ms, %.1f TFLOP/s; wgrad: %.2f ms, %.1f TFLOP/s" % (ms[0], fl[0] / ms[0] / 1e9, ms[1], fl[1] / max(ms[1], 1e-9) / 1e9))

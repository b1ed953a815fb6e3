"""GPU aid: kernel timeline of the GRAPHED training step (what bench.py times) from torch.profiler (CUPTI activity
records, no replay / cache flush as under ncu): per-kernel busy time, idle gaps between consecutive kernels, and the
step span.  Usage: python tools/timeline.py [batch] [bn_freeze] [out_prefix]"""
import collections
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "auditory-slow-fast_b200")):
    sys.path.insert(0, p)
import torch
from torch.profiler import ProfilerActivity, profile

from asf_b200.config.defaults import epic_slowfast_cfg
from asf_b200.models import build_model
from asf_b200.optim import FusedSGD
from asf_b200.train_step import TrainStep
from oracle import logmel_np as OL  # synthetic-input generator only

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
freeze = int(sys.argv[2]) if len(sys.argv) > 2 else 0
prefix = sys.argv[3] if len(sys.argv) > 3 else "gpurun_out/timeline"
cfg = epic_slowfast_cfg((97, 300), bool(freeze))
cfg.BN.FREEZE = bool(freeze)
cfg.NUM_GPUS = 1
torch.manual_seed(0)
model = build_model(cfg)
model.train()
if freeze:
    model.freeze_fn("bn_parameters")
    model.freeze_fn("bn_statistics")
bn_params = [p for n, p in model.named_parameters() if "bn" in n and p.requires_grad]
other = [p for n, p in model.named_parameters() if "bn" not in n and p.requires_grad]
opt = FusedSGD([{"params": bn_params, "weight_decay": 0.0}, {"params": other, "weight_decay": 1e-4}], lr=0.01,
               momentum=0.9, nesterov=True)
ts = TrainStep(cfg, model, opt, B, 47975, graph=True, segmented_update=os.environ.get("ASF_TL_SEGMENTED") == "1")
w = torch.from_numpy(OL.synthetic_waveforms(B, 47975, 1234)).cuda()
g = torch.Generator().manual_seed(1)
lab = {"verb": torch.randint(0, 97, (B,), generator=g).cuda(), "noun": torch.randint(0, 300, (B,), generator=g).cuda()}
for _ in range(5):
    ts(w, lab)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3):
        ts(w, lab)
    torch.cuda.synchronize()
kin = prof.profiler.kineto_results.events()
rows = []
for e in kin:
    if e.device_type() == torch.autograd.DeviceType.CUDA:
        s0 = e.start_ns() / 1e3
        rows.append((e.name(), s0, s0 + e.duration_ns() / 1e3, e.device_resource_id()))
rows.sort(key=lambda r: r[1])
# split into steps at the log-mel kernel
starts = [i for i, r in enumerate(rows) if "logmel" in r[0]]
print("cuda events", len(rows), "steps", len(starts))
if len(starts) >= 2:
    a, b = starts[-2], starts[-1]
    step = rows[a:b]
    span = rows[b][1] - rows[a][1]
    busy = collections.defaultdict(lambda: [0, 0.0])
    gaps = collections.defaultdict(lambda: [0, 0.0])
    prev_end = step[0][1]
    per_stream = collections.defaultdict(lambda: [0, 0.0, 1e18, 0.0])
    tot_busy = tot_gap = 0.0
    with open(prefix + "_events.csv", "w") as f:
        for name, s, e, sid in step:
            short = name.split("(")[0].replace("void ", "")[:60]
            gap = max(0.0, s - prev_end)
            busy[short][0] += 1
            busy[short][1] += e - s
            gaps[short][0] += 1
            gaps[short][1] += gap
            tot_busy += e - s
            tot_gap += gap
            f.write("%s,%.3f,%.3f,%.3f,%d\n" % (short, s - step[0][1], e - s, gap, sid))
            per_stream[sid][0] += 1
            per_stream[sid][1] += e - s
            per_stream[sid][2] = min(per_stream[sid][2], s - step[0][1])
            per_stream[sid][3] = max(per_stream[sid][3], e - step[0][1])
            prev_end = max(prev_end, e)
    with open(prefix + "_summary.txt", "w") as f:
        def out(*a):
            print(*a)
            print(*a, file=f)
        out("step span %.1f us, kernels %d, sum busy %.1f us, sum gaps %.1f us" % (span, len(step), tot_busy, tot_gap))
        for sid, v in sorted(per_stream.items()):
            out("stream %d: %d kernels, busy %.1f us, first start %.1f, last end %.1f" % (sid, v[0], v[1], v[2], v[3]))
        out("%-62s %5s %10s %8s %10s" % ("kernel", "n", "busy_us", "avg", "gap_before"))
        for k, v in sorted(busy.items(), key=lambda x: -x[1][1]):
            out("%-62s %5d %10.1f %8.2f %10.1f" % (k, v[0], v[1], v[1] / v[0], gaps[k][1]))

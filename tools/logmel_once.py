"""GPU aid: a few log-mel launches on a 256-clip batch (for ncu captures / quick timing)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "auditory-slow-fast_b200")):
    sys.path.insert(0, p)
import torch

from asf_b200.config.defaults import epic_slowfast_cfg
from asf_b200.datasets.audio_loader_epic import logmel_batch

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
cfg = epic_slowfast_cfg()
w = (0.1 * torch.randn(B, 47975)).clamp(-1, 1).cuda()
for _ in range(3):
    logmel_batch(cfg, w)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    logmel_batch(cfg, w)
e1.record()
torch.cuda.synchronize()
print("logmel_batch: %.3f ms per %d clips" % (e0.elapsed_time(e1) / 5, B))
from asf_b200.datasets.audio_loader_epic import logmel_cuda
import subprocess
for B2 in (64, 256):
    w2 = w[:B2].contiguous()
    ts = []
    for _ in range(8):
        a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        logmel_cuda(w2, num_frames=400)
        b_.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b_))
    print("logmel_cuda B=%d per-call ms:" % B2, ["%.2f" % t for t in ts])
print(subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active", "--format=csv,noheader"], capture_output=True, text=True).stdout)

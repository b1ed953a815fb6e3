"""GPU aid: one eager training step of bench's workload between cudaProfilerStart/Stop, for
    ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv ...
Usage: python tools/step_profile.py [batch] [bn_freeze]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "auditory-slow-fast_b200")):
    sys.path.insert(0, p)
import torch

from asf_b200.config.defaults import epic_slowfast_cfg
from asf_b200.models import build_model
from asf_b200.optim import FusedSGD
from asf_b200.train_step import TrainStep
from oracle import logmel_np as OL  # synthetic-input generator only

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
freeze = int(sys.argv[2]) if len(sys.argv) > 2 else 0
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
ts = TrainStep(cfg, model, opt, B, 47975, graph=False)
w = torch.from_numpy(OL.synthetic_waveforms(B, 47975, 1234)).cuda()
g = torch.Generator().manual_seed(1)
lab = {"verb": torch.randint(0, 97, (B,), generator=g).cuda(), "noun": torch.randint(0, 300, (B,), generator=g).cuda()}
for _ in range(3):
    ts(w, lab)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
ts(w, lab)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print("done")

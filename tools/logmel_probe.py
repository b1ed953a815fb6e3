import os, sys
ROOT = "/root/repo"
for p in (ROOT, os.path.join(ROOT, "auditory-slow-fast_b200")):
    sys.path.insert(0, p)
import torch
from asf_b200.datasets.audio_loader_epic import logmel_cuda
w = (0.1 * torch.randn(int(sys.argv[1]), 47975, device="cuda")).clamp_(-1, 1)
for _ in range(3):
    logmel_cuda(w, impl="tc")
torch.cuda.synchronize()

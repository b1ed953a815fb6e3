"""GPU aid: log-mel kernels timed alone (CUDA events around 20 back-to-back launches).  python tools/logmel_bench.py [B]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "auditory-slow-fast_b200")):
    sys.path.insert(0, p)
import torch

from asf_b200.datasets.audio_loader_epic import logmel_cuda

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
w = (0.1 * torch.randn(B, 47975, device="cuda")).clamp_(-1, 1)
for impl in ("tc", "simt"):
    logmel_cuda(w, impl=impl)
    torch.cuda.synchronize()
    for _ in range(30):  # spin the clocks up
        logmel_cuda(w, impl=impl)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        logmel_cuda(w, impl=impl)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / 20 * 1e3
    print("%-5s B=%d  %8.1f us  %8.0f clips/s  %6.2f TFLOP/s(dense fp32-equivalent)  %6.1f GB/s(alg)" %
          (impl, B, us, B / us * 1e6, B * 0.4986e9 / us / 1e6, B * 396700 / us / 1e3))
a = logmel_cuda(w, impl="tc")
b = logmel_cuda(w, impl="simt")
print("max |tc - simt| on log values:", float((a - b).abs().max()))

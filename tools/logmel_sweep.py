"""GPU aid: BASELINE configs[4] -- log-mel extraction throughput sweep, batch 1 ... 4096 clips x 1 ... 10 s, on one B200
(asf_logmel_fwd_tc through the Python API, device-resident waveforms, CUDA events) beside the CPU restatement of the
reference's librosa path (oracle/logmel_np.py; one core, and a process pool over all host cores like the reference's
DataLoader workers).   python tools/logmel_sweep.py [--cpu-seconds 2] > profiles/r02_logmel_sweep.txt"""
import os
import sys
import time
from concurrent.futures import ProcessPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "auditory-slow-fast_b200")):
    sys.path.insert(0, p)
import numpy as np
import torch

from asf_b200.datasets.audio_loader_epic import logmel_cuda
from oracle import logmel_np as OL


def _cpu_clip(args):
    n, seed = args
    w = OL.synthetic_waveforms(1, n, seed)[0]
    return OL.log_specgram(w).shape[0]


def cpu_rate(n, seconds, workers):
    if workers == 1:
        t0, k = time.perf_counter(), 0
        while time.perf_counter() - t0 < seconds or k < 2:
            _cpu_clip((n, k))
            k += 1
        return k / (time.perf_counter() - t0)
    with ProcessPoolExecutor(workers) as ex:
        list(ex.map(_cpu_clip, [(n, i) for i in range(workers)]))  # spin the pool up
        k = max(workers * 2, int(workers * seconds * 10))
        t0 = time.perf_counter()
        list(ex.map(_cpu_clip, [(n, i) for i in range(k)], chunksize=max(1, k // (workers * 4))))
        return k / (time.perf_counter() - t0)


def main():
    cpu_s = float(sys.argv[sys.argv.index("--cpu-seconds") + 1]) if "--cpu-seconds" in sys.argv else 2.0
    cores = os.cpu_count() or 1
    print("# log-mel extraction sweep (BASELINE configs[4]); GPU: %s; host cores: %d" % (torch.cuda.get_device_name(0), cores))
    print("# N = round(24000 (d - 0.001)) - 1 samples -> T = 200 d frames x 128 mels; algorithmic bytes = 4 N + 4 T 128 per clip")
    print("%5s %6s %8s %6s | %12s %10s %9s | %12s %12s | %9s" % ("d_s", "B", "N", "T", "gpu_clips/s", "us/launch", "GB/s_alg",
                                                            "cpu1_clips/s", "cpuN_clips/s", "gpu/cpuN"))
    g = torch.Generator(device="cuda").manual_seed(1234)
    for d in (1, 2, 4, 10):
        n = int(round(24000 * (d - 0.001))) - 1
        T = 1 + n // 120
        c1 = cpu_rate(n, cpu_s, 1)
        cn = cpu_rate(n, cpu_s, cores)
        for B in (1, 16, 256, 1024, 4096):
            w = (0.1 * torch.randn(B, n, device="cuda", generator=g)).clamp_(-1, 1)
            out = logmel_cuda(w)
            assert out.shape == (B, 1, T, 128)
            torch.cuda.synchronize()
            for _ in range(3):
                logmel_cuda(w)
            reps = int(max(3, min(200, 2e9 / (B * n * 4 + 1))))
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                logmel_cuda(w)
            e1.record()
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) / reps * 1e3
            rate = B / us * 1e6
            print("%5d %6d %8d %6d | %12.0f %10.1f %9.1f | %12.1f %12.1f | %9.0f" % (
                d, B, n, T, rate, us, B * (4 * n + 4 * T * 128) / us / 1e3, c1, cn, rate / cn))
            del w, out
            torch.cuda.empty_cache()


if __name__ == "__main__":
    main()

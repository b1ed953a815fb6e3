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
import numpy as np  # noqa: F401

from oracle import logmel_np as OL  # numpy only: the spawned host workers re-import this file, so no torch up here


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
    # spawn, not fork (this process may hold a CUDA context and BLAS threads), one BLAS thread per worker
    import multiprocessing as mp
    saved = {k: os.environ.get(k) for k in ("OMP_NUM_THREADS", "MKL_NUM_THREADS", "OPENBLAS_NUM_THREADS")}
    for k in saved:
        os.environ[k] = "1"
    try:
        with ProcessPoolExecutor(workers, mp_context=mp.get_context("spawn")) as ex:
            list(ex.map(_cpu_clip, [(n, i) for i in range(workers)]))  # spin the pool up
            k = workers * 4
            t0 = time.perf_counter()
            list(ex.map(_cpu_clip, [(n, i) for i in range(k)], chunksize=1))
            return k / (time.perf_counter() - t0)
    finally:
        for k, v in saved.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


def main():
    import torch

    from asf_b200.datasets.audio_loader_epic import logmel_cuda
    cpu_s = float(sys.argv[sys.argv.index("--cpu-seconds") + 1]) if "--cpu-seconds" in sys.argv else 1.0
    cores = os.cpu_count() or 1
    print("# log-mel extraction sweep (BASELINE configs[4]); GPU: %s; host cores: %d" % (
        torch.cuda.get_device_name(0) if torch.cuda.is_available() else "none", cores))
    print("# N = round(24000 (d - 0.001)) - 1 samples -> T = 200 d frames x 128 mels; algorithmic bytes = 4 N + 4 T 128 per clip")
    print("%5s %6s %8s %6s | %12s %10s %9s | %12s %12s | %9s" % ("d_s", "B", "N", "T", "gpu_clips/s", "us/launch", "GB/s_alg",
                                                            "cpu1_clips/s", "cpuN_clips/s", "gpu/cpuN"))
    durations = (1, 2, 4, 10)
    lens = {d: int(round(24000 * (d - 0.001))) - 1 for d in durations}
    # host rates first, before this process touches CUDA
    t0 = time.perf_counter()
    cpu1 = {d: cpu_rate(lens[d], cpu_s, 1) for d in durations}
    cpun = {d: cpu_rate(lens[d], cpu_s, cores) for d in durations}
    print("# host rates measured in %.0f s: 1 core %s, %d cores %s" % (
        time.perf_counter() - t0, {d: round(v, 1) for d, v in cpu1.items()}, cores, {d: round(v, 1) for d, v in cpun.items()}),
        flush=True)
    if "--cpu-only" in sys.argv:
        return
    g = torch.Generator(device="cuda").manual_seed(1234)
    for d in durations:
        n = lens[d]
        T = 1 + n // 120
        c1, cn = cpu1[d], cpun[d]
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

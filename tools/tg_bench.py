"""GPU aid: single tap-GEMM layers of the real network (B = 64 shapes) timed in isolation with CUDA events.

    python tools/tg_bench.py [case-substring] [--reps N] [--once]

--once launches every selected case exactly once after one warm-up (what an `ncu --set full` capture wants).
Modes per case: fwd_frozen (shift + residual + ReLU epilogue), fwd_stats (raw output + per-channel sums),
dgrad (transposed weights, mask epilogue), wgrad."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "auditory-slow-fast_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch

from asf_b200.engine import Act
from gpu_util import Harness, make_unit

# (name, B, T, F, cin, cout, kernel, stride, pad)
CASES = [
    ("s4_a_3x1_1024_256", 64, 25, 8, 1024, 256, (3, 1), (1, 1), (1, 0)),
    ("s4_b_1x3_256_256", 64, 25, 8, 256, 256, (1, 3), (1, 1), (0, 1)),
    ("s4_c_1x1_256_1024", 64, 25, 8, 256, 1024, (1, 1), (1, 1), (0, 0)),
    ("s5_a_3x1_2048_512", 64, 25, 4, 2048, 512, (3, 1), (1, 1), (1, 0)),
    ("s5_c_1x1_512_2048", 64, 25, 4, 512, 2048, (1, 1), (1, 1), (0, 0)),
    ("s3_c_1x1_128_512", 64, 25, 16, 128, 512, (1, 1), (1, 1), (0, 0)),
    ("s3_a_1x1_512_128", 64, 25, 16, 512, 128, (1, 1), (1, 1), (0, 0)),
    ("s2_c_1x1_64_256", 64, 25, 32, 64, 256, (1, 1), (1, 1), (0, 0)),
    ("f2_a_3x1_32_8", 64, 100, 32, 32, 8, (3, 1), (1, 1), (1, 0)),
    ("f2_b_1x3_8_8", 64, 100, 32, 8, 8, (1, 3), (1, 1), (0, 1)),
    ("f2_c_1x1_8_32", 64, 100, 32, 8, 32, (1, 1), (1, 1), (0, 0)),
    ("f3_a_3x1_64_16", 64, 100, 16, 64, 16, (3, 1), (1, 1), (1, 0)),
    ("f4_a_3x1_128_32", 64, 100, 8, 128, 32, (3, 1), (1, 1), (1, 0)),
    ("f4_c_1x1_32_128", 64, 100, 8, 32, 128, (1, 1), (1, 1), (0, 0)),
    ("f5_a_3x1_256_64", 64, 100, 4, 256, 64, (3, 1), (1, 1), (1, 0)),
]


def main():
    argv = sys.argv[1:]
    args = [a for i, a in enumerate(argv) if not a.startswith("--") and (i == 0 or argv[i - 1] not in ("--reps", "--modes"))]
    sel = args[0] if args else ""
    once = "--once" in sys.argv
    reps = 20
    if "--reps" in sys.argv:
        reps = int(sys.argv[sys.argv.index("--reps") + 1])
    modes = ["fwd_frozen", "fwd_stats", "dgrad", "wgrad"]
    if "--modes" in sys.argv:
        modes = sys.argv[sys.argv.index("--modes") + 1].split(",")
    h = Harness(torch.bfloat16)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    print("%-22s %-10s %9s %9s %9s" % ("case", "mode", "us", "TFLOP/s", "GB/s(alg)"))
    for case in CASES:
        name, B, T, Fq, cin, cout, k, s, p = case
        if sel and not any(t in name for t in sel.split(',')):
            continue
        u, x, w = make_unit(h, B, T, Fq, cin, cout, k, s, p, seed=1)
        res = Act(u.y.B, u.y.T, u.y.F, cout, h.dtype, h.device)
        res.buf.normal_()
        dy = Act(u.y.B, u.y.T, u.y.F, cout, h.dtype, h.device)
        dy.buf.normal_()
        g = torch.zeros_like(w)
        rows_out = u.y.rows
        flops = 2.0 * rows_out * cin * cout * k[0] * k[1]
        by_fwd = 2.0 * (u.x.rows * cin + rows_out * cout)

        def fwd_frozen():
            u.bn.eval()
            u.res = res
            h._unit_forward(u, training=False)

        def fwd_stats():
            u.bn.train()
            u.res = None
            h._prepare_unit(u, True)
            h._conv_fwd(u, u.y, dict(stats=u.stats, stats_stride=u.Cout))

        def dgrad():
            u.bn.eval()
            h._prepare_unit(u, True)
            h._conv_dgrad(u, dy, u.x.grad, accumulate=False, mask=u.x)

        def wgrad():
            h._conv_wgrad(u, dy, None, g, accumulate=False)

        for mode in modes:
            fn = {"fwd_frozen": fwd_frozen, "fwd_stats": fwd_stats, "dgrad": dgrad, "wgrad": wgrad}[mode]
            fn()
            torch.cuda.synchronize()
            if once:
                fn()
                torch.cuda.synchronize()
                continue
            tot = 0.0
            for _ in range(reps):
                flush.zero_()  # evict L2 between timed launches
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                fn()
                e1.record()
                torch.cuda.synchronize()
                tot += e0.elapsed_time(e1)
            us = tot / reps * 1e3
            print("%-22s %-10s %9.1f %9.1f %9.1f" % (name, mode, us, flops / us / 1e6, by_fwd / us / 1e3))


if __name__ == "__main__":
    main()

"""GPU aid: where a tap-GEMM forward / dgrad launch spends its time (asf_tapgemm_probe stall counters), per layer
shape of the real network.   python tools/tg_probe.py [case-substring[,..]] [--modes fwd_stats,fwd_frozen,dgrad]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "auditory-slow-fast_b200"), os.path.join(ROOT, "tests"), os.path.join(ROOT, "tools")):
    sys.path.insert(0, p)
import torch

from asf_b200 import _lib as L
from asf_b200.engine import Act
from gpu_util import Harness, make_unit
from tg_bench import CASES

EXTRA = [
    ("f3_b_1x3_16_16", 64, 100, 16, 16, 16, (1, 3), (1, 1), (0, 1)),
    ("f4_b_1x3_32_32", 64, 100, 8, 32, 32, (1, 3), (1, 1), (0, 1)),
    ("s3_b_1x3_128_128", 64, 25, 16, 128, 128, (1, 3), (1, 1), (0, 1)),
]
NAMES = ["body", "prod_wait_stage", "mma_wait_ops", "mma_wait_acc", "epi_wait_acc", "epi_busy", "epi_wait_store",
         "drain"]


def main():
    argv = sys.argv[1:]
    sel = argv[0] if argv and not argv[0].startswith("--") else ""
    modes = ["fwd_stats", "fwd_frozen", "dgrad"]
    if "--modes" in argv:
        modes = argv[argv.index("--modes") + 1].split(",")
    h = Harness(torch.bfloat16)
    cnt = torch.zeros(16, dtype=torch.int64, device="cuda")
    print("%-20s %-10s %7s | kcycles per CTA: %s" % ("case", "mode", "us", " ".join("%15s" % n for n in NAMES)))
    for case in CASES + EXTRA:
        name, B, T, Fq, cin, cout, k, s, p = case
        if sel and not any(t in name for t in sel.split(",")):
            continue
        u, x, w = make_unit(h, B, T, Fq, cin, cout, k, s, p, seed=1)
        res = Act(u.y.B, u.y.T, u.y.F, cout, h.dtype, h.device)
        res.buf.normal_()
        dy = Act(u.y.B, u.y.T, u.y.F, cout, h.dtype, h.device)
        dy.buf.normal_()
        xres = Act(u.x.B, u.x.T, u.x.F, cin, h.dtype, h.device)
        xres.buf.normal_()

        def fwd_frozen():
            u.bn.eval()
            u.res = res
            h._unit_forward(u, training=False)

        def fwd_stats():
            u.bn.train()
            u.res = None
            h._prepare_unit(u, True)
            ws, tk = h._stat_bufs(u)
            h._conv_fwd(u, u.y, dict(stats=u.stats, stats_stride=u.Cout, stats_ws=ws, stats_ticket=tk))

        def fwd_plain():
            u.bn.eval()
            u.res = None
            h._unit_forward(u, training=False)

        def dgrad_res():
            u.bn.eval()
            h._prepare_unit(u, True)
            h._conv_dgrad(u, dy, u.x.grad, accumulate=False, res=xres, mask=u.x)

        def dgrad():
            u.bn.eval()
            h._prepare_unit(u, True)
            h._conv_dgrad(u, dy, u.x.grad, accumulate=False, mask=u.x)

        for mode in modes:
            fn = {"fwd_frozen": fwd_frozen, "fwd_stats": fwd_stats, "dgrad": dgrad, "fwd_plain": fwd_plain,
                  "dgrad_res": dgrad_res}[mode]
            if mode == "dgrad_res" and (s != (1, 1)):
                continue
            fn()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5):
                fn()
            e1.record()
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) / 5 * 1e3
            cnt.zero_()
            cnt[11] = 2 ** 62
            L.call("asf_tapgemm_probe", L.ptr(cnt))
            fn()
            torch.cuda.synchronize()
            L.call("asf_tapgemm_probe", None)
            c = cnt.cpu().tolist()
            n = max(c[8], 1)
            print("%-20s %-10s %7.1f | %s   (CTAs %d)  entry->body %.2f us, first entry->last exit %.2f us"
                  % (name, mode, us, " ".join("%15.1f" % (c[i] / n / 1e3) for i in range(8)), c[8], c[9] / n / 1e3,
                     (c[12] - c[11]) / 1e3))


if __name__ == "__main__":
    main()

"""GPU aid: stem weight-gradient kernel (asf_stem_wgrad) timing and agreement with torch on the two stems of the
network at the bench batch.  Usage: python tools/stem_wgrad_bench.py [batch]"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "auditory-slow-fast_b200")):
    sys.path.insert(0, p)
import torch
import torch.nn.functional as F
from asf_b200 import _lib as L

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
for name, T, cout, k, s, p in (("slow", 100, 64, (1, 7), (2, 2), (0, 3)), ("fast", 400, 8, (5, 7), (2, 2), (2, 3))):
    torch.manual_seed(0)
    x = torch.randn(B, 1, T, 128, device="cuda")
    To, Fo = (T + 2 * p[0] - k[0]) // s[0] + 1, (128 + 2 * p[1] - k[1]) // s[1] + 1
    dy = torch.randn(B, To, Fo, cout, device="cuda").to(torch.bfloat16)
    g = L.ConvGeom()
    g.B, g.Tin, g.Fin, g.Cin, g.Tout, g.Fout, g.Cout = B, T, 128, 1, To, Fo, cout
    g.kt, g.kf, g.st, g.sf, g.pt, g.pf = k[0], k[1], s[0], s[1], p[0], p[1]
    g.ld_in, g.ld_out = 1, cout
    dw = torch.zeros(cout, 1, k[0], k[1], device="cuda")
    st = L.stream_ptr()

    def run():
        L.call("asf_stem_wgrad", L.ptr(x), L.ptr(dy), L.BF16, L.ptr(dw), C.byref(g), 0, st)

    run()
    torch.cuda.synchronize()
    w = torch.zeros(cout, 1, k[0], k[1], device="cuda", requires_grad=True)
    F.conv2d(x, w, stride=s, padding=p).backward(dy.float().permute(0, 3, 1, 2))
    err = float((dw - w.grad).norm() / w.grad.norm())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(3):
        run()
    e0.record()
    for _ in range(20):
        run()
    e1.record()
    torch.cuda.synchronize()
    print("%s stem: rel err vs torch %.2e, %.1f us per launch" % (name, err, e0.elapsed_time(e1) * 1e3 / 20))
    assert err < 2e-3

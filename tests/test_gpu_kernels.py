"""GPU (-m gpu): every CUDA kernel through the C ABI against a plain PyTorch fp32 reference of the same op
on the same (rounded) inputs.  Conv classes follow SURVEY.md 8(a): 1x1, 1x1/s(1,2), 1x3, 1x3/s(1,2), 3x1,
7x1/s(4,1) at the channel counts / spatial sizes of the real network (incl. ragged tiles and channel
counts that are not multiples of the 64-wide K chunk or the 16-wide N chunk)."""
import ctypes as C
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from asf_b200 import _lib as L  # noqa: E402
from asf_b200.engine import Act  # noqa: E402
from gpu_util import Harness, make_unit, max_err, nchw, nhwc, rel_err  # noqa: E402

# (name, B, T, F, cin, cout, kernel, stride, pad)
CONV_CASES = [
    ("1x1_s2a_slow", 2, 25, 32, 80, 64, (1, 1), (1, 1), (0, 0)),
    ("1x1_c_slow_s5", 3, 25, 4, 512, 2048, (1, 1), (1, 1), (0, 0)),
    ("1x1_fast_c8", 2, 100, 32, 8, 32, (1, 1), (1, 1), (0, 0)),
    ("1x1_fast_to8", 2, 100, 32, 32, 8, (1, 1), (1, 1), (0, 0)),
    ("1x1s2_short_slow", 2, 25, 32, 320, 512, (1, 1), (1, 2), (0, 0)),
    ("1x1s2_short_fast", 2, 100, 8, 128, 256, (1, 1), (1, 2), (0, 0)),
    ("1x3_slow_s3", 2, 25, 16, 128, 128, (1, 3), (1, 1), (0, 1)),
    ("1x3_slow_s5_F4", 3, 25, 4, 512, 512, (1, 3), (1, 1), (0, 1)),
    ("1x3_fast_c8", 2, 100, 32, 8, 8, (1, 3), (1, 1), (0, 1)),
    ("1x3s2_slow_s4", 2, 25, 16, 256, 256, (1, 3), (1, 2), (0, 1)),
    ("1x3s2_fast_s5", 2, 100, 8, 64, 64, (1, 3), (1, 2), (0, 1)),
    ("3x1_slow_s4", 2, 25, 8, 1024, 256, (3, 1), (1, 1), (1, 0)),
    ("3x1_slow_s5", 3, 25, 4, 1280, 512, (3, 1), (1, 1), (1, 0)),
    ("3x1_fast_s2", 2, 100, 32, 32, 8, (3, 1), (1, 1), (1, 0)),
    ("3x1_fast_s4", 1, 100, 8, 128, 32, (3, 1), (1, 1), (1, 0)),
    ("7x1s4_lat_s1", 2, 100, 32, 8, 16, (7, 1), (4, 1), (3, 0)),
    ("7x1s4_lat_s4", 2, 100, 8, 128, 256, (7, 1), (4, 1), (3, 0)),
]
IDS = [c[0] for c in CONV_CASES]

# The same conv classes at the BENCHMARKED size (64 clips per GPU): M = 6 400 ... 204 800 rows, several tiles per
# persistent CTA, split-K weight gradients over many CTAs, the largest layers of both pathways.
BENCH_CASES = [
    ("B64_3x1_slow_s4a", 64, 25, 8, 1024, 256, (3, 1), (1, 1), (1, 0)),
    ("B64_3x1_slow_s5a", 64, 25, 4, 2048, 512, (3, 1), (1, 1), (1, 0)),
    ("B64_1x1_slow_s5c", 64, 25, 4, 512, 2048, (1, 1), (1, 1), (0, 0)),
    ("B64_1x3_slow_s4b", 64, 25, 8, 256, 256, (1, 3), (1, 1), (0, 1)),
    ("B64_1x3s2_slow_s5b", 64, 25, 8, 512, 512, (1, 3), (1, 2), (0, 1)),
    ("B64_1x1s2_short_s5", 64, 25, 8, 1280, 2048, (1, 1), (1, 2), (0, 0)),
    ("B64_1x1_slow_s2a", 64, 25, 32, 80, 64, (1, 1), (1, 1), (0, 0)),
    ("B64_7x1s4_lat_s4", 64, 100, 8, 128, 256, (7, 1), (4, 1), (3, 0)),
    ("B64_3x1_fast_s2a", 64, 100, 32, 32, 8, (3, 1), (1, 1), (1, 0)),
    ("B64_1x3_fast_s3b", 64, 100, 16, 16, 16, (1, 3), (1, 1), (0, 1)),
    ("B64_1x1_fast_s5c", 64, 100, 4, 64, 256, (1, 1), (1, 1), (0, 0)),
]
BENCH_IDS = [c[0] for c in BENCH_CASES]


def _ref_conv(x, w, s, p):
    return F.conv2d(x, w, None, stride=s, padding=p)


@pytest.fixture(scope="module")
def h_tc():
    return Harness(torch.bfloat16)


@pytest.fixture(scope="module")
def h_f32():
    return Harness(torch.float32)


@pytest.mark.parametrize("case", CONV_CASES + BENCH_CASES, ids=IDS + BENCH_IDS)
def test_tapgemm_forward_fused_affine(h_tc, case):
    """frozen/eval BN: conv with BN scale folded into the weights, shift + residual + ReLU in the epilogue."""
    _, B, T, Fq, cin, cout, k, s, p = case
    u, x, w = make_unit(h_tc, B, T, Fq, cin, cout, k, s, p, seed=1)
    u.bn.eval()
    res = Act(u.y.B, u.y.T, u.y.F, cout, h_tc.dtype, h_tc.device)
    res.buf.copy_(torch.randn(res.buf.shape, device="cuda"))
    u.res = res
    h_tc._unit_forward(u, training=False)
    torch.cuda.synchronize()
    scale = u.bn.weight / torch.sqrt(u.bn.running_var + u.bn.eps)
    shift = u.bn.bias - u.bn.running_mean * scale
    w_seen = (w * scale.view(-1, 1, 1, 1)).to(torch.bfloat16).float()
    ref = _ref_conv(x, w_seen, s, p) + shift.view(1, -1, 1, 1) + nchw(res.view().float())
    ref = F.relu(ref)
    out = nchw(u.y.view().float())
    tol = 2e-2 * float(ref.abs().max())  # bf16 output rounding (2^-9 relative) dominates
    assert max_err(out, ref) <= tol, (max_err(out, ref), tol)
    assert rel_err(out, ref) < 6e-3, rel_err(out, ref)


_STAT_CASES = [CONV_CASES[i] for i in (0, 4, 6, 9, 11, 13, 15, 16)] + [BENCH_CASES[i] for i in (0, 2, 5, 8, 10)]


@pytest.mark.parametrize("case", _STAT_CASES, ids=[c[0] for c in _STAT_CASES])
def test_tapgemm_forward_batch_stats(h_tc, case):
    """training BN: raw conv output + per-channel sum / sum-of-squares from the epilogue (ordered two-stage reduction in
    double, bit-reproducible), then normalise."""
    _, B, T, Fq, cin, cout, k, s, p = case
    u, x, w = make_unit(h_tc, B, T, Fq, cin, cout, k, s, p, seed=2)
    u.bn.train()
    rm0, rv0 = u.bn.running_mean.clone(), u.bn.running_var.clone()
    h_tc._unit_forward(u, training=True)
    torch.cuda.synchronize()
    stats1, y1 = u.stats.clone(), u.y.buf.clone()
    rm1, rv1 = u.bn.running_mean.clone(), u.bn.running_var.clone()
    u.stats.fill_(-1.0)
    h_tc._unit_forward(u, training=True)  # a second launch on the same workspace / tickets: same bits
    torch.cuda.synchronize()
    assert torch.equal(u.stats, stats1) and torch.equal(u.y.buf, y1)
    assert int(u.stats_ticket.abs().sum()) == 0, "the tickets reset themselves"
    with torch.no_grad():  # undo the second running-statistics update
        u.bn.running_mean.copy_(rm1)
        u.bn.running_var.copy_(rv1)
        u.bn.num_batches_tracked.fill_(1)
    stored = u.raw.view().double()  # the bf16 values BN reads: the sums are over exactly these
    # (fp32 partial sums inside a CTA, double across CTAs: ~1e-6 of the sum of magnitudes)
    assert torch.allclose(stats1[:cout], stored.sum((0, 1, 2)), rtol=0, atol=1e-5 * float(stored.abs().sum((0, 1, 2)).max()))
    assert torch.allclose(stats1[cout:], (stored * stored).sum((0, 1, 2)), rtol=1e-5, atol=0)
    conv = _ref_conv(x, w.to(torch.bfloat16).float(), s, p)
    mean = conv.mean((0, 2, 3))
    var = conv.var((0, 2, 3), unbiased=False)
    n = conv.numel() / cout
    assert torch.allclose(u.mean, mean, atol=2e-3 * float(conv.abs().max()))
    assert torch.allclose(u.invstd, 1.0 / torch.sqrt(var + u.bn.eps), rtol=5e-3)
    ref = F.relu(F.batch_norm(conv, None, None, u.bn.weight, u.bn.bias, True, 0.1, u.bn.eps))
    out = nchw(u.y.view().float())
    assert rel_err(out, ref) < 1e-2, rel_err(out, ref)
    assert torch.allclose(u.bn.running_mean, 0.9 * rm0 + 0.1 * mean, atol=1e-3)
    assert torch.allclose(u.bn.running_var, 0.9 * rv0 + 0.1 * var * n / (n - 1), rtol=1e-2, atol=1e-4)
    assert int(u.bn.num_batches_tracked) == 1


@pytest.mark.parametrize("case", CONV_CASES + BENCH_CASES, ids=IDS + BENCH_IDS)
def test_tapgemm_dgrad(h_tc, case):
    _, B, T, Fq, cin, cout, k, s, p = case
    u, x, w = make_unit(h_tc, B, T, Fq, cin, cout, k, s, p, seed=3)
    u.bn.eval()
    h_tc._prepare_unit(u, training=True)  # frozen-BN training: folded weights
    dy = Act(u.y.B, u.y.T, u.y.F, cout, h_tc.dtype, h_tc.device)
    dy.buf.copy_(torch.randn(dy.buf.shape, device="cuda"))
    dx = u.x.grad
    prev = torch.randn(dx.buf.shape, device="cuda").to(h_tc.dtype)
    dx.buf.copy_(prev)
    # a strided 1x1 leaves odd input columns untouched, so in the network it is never the last (masking) writer
    covered = not (k == (1, 1) and s != (1, 1))
    h_tc._conv_dgrad(u, dy, dx, accumulate=True, mask=u.x if covered else None)
    torch.cuda.synchronize()
    scale = u.bn.weight / torch.sqrt(u.bn.running_var + u.bn.eps)
    w_seen = (w * scale.view(-1, 1, 1, 1)).to(torch.bfloat16).float()
    xr = x.clone().requires_grad_(True)
    _ref_conv(xr, w_seen, s, p).backward(nchw(dy.view().float()))
    ref = xr.grad + nchw(prev.float())
    if covered:
        ref = ref * (x > 0)
    out = nchw(dx.view().float())
    assert rel_err(out, ref) < 1e-2, rel_err(out, ref)


@pytest.mark.parametrize("case", CONV_CASES + BENCH_CASES, ids=IDS + BENCH_IDS)
def test_tapgemm_wgrad(h_tc, case):
    _, B, T, Fq, cin, cout, k, s, p = case
    u, x, w = make_unit(h_tc, B, T, Fq, cin, cout, k, s, p, seed=4)
    dy = Act(u.y.B, u.y.T, u.y.F, cout, h_tc.dtype, h_tc.device)
    dy.buf.copy_(torch.randn(dy.buf.shape, device="cuda"))
    g = torch.zeros_like(w)
    rscale = 0.5 + torch.rand(cout, device="cuda")
    h_tc._conv_wgrad(u, dy, rscale, g, accumulate=False)
    torch.cuda.synchronize()
    wr = w.clone().requires_grad_(True)
    _ref_conv(x, wr, s, p).backward(nchw(dy.view().float()))
    ref = wr.grad * rscale.view(-1, 1, 1, 1)
    assert rel_err(g, ref) < 5e-3, rel_err(g, ref)
    # accumulate into an existing gradient
    h_tc._conv_wgrad(u, dy, rscale, g, accumulate=True)
    torch.cuda.synchronize()
    assert rel_err(g, 2 * ref) < 5e-3


# Narrow layers with >= 16384 output positions take the CUDA-core weight-gradient kernel (csrc/narrow_wgrad.cuh): same
# entry point, same partial-tile workspace, every conv class of the fast pathway incl. strides, laterals and a
# concat-slice input pitch.
NARROW_CASES = [
    ("1x3_c8", 8, 100, 32, 8, 8, (1, 3), (1, 1), (0, 1), None),
    ("3x1_32to8", 8, 100, 32, 32, 8, (3, 1), (1, 1), (1, 0), None),
    ("1x1_8to32", 8, 100, 32, 8, 32, (1, 1), (1, 1), (0, 0), None),
    ("1x3_c16", 16, 100, 16, 16, 16, (1, 3), (1, 1), (0, 1), None),
    ("1x1_64to16", 16, 100, 16, 64, 16, (1, 1), (1, 1), (0, 0), None),
    ("3x1_16to32_pitch80", 16, 100, 16, 16, 32, (3, 1), (1, 1), (1, 0), 80),
    ("1x1s2_32to64", 16, 100, 32, 32, 64, (1, 1), (1, 2), (0, 0), None),
    ("1x3s2_c16", 12, 100, 32, 16, 16, (1, 3), (1, 2), (0, 1), None),
    ("7x1s4_lat_8to16", 24, 100, 32, 8, 16, (7, 1), (4, 1), (3, 0), None),
]


@pytest.mark.parametrize("case", NARROW_CASES, ids=[c[0] for c in NARROW_CASES])
def test_narrow_wgrad(h_tc, case):
    _, B, T, Fq, cin, cout, k, s, p, ld_in = case
    u, x, w = make_unit(h_tc, B, T, Fq, cin, cout, k, s, p, seed=6, ld_in=ld_in)
    dy = Act(u.y.B, u.y.T, u.y.F, cout, h_tc.dtype, h_tc.device)
    dy.buf.copy_(torch.randn(dy.buf.shape, device="cuda"))
    g = torch.zeros_like(w)
    h_tc._conv_wgrad(u, dy, None, g, accumulate=False)
    torch.cuda.synchronize()
    rows = u.y.B * u.y.T * u.y.F
    assert u.wg_ksplit == -(-rows // max(512, -(-rows // (2 * 148 * 128)) * 128)), \
        "expected the CUDA-core narrow-layer kernel's CTA count as ksplit, got %d" % u.wg_ksplit
    wr = w.clone().requires_grad_(True)
    _ref_conv(x, wr, s, p).backward(nchw(dy.view().float()))
    assert rel_err(g, wr.grad) < 5e-3, rel_err(g, wr.grad)
    h_tc._conv_wgrad(u, dy, None, g, accumulate=True)
    torch.cuda.synchronize()
    assert rel_err(g, 2 * wr.grad) < 5e-3


def _narrow_count():
    from asf_b200 import _lib as L
    return int(L.lib().asf_narrow_conv_launches())


@pytest.mark.parametrize("case", NARROW_CASES, ids=[c[0] for c in NARROW_CASES])
def test_narrow_conv_forward_and_dgrad(h_tc, case, monkeypatch):
    """csrc/narrow_conv.cuh behind asf_tapgemm mode 0: fused-affine forward (+ residual, ReLU), batch-statistics
    forward and dgrad (+ accumulate, ReLU mask) against torch, and against the tensor-core kernel on the same inputs
    (both accumulate in fp32 and round once to bf16, so they agree to a bf16 ulp)."""
    _, B, T, Fq, cin, cout, k, s, p, ld_in = case

    def run_all(tag):
        out = {}
        u, x, w = make_unit(h_tc, B, T, Fq, cin, cout, k, s, p, seed=7, ld_in=ld_in)
        u.bn.eval()
        res = Act(u.y.B, u.y.T, u.y.F, cout, h_tc.dtype, h_tc.device)
        res.buf.copy_(torch.randn(res.buf.shape, generator=torch.Generator("cuda").manual_seed(1), device="cuda"))
        u.res = res
        h_tc._unit_forward(u, training=False)
        torch.cuda.synchronize()
        out["fwd"] = nchw(u.y.view().float()).clone()
        scale = u.bn.weight / torch.sqrt(u.bn.running_var + u.bn.eps)
        shift = u.bn.bias - u.bn.running_mean * scale
        w_seen = (w * scale.view(-1, 1, 1, 1)).to(torch.bfloat16).float()
        out["fwd_ref"] = F.relu(_ref_conv(x, w_seen, s, p) + shift.view(1, -1, 1, 1) + nchw(res.view().float()))
        # dgrad with accumulate + mask
        h_tc._prepare_unit(u, training=True)
        dy = Act(u.y.B, u.y.T, u.y.F, cout, h_tc.dtype, h_tc.device)
        dy.buf.copy_(torch.randn(dy.buf.shape, generator=torch.Generator("cuda").manual_seed(2), device="cuda"))
        dx = u.x.grad
        dx.buf.copy_(torch.randn(dx.buf.shape, generator=torch.Generator("cuda").manual_seed(3), device="cuda"))
        prev = nchw(dx.view().float()).clone()
        covered = not (k == (1, 1) and s != (1, 1))
        h_tc._conv_dgrad(u, dy, dx, accumulate=True, mask=u.x if covered else None)
        torch.cuda.synchronize()
        out["dgrad"] = nchw(dx.view().float()).clone()
        xr = x.clone().requires_grad_(True)
        _ref_conv(xr, w_seen, s, p).backward(nchw(dy.view().float()))
        ref = xr.grad + prev
        out["dgrad_ref"] = ref * (x > 0) if covered else ref
        # batch statistics
        u2, x2, w2 = make_unit(h_tc, B, T, Fq, cin, cout, k, s, p, seed=8, ld_in=ld_in)
        u2.bn.train()
        h_tc._unit_forward(u2, training=True)
        torch.cuda.synchronize()
        conv = _ref_conv(x2, w2.to(torch.bfloat16).float(), s, p)
        out["mean"], out["invstd"] = u2.mean.clone(), u2.invstd.clone()
        out["mean_ref"] = conv.mean((0, 2, 3))
        out["invstd_ref"] = 1.0 / torch.sqrt(conv.var((0, 2, 3), unbiased=False) + u2.bn.eps)
        out["bn"] = nchw(u2.y.view().float()).clone()
        out["bn_ref"] = F.relu(F.batch_norm(conv, None, None, u2.bn.weight, u2.bn.bias, True, 0.1, u2.bn.eps))
        out["amax"] = float(conv.abs().max())
        return out

    monkeypatch.setenv("ASF_NARROW_CONV", "1")
    n0 = _narrow_count()
    a = run_all("narrow")
    taken = _narrow_count() - n0
    assert taken >= 3, "forward, dgrad and batch-stat forward should take the CUDA-core narrow kernel (%d did)" % taken
    monkeypatch.setenv("ASF_NARROW_CONV", "0")
    n1 = _narrow_count()
    b = run_all("tensor-core")
    assert _narrow_count() == n1
    assert rel_err(a["fwd"], a["fwd_ref"]) < 6e-3
    assert max_err(a["fwd"], a["fwd_ref"]) <= 2e-2 * float(a["fwd_ref"].abs().max())
    assert rel_err(a["dgrad"], a["dgrad_ref"]) < 1e-2
    assert torch.allclose(a["mean"], a["mean_ref"], atol=2e-3 * a["amax"])
    assert torch.allclose(a["invstd"], a["invstd_ref"], rtol=5e-3)
    assert rel_err(a["bn"], a["bn_ref"]) < 1e-2
    # the two kernels differ only in fp32 summation order before one bf16 rounding
    for key in ("fwd", "dgrad", "bn"):
        assert rel_err(a[key], b[key]) < 3e-3, (key, rel_err(a[key], b[key]))
    assert torch.allclose(a["mean"], b["mean"], atol=1e-4 * a["amax"])


def test_narrow_conv_dgrad_k128(h_tc, monkeypatch):
    """dgrad of a 1x1 expansion conv 32 -> 128 channels (fast pathway s4, conv c): contraction over 128 channels into
    32, the widest K the narrow kernel takes; against torch and against the tensor-core kernel."""
    B, T, Fq, cin, cout = 16, 100, 16, 32, 128
    got = {}
    for tag, env in (("narrow", "1"), ("tc", "0")):
        monkeypatch.setenv("ASF_NARROW_CONV", env)
        u, x, w = make_unit(h_tc, B, T, Fq, cin, cout, (1, 1), (1, 1), (0, 0), seed=9)
        u.bn.eval()
        h_tc._prepare_unit(u, training=True)
        dy = Act(u.y.B, u.y.T, u.y.F, cout, h_tc.dtype, h_tc.device)
        dy.buf.copy_(torch.randn(dy.buf.shape, generator=torch.Generator("cuda").manual_seed(2), device="cuda"))
        dx = u.x.grad
        n0 = _narrow_count()
        h_tc._conv_dgrad(u, dy, dx, accumulate=False, mask=u.x)
        torch.cuda.synchronize()
        assert _narrow_count() - n0 == (1 if env == "1" else 0)
        got[tag] = nchw(dx.view().float()).clone()
        if tag == "narrow":
            scale = u.bn.weight / torch.sqrt(u.bn.running_var + u.bn.eps)
            w_seen = (w * scale.view(-1, 1, 1, 1)).to(torch.bfloat16).float()
            xr = x.clone().requires_grad_(True)
            _ref_conv(xr, w_seen, (1, 1), (0, 0)).backward(nchw(dy.view().float()))
            ref = xr.grad * (x > 0)
    assert rel_err(got["narrow"], ref) < 1e-2, rel_err(got["narrow"], ref)
    assert rel_err(got["narrow"], got["tc"]) < 3e-3, rel_err(got["narrow"], got["tc"])


def test_tapgemm_concat_slice_output(h_tc):
    """the lateral conv writes into channels [64:80) of the slow buffer (no torch.cat)."""
    u, x, w = make_unit(h_tc, 2, 100, 32, 8, 16, (7, 1), (4, 1), (3, 0), seed=5, ld_out=80)
    u.bn.eval()
    u.y.buf.fill_(7.0)
    h_tc._unit_forward(u, training=False)
    torch.cuda.synchronize()
    assert float((u.y.buf[..., :64].float() - 7.0).abs().max()) == 0.0
    scale = u.bn.weight / torch.sqrt(u.bn.running_var + u.bn.eps)
    shift = u.bn.bias - u.bn.running_mean * scale
    ref = F.relu(_ref_conv(x, (w * scale.view(-1, 1, 1, 1)).to(torch.bfloat16).float(), (4, 1), (3, 0)) + shift.view(1, -1, 1, 1))
    assert rel_err(nchw(u.y.view().float()), ref) < 6e-3


@pytest.mark.parametrize("case", [CONV_CASES[i] for i in (0, 4, 9, 11, 15)], ids=[IDS[i] for i in (0, 4, 9, 11, 15)])
def test_simt_conv_fp32(h_f32, case):
    """CUDA-core fp32 parity path: forward, dgrad, wgrad against torch to float rounding."""
    _, B, T, Fq, cin, cout, k, s, p = case
    u, x, w = make_unit(h_f32, B, T, Fq, cin, cout, k, s, p, seed=6)
    u.bn.eval()
    h_f32._unit_forward(u, training=False)
    scale = u.bn.weight / torch.sqrt(u.bn.running_var + u.bn.eps)
    shift = u.bn.bias - u.bn.running_mean * scale
    scale, shift = scale.detach(), shift.detach()
    ws = w * scale.view(-1, 1, 1, 1)
    ref = F.relu(_ref_conv(x, ws, s, p) + shift.view(1, -1, 1, 1))
    assert rel_err(nchw(u.y.view()), ref) < 1e-5
    dy = Act(u.y.B, u.y.T, u.y.F, cout, h_f32.dtype, h_f32.device)
    dy.buf.copy_(torch.randn(dy.buf.shape, device="cuda"))
    h_f32._conv_dgrad(u, dy, u.x.grad, accumulate=False)
    g = torch.zeros_like(w)
    h_f32._conv_wgrad(u, dy, u.scale, g, accumulate=False)
    torch.cuda.synchronize()
    xr, wr = x.clone().requires_grad_(True), ws.clone().requires_grad_(True)
    _ref_conv(xr, wr, s, p).backward(nchw(dy.view()))
    assert rel_err(nchw(u.x.grad.view()), xr.grad) < 1e-5
    assert rel_err(g, wr.grad * scale.view(-1, 1, 1, 1)) < 1e-4


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_bn_stats_apply_backward(dtype):
    rows, Cn = 3000, 64
    g = torch.Generator().manual_seed(0)
    x = (torch.randn(rows, Cn, generator=g) * 2 + 0.5).cuda().to(dtype)
    dz = torch.randn(rows, Cn, generator=g).cuda().to(dtype)
    gamma = (0.5 + torch.rand(Cn, generator=g)).cuda()
    beta = torch.randn(Cn, generator=g).cuda()
    st = L.stream_ptr()
    code = L.dtype_code(dtype)
    # statistics come out as doubles from an ordered two-stage reduction: garbage in, no zeroing needed, bit-reproducible
    stats = torch.full((2 * Cn,), 123.0, dtype=torch.float64, device="cuda")
    ws = torch.empty(L.stats_ws_floats(Cn), device="cuda")
    tk = torch.zeros(L.STATS_TICKETS, dtype=torch.int32, device="cuda")
    L.call("asf_channel_stats", L.ptr(x), code, rows, Cn, Cn, L.ptr(stats), L.ptr(ws), L.ptr(tk), st)
    stats_again = torch.empty_like(stats)
    L.call("asf_channel_stats", L.ptr(x), code, rows, Cn, Cn, L.ptr(stats_again), L.ptr(ws), L.ptr(tk), st)
    xd = x.double()
    assert torch.equal(stats, stats_again) and int(tk.abs().sum()) == 0
    assert torch.allclose(stats[:Cn], xd.sum(0), rtol=1e-6 if dtype == torch.bfloat16 else 1e-12)
    assert torch.allclose(stats[Cn:], (xd * xd).sum(0), rtol=1e-6 if dtype == torch.bfloat16 else 1e-12)
    scale, shift, mean, invstd = (torch.empty(Cn, device="cuda") for _ in range(4))
    rm, rv, nbt = torch.zeros(Cn, device="cuda"), torch.ones(Cn, device="cuda"), torch.zeros((), dtype=torch.int64, device="cuda")
    L.call("asf_bn_finalize", L.ptr(stats), Cn, C.c_double(rows), L.ptr(gamma), L.ptr(beta), C.c_float(1e-5), C.c_float(0.1),
           L.ptr(rm), L.ptr(rv), L.ptr(nbt), L.ptr(scale), L.ptr(shift), L.ptr(mean), L.ptr(invstd), st)
    y = torch.empty_like(x)
    L.call("asf_bn_apply", L.ptr(x), Cn, None, 0, L.ptr(y), Cn, code, rows, Cn, L.ptr(scale), L.ptr(shift), 1, st)
    xf = x.float().requires_grad_(True)
    ref = F.relu(F.batch_norm(xf, None, None, gamma, beta, True, 0.1, 1e-5))
    tol = 1e-5 if dtype == torch.float32 else 1e-2
    assert rel_err(y, ref) < tol
    assert torch.allclose(mean, xf.mean(0), atol=1e-4) and int(nbt) == 1
    # backward (no relu): dx
    ref2 = F.batch_norm(xf, None, None, gamma, beta, True, 0.1, 1e-5)
    ref2.backward(dz.float())
    red = torch.full((2 * Cn,), -5.0, dtype=torch.float64, device="cuda")
    L.call("asf_bn_bwd_reduce", L.ptr(dz), Cn, L.ptr(x), Cn, code, rows, Cn, L.ptr(mean), L.ptr(invstd), L.ptr(red),
           L.ptr(ws), L.ptr(tk), st)
    red_again = torch.empty_like(red)
    L.call("asf_bn_bwd_reduce", L.ptr(dz), Cn, L.ptr(x), Cn, code, rows, Cn, L.ptr(mean), L.ptr(invstd), L.ptr(red_again),
           L.ptr(ws), L.ptr(tk), st)
    assert torch.equal(red, red_again)
    dx, dg, db = torch.empty_like(x), torch.empty(Cn, device="cuda"), torch.empty(Cn, device="cuda")
    L.call("asf_bn_bwd_apply", L.ptr(dz), Cn, L.ptr(x), Cn, L.ptr(dx), Cn, code, rows, Cn, L.ptr(gamma), L.ptr(mean),
           L.ptr(invstd), L.ptr(red), C.c_double(rows), L.ptr(dg), L.ptr(db), 0, st)
    torch.cuda.synchronize()
    assert rel_err(dx, xf.grad) < (1e-4 if dtype == torch.float32 else 1e-2)
    assert torch.allclose(db, dz.float().sum(0), rtol=1e-3, atol=1e-2)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_stem_pool_forward_backward(dtype):
    B, T, Fq, Cn = 2, 50, 64, 8
    g = torch.Generator().manual_seed(1)
    x = torch.randn(B, Cn, T, Fq, generator=g).cuda()
    xa = nhwc(x).to(dtype)
    scale = (0.5 + torch.rand(Cn, generator=g)).cuda()
    shift = (0.3 * torch.randn(Cn, generator=g)).cuda()
    To, Fo = 25, 32
    y = torch.empty(B, To, Fo, Cn, dtype=dtype, device="cuda")
    st = L.stream_ptr()
    code = L.dtype_code(dtype)
    L.call("asf_bn_relu_maxpool_fwd", L.ptr(xa), L.ptr(y), Cn, code, B, T, Fq, Cn, L.ptr(scale), L.ptr(shift), None, st)
    xs = nchw(xa.float()).requires_grad_(True)
    ref = F.max_pool2d(F.relu(xs * scale.view(1, -1, 1, 1) + shift.view(1, -1, 1, 1)), 3, 2, 1)
    assert rel_err(nchw(y.float()), ref) < (1e-6 if dtype == torch.float32 else 4e-3)
    dy = torch.randn(B, To, Fo, Cn, generator=g).cuda().to(dtype)
    dz = torch.empty_like(xa)
    L.call("asf_bn_relu_maxpool_bwd", L.ptr(xa), L.ptr(dy), Cn, L.ptr(dz), code, B, T, Fq, Cn, L.ptr(scale), L.ptr(shift),
           None, st)
    torch.cuda.synchronize()
    ref.backward(nchw(dy.float()))
    if dtype == torch.bfloat16:
        # the production path: arg-max codes written by the forward, the backward reads only codes + dy -- bit-equal
        # to the recomputing kernel (same first-maximum rule)
        am = torch.empty(B, To, Fo, Cn // 8, dtype=torch.int32, device="cuda")
        y2, dz2 = torch.empty_like(y), torch.empty_like(dz)
        L.call("asf_bn_relu_maxpool_fwd", L.ptr(xa), L.ptr(y2), Cn, code, B, T, Fq, Cn, L.ptr(scale), L.ptr(shift),
               L.ptr(am), st)
        L.call("asf_bn_relu_maxpool_bwd", None, L.ptr(dy), Cn, L.ptr(dz2), code, B, T, Fq, Cn, None, None, L.ptr(am), st)
        torch.cuda.synchronize()
        assert torch.equal(y2, y) and torch.equal(dz2, dz)
    # dz is the gradient w.r.t. the BN output (before the scale): compare with grad_x / scale
    refdz = xs.grad / scale.view(1, -1, 1, 1)
    assert rel_err(nchw(dz.float()), refdz) < (1e-5 if dtype == torch.float32 else 1e-2)


@pytest.mark.parametrize("name,T,cout,k,s,p", [("slow", 100, 64, (1, 7), (2, 2), (0, 3)),
                                               ("fast", 400, 8, (5, 7), (2, 2), (2, 3)),
                                               ("ragged", 37, 16, (3, 5), (2, 2), (1, 2)),
                                               ("wide", 21, 8, (5, 7), (2, 2), (2, 3))])
def test_stem_wgrad(name, T, cout, k, s, p):
    """asf_stem_wgrad (C_in = 1 stems, stem_helper.py:150-157) against torch's conv weight gradient, overwrite and
    accumulate; 'ragged' has row counts that do not divide the warp / slice tiling."""
    B, Fq = 3, {"ragged": 50, "wide": 192}.get(name, 128)  # wide: 96 output columns = three 32-lane steps per line
    torch.manual_seed(3)
    x = torch.randn(B, 1, T, Fq, device="cuda")
    To, Fo = (T + 2 * p[0] - k[0]) // s[0] + 1, (Fq + 2 * p[1] - k[1]) // s[1] + 1
    dy = torch.randn(B, To, Fo, cout, device="cuda").to(torch.bfloat16)
    g = L.ConvGeom()
    g.B, g.Tin, g.Fin, g.Cin, g.Tout, g.Fout, g.Cout = B, T, Fq, 1, To, Fo, cout
    g.kt, g.kf, g.st, g.sf, g.pt, g.pf = k[0], k[1], s[0], s[1], p[0], p[1]
    g.ld_in, g.ld_out = 1, cout
    dw = torch.full((cout, 1, k[0], k[1]), 7.0, device="cuda")
    st = L.stream_ptr()
    ks = L.lib().asf_stem_wgrad_ksplit(C.byref(g), L.BF16)
    assert ks >= 1
    n = cout * k[0] * k[1]
    dwp = torch.full((ks, n), float("nan"), device="cuda")  # every partial row is written, none accumulated into
    L.call("asf_stem_wgrad", L.ptr(x), L.ptr(dy), L.BF16, L.ptr(dwp), C.byref(g), st)
    L.call("asf_unpack_wgrad", L.ptr(dwp), L.ptr(dw), 1, n, 1, None, 0, ks, st)
    torch.cuda.synchronize()
    w = torch.zeros(cout, 1, k[0], k[1], device="cuda", requires_grad=True)
    F.conv2d(x, w, stride=s, padding=p).backward(nchw(dy.float()))
    assert rel_err(dw, w.grad) < 1e-4, rel_err(dw, w.grad)
    dwp2 = torch.empty_like(dwp)
    L.call("asf_stem_wgrad", L.ptr(x), L.ptr(dy), L.BF16, L.ptr(dwp2), C.byref(g), st)
    L.call("asf_unpack_wgrad", L.ptr(dwp2), L.ptr(dw), 1, n, 1, None, 1, ks, st)
    torch.cuda.synchronize()
    assert torch.equal(dwp, dwp2), "the stem weight gradient must be bit-reproducible (no atomics)"
    assert rel_err(dw, 2 * w.grad) < 1e-4
    # the fused form: input lines staged in shared memory, the cross-block sum folded into the kernel
    rows = L.lib().asf_stem_wgrad_fused_rows(C.byref(g), L.BF16)
    assert rows >= 1
    ws = torch.full((rows, n), float("nan"), device="cuda")
    tk = torch.zeros(1, dtype=torch.int32, device="cuda")
    dwf = torch.full_like(dw, 5.0)
    L.call("asf_stem_wgrad_fused", L.ptr(x), L.ptr(dy), L.BF16, L.ptr(ws), L.ptr(dwf), 0, L.ptr(tk), C.byref(g), st)
    dwf1 = dwf.clone()
    L.call("asf_stem_wgrad_fused", L.ptr(x), L.ptr(dy), L.BF16, L.ptr(ws), L.ptr(dwf), 1, L.ptr(tk), C.byref(g), st)
    dwf2 = torch.full_like(dw, -1.0)
    L.call("asf_stem_wgrad_fused", L.ptr(x), L.ptr(dy), L.BF16, L.ptr(ws), L.ptr(dwf2), 0, L.ptr(tk), C.byref(g), st)
    torch.cuda.synchronize()
    assert int(tk) == 0
    assert rel_err(dwf1, w.grad) < 1e-4, rel_err(dwf1, w.grad)
    assert rel_err(dwf, 2 * w.grad) < 1e-4
    assert torch.equal(dwf1, dwf2), "the fused stem weight gradient must be bit-reproducible"
    # fp32 dout (parity mode): the scalar kernel, double accumulators
    dyf = dy.float()
    ksf = L.lib().asf_stem_wgrad_ksplit(C.byref(g), L.F32)
    dwpf = torch.empty((ksf, n), device="cuda")
    L.call("asf_stem_wgrad", L.ptr(x), L.ptr(dyf), L.F32, L.ptr(dwpf), C.byref(g), st)
    L.call("asf_unpack_wgrad", L.ptr(dwpf), L.ptr(dw), 1, n, 1, None, 0, ksf, st)
    torch.cuda.synchronize()
    assert rel_err(dw, w.grad) < 2e-6, rel_err(dw, w.grad)


def test_head_kernels():
    B, P, Cn = 4, 100, 256
    g = torch.Generator().manual_seed(2)
    x = torch.randn(B, P, Cn, generator=g).cuda().to(torch.bfloat16)
    feat = torch.zeros(B, 300, device="cuda")
    st = L.stream_ptr()
    L.call("asf_avgpool_fwd", L.ptr(x), L.BF16, B, P, Cn, L.ptr(feat), 300, 20, st)
    assert torch.allclose(feat[:, 20:276], x.float().mean(1), atol=1e-5)
    W = torch.randn(97, 300, generator=g).cuda()
    bias = torch.randn(97, generator=g).cuda()
    out = torch.empty(B, 97, device="cuda")
    L.call("asf_sgemm", L.ptr(feat), L.ptr(W), L.ptr(out), B, 97, 300, 0, 1, 300, 300, 97, C.c_float(1.0), C.c_float(0.0), L.ptr(bias), st)
    assert torch.allclose(out, feat @ W.t() + bias, atol=1e-4)
    dl = torch.randn(B, 97, generator=g).cuda()
    dW = torch.ones(97, 300, device="cuda")
    L.call("asf_sgemm", L.ptr(dl), L.ptr(feat), L.ptr(dW), 97, 300, B, 1, 0, 97, 300, 300, C.c_float(1.0), C.c_float(1.0), None, st)
    assert torch.allclose(dW, 1 + dl.t() @ feat, atol=1e-4)
    df = torch.zeros(B, 300, device="cuda")
    L.call("asf_sgemm", L.ptr(dl), L.ptr(W), L.ptr(df), B, 300, 97, 0, 0, 97, 300, 300, C.c_float(1.0), C.c_float(1.0), None, st)
    assert torch.allclose(df, dl @ W, atol=1e-4)
    sm = out.clone()
    L.call("asf_softmax_rows", L.ptr(sm), B, 97, 97, st)
    assert torch.allclose(sm, torch.softmax(out, 1), atol=1e-6)
    dx = torch.empty_like(x)
    L.call("asf_avgpool_bwd", L.ptr(df), 300, 20, L.ptr(x), L.ptr(dx), L.BF16, B, P, Cn, st)
    torch.cuda.synchronize()
    ref = (df[:, None, 20:276] / P).expand(B, P, Cn) * (x.float() > 0)
    assert rel_err(dx, ref) < 5e-3


def test_weight_pack_roundtrip():
    O, I, taps = 24, 40, 3
    w = torch.randn(O, I, taps, 1, device="cuda")
    rs = torch.rand(O, device="cuda") + 0.5
    p = torch.empty(O, taps, I, device="cuda")
    pt = torch.empty(I, taps, O, device="cuda")
    st = L.stream_ptr()
    L.call("asf_pack_weight", L.ptr(w), L.ptr(p), L.F32, O, I, taps, L.ptr(rs), st)
    L.call("asf_pack_weight_t", L.ptr(w), L.ptr(pt), L.F32, O, I, taps, L.ptr(rs), st)
    ws = w[..., 0] * rs.view(-1, 1, 1)
    assert torch.equal(p, ws.permute(0, 2, 1).contiguous())
    assert torch.equal(pt, ws.permute(1, 2, 0).contiguous())
    back = torch.ones_like(w)
    L.call("asf_unpack_wgrad", L.ptr(p), L.ptr(back), O, I, taps, None, 1, 1, st)
    torch.cuda.synchronize()
    assert torch.allclose(back[..., 0], 1 + ws)


def test_weight_pack_multi():
    """asf_pack_weights_multi (every conv of the network in one launch, what the training step runs after each
    optimiser step): both packed layouts, bf16 with the folded BN scale and the fp32 stem layout, ragged channel counts,
    1..7 taps through the shared-memory tiles and the 35-tap stem through the element-wise loop -- bit-exact."""
    shapes = [(8, 8, 3, True, True), (64, 1, 7, False, False), (8, 1, 35, False, False), (256, 512, 3, True, True),
              (16, 8, 7, True, False), (80, 40, 1, True, True), (97, 33, 5, True, False), (2048, 64, 1, True, True)]
    g = torch.Generator().manual_seed(4)
    arr = (L.PackDesc * len(shapes))()
    keep, chunk = [], 0
    for i, (O, I, taps, bf, scaled) in enumerate(shapes):
        w = torch.randn(O, I, taps, generator=g).cuda()
        rs = (torch.rand(O, generator=g) + 0.5).cuda() if scaled else None
        dt = torch.bfloat16 if bf else torch.float32
        pk = torch.full((O, taps, I), 9.0, dtype=dt, device="cuda")
        pt = torch.full((I, taps, O), 9.0, dtype=dt, device="cuda") if bf else None
        d = arr[i]
        d.w, d.packed, d.packed_t = w.data_ptr(), pk.data_ptr(), (pt.data_ptr() if pt is not None else None)
        d.rscale = rs.data_ptr() if rs is not None else None
        d.O, d.I, d.taps, d.dtype, d.chunk0 = O, I, taps, L.dtype_code(dt), chunk
        chunk += -(-(O * I * taps) // L.PACK_CHUNK)
        keep.append((w, rs, pk, pt, dt))
    table = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8).cuda()
    L.call("asf_pack_weights_multi", L.ptr(table), len(shapes), chunk, L.stream_ptr())
    torch.cuda.synchronize()
    for (w, rs, pk, pt, dt), shp in zip(keep, shapes):
        ws = w * rs.view(-1, 1, 1) if rs is not None else w
        assert torch.equal(pk, ws.permute(0, 2, 1).contiguous().to(dt)), shp
        if pt is not None:
            assert torch.equal(pt, ws.permute(1, 2, 0).contiguous().to(dt)), shp


def test_bad_arguments_raise():
    with pytest.raises(RuntimeError):
        L.call("asf_channel_stats", None, L.F32, 10, 8, 8, None, None, None, L.stream_ptr())
    x = torch.zeros(4, 8, device="cuda")
    with pytest.raises(RuntimeError):
        L.call("asf_channel_stats", L.ptr(x), 7, 4, 8, 8, L.ptr(x), L.ptr(x), L.ptr(x), L.stream_ptr())

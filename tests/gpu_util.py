"""Helpers shared by the -m gpu tests: run single convs through the engine's launch code."""
import ctypes as C

import torch
import torch.nn as nn
import torch.nn.functional as F

from asf_b200 import _lib as L
from asf_b200.engine import Act, ConvUnit, Engine


def nhwc(t):
    """NCHW tensor -> contiguous channels-last (B, T, F, C) copy."""
    return t.permute(0, 2, 3, 1).contiguous()


def nchw(t):
    return t.permute(0, 3, 1, 2).contiguous()


class Harness(Engine):
    """Engine without a model: enough state to launch single conv units."""

    def __init__(self, dtype, force_simt=False):
        L.require_cuda_sm100()
        self.dtype = dtype
        self.code = L.dtype_code(dtype)
        self.use_tc = dtype == torch.bfloat16
        self.force_simt = force_simt
        self.device = torch.device("cuda", torch.cuda.current_device())
        self._ws = {}
        self.grad_hook = None
        self.sync_bn = None
        self.plan = None
        self._plans = {}
        self.two_stream = False
        self.sm_split = (0, 0)
        self._side = None
        self._slow = None
        self._wg = {}
        self.wg_per_path = 1
        self._wg_rr = [0, 0]
        self.wgrad_streams = False
        self.defer_unpack = False
        self._pending_unpack = []


def make_unit(h, B, T, Fq, cin, cout, k, s, p, seed=0, ld_in=None, ld_out=None, in_scale=1.0):
    """Random conv unit on random input; returns (unit, x_nchw_fp32, w_oihw_fp32)."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    conv = nn.Conv2d(cin, cout, k, stride=s, padding=p, bias=False).cuda()
    with torch.no_grad():
        conv.weight.copy_(torch.randn(conv.weight.shape, generator=g) / (cin * k[0] * k[1]) ** 0.5)
    bn = nn.BatchNorm2d(cout).cuda()
    with torch.no_grad():
        bn.weight.copy_(0.5 + torch.rand(cout, generator=g))
        bn.bias.copy_(0.2 * torch.randn(cout, generator=g))
        bn.running_mean.copy_(0.1 * torch.randn(cout, generator=g))
        bn.running_var.copy_(0.5 + torch.rand(cout, generator=g))
    x = (in_scale * torch.randn(B, cin, T, Fq, generator=g)).cuda()
    xa = Act(B, T, Fq, cin, h.dtype, h.device) if ld_in is None else Act(B, T, Fq, ld_in, h.dtype, h.device).slice(0, cin)
    if ld_in is not None:
        xa.buf.zero_()
    xa.view().copy_(nhwc(x))
    To = (T + 2 * p[0] - k[0]) // s[0] + 1
    Fo = (Fq + 2 * p[1] - k[1]) // s[1] + 1
    if ld_out is None:
        ya = Act(B, To, Fo, cout, h.dtype, h.device)
    else:
        ya = Act(B, To, Fo, ld_out, h.dtype, h.device).slice(ld_out - cout, cout)
        ya.buf.zero_()
    u = ConvUnit("test", conv, bn, xa, ya, relu=True)
    u.is_stem = False
    u.bn_name = "bn"
    dev = h.device
    u.scale = torch.empty(cout, dtype=torch.float32, device=dev)
    u.shift = torch.empty(cout, dtype=torch.float32, device=dev)
    u.mean = torch.empty(cout, dtype=torch.float32, device=dev)
    u.invstd = torch.empty(cout, dtype=torch.float32, device=dev)
    u.stats = torch.zeros(2 * cout, dtype=torch.float64, device=dev)
    u.red = torch.zeros(2 * cout, dtype=torch.float64, device=dev)
    # what the kernels really see: the rounded input
    x_seen = nchw(xa.view().float())
    return u, x_seen, conv.weight.detach().clone()


def rel_err(a, b):
    a, b = a.float(), b.float()
    return float((a - b).norm() / (b.norm() + 1e-20))


def max_err(a, b):
    return float((a.float() - b.float()).abs().max())


def engine_decisions(model):
    """The ReLU masks and stem max-pool selections of the engine's LAST forward, in the ``forced`` format of
    oracle.model_torch.Net (NCHW masks named like the units, flat (t, f) pooling indices).  Read from the plan's
    activation buffers, which the engine keeps for its backward pass."""
    eng = model.engine
    plan = eng.plan
    forced = {}
    for u in plan.units:
        if u.relu:
            forced[u.name] = (u.y.view().float() > 0).permute(0, 3, 1, 2).contiguous()
    for u in plan.stems:
        raw = u.y.view().double()  # (B, To, Fo, C) raw conv output; BN + ReLU + max-pool are one fused kernel
        # the kernels evaluate fma(raw, scale, shift) in fp32 = the exact value rounded once: form it in double and
        # round, so that the decisions read here are the kernel's own even for pre-activations within an ulp of zero
        # (ONE flipped stem element moves the fast stem's bias gradient, a heavily cancelling sum, by 8e-3)
        a = (raw * u.scale.double().view(1, 1, 1, -1) + u.shift.double().view(1, 1, 1, -1)).float()
        forced[u.name] = (a > 0).permute(0, 3, 1, 2).contiguous()
        B, To, Fo, Cn = raw.shape
        if u.pool_argmax is not None:
            codes = u.pool_argmax  # (B, Tp, Fp, C/8) one nibble per channel: (dt+1)*3 + (df+1), 15 = nothing positive
            Tp, Fp = codes.shape[1], codes.shape[2]
            sh = torch.arange(8, device=codes.device, dtype=torch.int32) * 4
            nib = (codes.unsqueeze(-1) >> sh) & 15                       # (B, Tp, Fp, C/8, 8)
            nib = nib.reshape(B, Tp, Fp, Cn).long()
            none = nib == 15
            nib = torch.where(none, torch.full_like(nib, 4), nib)        # window centre: always inside the plane
            tp = torch.arange(Tp, device=codes.device).view(1, Tp, 1, 1)
            fp = torch.arange(Fp, device=codes.device).view(1, 1, Fp, 1)
            t = 2 * tp + nib // 3 - 1
            f = 2 * fp + nib % 3 - 1
            idx = (t * Fo + f).permute(0, 3, 1, 2).contiguous()
        else:
            act = F.relu(a).permute(0, 3, 1, 2).contiguous()
            _, idx = F.max_pool2d(act, kernel_size=[3, 3], stride=[2, 2], padding=[1, 1], return_indices=True)
        forced[u.name + ".pool"] = idx
    return forced

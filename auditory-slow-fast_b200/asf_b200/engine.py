"""The fused executor behind ``AudioSlowFast.forward`` / ``ResNet.forward`` and their backward.

What the reference does with 110 nn.Conv2d + 110 nn.BatchNorm2d + ReLU / add / cat / pool modules
(/root/reference/audio_slowfast/models/audio_model_builder.py:311-325 and the helpers it calls) is laid
out here ONCE per batch size as a static plan over pre-allocated channels-last buffers and executed as
a flat sequence of C-ABI kernel launches on the current CUDA stream:

  * every residual-stage / lateral conv  -> asf_tapgemm (tcgen05 + TMEM + TMA implicit GEMM, bf16 in,
    fp32 accumulate) with BN scale/shift, ReLU, residual add and the lateral's concat-slice write fused
    into the epilogue (frozen / eval BN), or per-channel batch statistics emitted by the epilogue followed
    by one fused normalise(+residual)(+ReLU) pass (training BN);
  * stems (C_in = 1)                     -> CUDA-core direct conv + fused BN/ReLU/max-pool;
  * head                                 -> warp-reduction average pool + CUDA-core GEMMs;
  * backward                             -> the same tap-GEMM kernel for dgrad (flipped taps, transposed
    weights, ReLU mask / gradient accumulation fused in the epilogue) and its MN-major mode for wgrad.

``Engine.run`` registers one autograd node for the whole network; its backward writes parameter gradients
straight into ``p.grad`` (views of one flat fp32 buffer that the data-parallel wrapper all-reduces in
buckets while the remaining backward is still running).

With compute dtype float32 (cfg.B200.COMPUTE_DTYPE) every conv runs on the CUDA-core direct kernels
instead: that is the fp32 parity mode (1e-3 logits criterion), not a fallback -- it needs the GPU too.
"""
from __future__ import annotations

import contextlib
import ctypes as C
import math
import os

import torch
import torch.distributed as dist
import torch.nn as nn

from . import _lib as L

BN_UNFROZEN = ("s1.pathway0_stem.bn", "s1.pathway1_stem.bn", "s1_fuse.bn")


class Act:
    """Channels-last activation: logical (B, T, F, C) living in ``buf`` (B, T, F, Ctot) at channel offset."""

    def __init__(self, B, T, F, C, dtype, device, parent=None, c_off=0, name=""):
        self.B, self.T, self.F, self.C = B, T, F, C
        self.name = name
        if parent is None:
            self.buf = torch.empty((B, T, F, C), dtype=dtype, device=device)
            self.c_off = 0
        else:
            assert (parent.B, parent.T, parent.F) == (B, T, F) and c_off + C <= parent.buf.shape[3]
            self.buf = parent.buf
            self.c_off = parent.c_off + c_off
        self.parent = parent
        self.ld = self.buf.shape[3]
        self.rows = B * T * F
        self.dtype = dtype
        self.code = L.dtype_code(dtype)
        self.esize = self.buf.element_size()
        self._grad = None
        self.grad_state = 0  # backward bookkeeping: 0 = nothing written yet this pass

    @property
    def ptr(self):
        return self.buf.data_ptr() + self.c_off * self.esize

    def row_ptr(self, row):
        return self.ptr + row * self.ld * self.esize

    def view(self):
        return self.buf[..., self.c_off:self.c_off + self.C]

    def slice(self, c_off, C, name=""):
        return Act(self.B, self.T, self.F, C, self.dtype, self.buf.device, parent=self, c_off=c_off, name=name)

    @property
    def grad(self):
        """Gradient buffer with the same geometry (slices share their parent's gradient buffer)."""
        if self._grad is None:
            if self.parent is not None:
                pg = self.parent.grad
                self._grad = Act(self.B, self.T, self.F, self.C, self.dtype, self.buf.device, parent=pg,
                                 c_off=self.c_off - self.parent.c_off, name=self.name + ".grad")
            else:
                self._grad = Act(self.B, self.T, self.F, self.C, self.dtype, self.buf.device, name=self.name + ".grad")
        return self._grad


class ConvUnit:
    """conv (+BN) (+residual) (+ReLU) with its static geometry and buffers."""

    def __init__(self, name, conv, bn, x, y, relu, res=None):
        self.name, self.conv, self.bn, self.x, self.y, self.relu, self.res = name, conv, bn, x, y, relu, res
        self.kt, self.kf = conv.kernel_size
        self.st, self.sf = conv.stride
        self.pt, self.pf = conv.padding
        self.Cin, self.Cout = conv.in_channels, conv.out_channels
        self.taps = self.kt * self.kf
        assert x.C == self.Cin and y.C == self.Cout, (name, x.C, self.Cin, y.C, self.Cout)
        To = (x.T + 2 * self.pt - self.kt) // self.st + 1
        Fo = (x.F + 2 * self.pf - self.kf) // self.sf + 1
        assert (To, Fo) == (y.T, y.F), (name, To, Fo, y.T, y.F)
        self.raw = None          # raw conv output (training-BN mode only)
        self.scale = self.shift = self.mean = self.invstd = self.stats = self.red = None
        self.w_pack = self.w_pack_t = None
        self.w_key = None
        self.batch_mode = False
        # conv class for the tap-GEMM path
        if self.kt == 1 and self.kf == 1 and self.st == 1 and self.sf == 1:
            self.axis, self.k, self.s, self.p = "flat", 1, 1, 0
        elif self.kf == 1 and self.sf == 1 and self.pf == 0:
            self.axis, self.k, self.s, self.p = "T", self.kt, self.st, self.pt
        elif self.kt == 1 and self.st == 1 and self.pt == 0:
            self.axis, self.k, self.s, self.p = "F", self.kf, self.sf, self.pf
        else:
            self.axis = None

    def geom(self):
        g = L.ConvGeom()
        g.B, g.Tin, g.Fin, g.Cin = self.x.B, self.x.T, self.x.F, self.Cin
        g.Tout, g.Fout, g.Cout = self.y.T, self.y.F, self.Cout
        g.kt, g.kf, g.st, g.sf, g.pt, g.pf = self.kt, self.kf, self.st, self.sf, self.pt, self.pf
        g.ld_in, g.ld_out = self.x.ld, self.y.ld
        return g


def _epilogue(scale=None, shift=None, res=None, mask=None, stats=None, stats_stride=0, relu=False, accumulate=False,
              row_off=0, stats_ws=None, stats_ticket=None):
    e = L.Epilogue()
    e.scale = scale.data_ptr() if scale is not None else None
    e.shift = shift.data_ptr() if shift is not None else None
    if res is not None:
        e.res, e.ld_res = res.row_ptr(row_off), res.ld
    if mask is not None:
        e.mask, e.ld_mask = mask.row_ptr(row_off), mask.ld
    if stats is not None:
        e.stats, e.stats_stride = stats.data_ptr(), stats_stride
        e.stats_ws, e.stats_ticket = stats_ws.data_ptr(), stats_ticket.data_ptr()
    e.relu, e.accumulate = int(relu), int(accumulate)
    return e


def _pow2_box(first):
    assert first <= 128 and (first & (first - 1)) == 0, "frequency extent %d must be a power of two <= 128" % first
    return (first, 128 // first, 1)


def _time_box(F, T, B):
    """128-row tile of a temporal conv's (F, T, B) position space: (F, t, b) with F * t * b = 128, powers of two,
    chosen to minimise the number of tiles.  A tile must not wrap around a clip, so with the obvious (F, 128 / F, 1)
    the T = 25 frames of the slow pathway fill 25 of 32 rows: 22 % of the tensor-core work of the largest layers
    (the 3x1 convs are half of all FLOPs) went into padding.  (8, 1, 16) -- one frame of 16 clips -- covers
    64 x 25 x 8 positions with exactly 100 tiles instead of 128."""
    assert F <= 128 and (F & (F - 1)) == 0, "frequency extent %d must be a power of two <= 128" % F
    best = None
    t = 128 // F
    while t >= 1:
        b = 128 // (F * t)
        tiles = -(-T // t) * -(-B // b)
        if best is None or tiles < best[0]:
            best = (tiles, (F, t, b))
        t //= 2
    return best[1]


def _set_m(tg, ext, box, m_to_a):
    for j in range(3):
        tg.m_ext[j], tg.m_box[j], tg.m_to_a[j] = ext[j], box[j], m_to_a[j]


def _set_a(tg, act, dims, strides):
    tg.a = act.ptr
    tg.nd = len(dims)
    for i, (d, s) in enumerate(zip(dims, strides)):
        tg.a_dim[i], tg.a_stride[i] = d, s


def tg_forward_geometry(u, tg, x):
    """Fills the A-side (tensor map, M-space, taps) of a TapGemm for unit ``u`` reading activation ``x``
    over the OUTPUT position space.  Returns the o_stride triple (in output rows)."""
    B, T, F, ld = x.B, x.T, x.F, x.ld
    k, s, p = u.k, u.s, u.p
    tg.K = u.Cin
    tg.ntaps = k
    tg.ntaps_w = k
    for i in range(k):
        tg.b_tap[i] = i
    if u.axis == "flat":
        _set_a(tg, x, [B * T * F], [ld])
        _set_m(tg, (B * T * F, 1, 1), (128, 1, 1), (1, 0, 0))
        return (1, 0, 0)
    if u.axis == "T" and s == 1:
        _set_a(tg, x, [F, T, B], [ld, F * ld, T * F * ld])
        _set_m(tg, (F, T, B), _time_box(F, T, B), (1, 2, 3))
        for i in range(k):
            tg.a_off[i][1] = i - p
        return (1, F, T * F)
    if u.axis == "T":
        assert T % s == 0 and u.y.T == T // s, "time-strided conv needs T divisible by the stride"
        To = T // s
        _set_a(tg, x, [F, s, To, B], [ld, F * ld, s * F * ld, T * F * ld])
        _set_m(tg, (F, To, B), _time_box(F, To, B), (1, 3, 4))
        for i in range(k):
            q = i - p
            tg.a_off[i][1] = q % s
            tg.a_off[i][2] = (q - q % s) // s
        return (1, F, To * F)
    if u.axis == "F" and s == 1:
        _set_a(tg, x, [F, B * T], [ld, F * ld])
        _set_m(tg, (F, B * T, 1), _pow2_box(F), (1, 2, 0))
        for i in range(k):
            tg.a_off[i][0] = i - p
        return (1, F, 0)
    if u.axis == "F":
        assert F % s == 0 and u.y.F == F // s, "frequency-strided conv needs F divisible by the stride"
        Fo = F // s
        _set_a(tg, x, [s, Fo, B * T], [ld, s * ld, F * ld])
        _set_m(tg, (Fo, B * T, 1), _pow2_box(Fo), (2, 3, 0))
        for i in range(k):
            q = i - p
            tg.a_off[i][0] = q % s
            tg.a_off[i][1] = (q - q % s) // s
        return (1, Fo, 0)
    raise NotImplementedError("conv %s is not a tap-GEMM shape" % u.name)


def tg_output_space(u, y):
    """(dims, strides, m_to_a) of a tensor map over an OUTPUT-shaped activation ``y`` matching the M-space
    that tg_forward_geometry defines for ``u``."""
    B, T, F, ld = y.B, y.T, y.F, y.ld
    if u.axis == "flat":
        return [B * T * F], [ld], (1, 0, 0)
    if u.axis == "T":
        return [F, T, B], [ld, F * ld, T * F * ld], (1, 2, 3)
    return [F, B * T], [ld, F * ld], (1, 2, 0)


class Engine:
    def __init__(self, model):
        L.require_cuda_sm100()
        self.model = model
        self.dtype = model.compute_dtype
        self.code = L.dtype_code(self.dtype)
        self.use_tc = self.dtype == torch.bfloat16
        self.device = next(model.parameters()).device
        if self.device.type != "cuda":
            raise RuntimeError("asf_b200 models must live on a CUDA device (call .cuda()); there is no CPU path")
        self.plan = None
        # plans are cached per (batch, input shapes) and never freed: a captured CUDA graph keeps raw pointers into its
        # plan's buffers, so running another batch size (e.g. the partial last eval batch) must not release them
        self._plans = {}
        self.grad_hook = None        # callable(engine, names_done) used by the data-parallel wrapper
        # callable(engine, lo, hi): elements [lo, hi) of the flat gradient buffer are FINAL (reduced over the ranks when
        # data parallel) on the current stream -- the training step updates those parameters there, under the rest of
        # the backward pass, instead of in one launch after it
        self.update_hook = None
        self.stage_hook = None       # callable(engine, unit_name) at every stage boundary of the backward pass
        self._upd_done = 0
        self.force_simt = False      # debugging: run convs on the CUDA-core kernels even in bf16
        # (process_group, world_size) set by the data-parallel wrapper: train-mode BatchNorm then normalises with the
        # statistics of the GLOBAL batch, like the torch.nn.SyncBatchNorm the reference converts to when NUM_GPUS > 1
        # (/root/reference/audio_slowfast/models/build.py:45-46)
        self.sync_bn = None
        self._flat_grad = None
        self._grad_views = {}
        self._ws = {}
        # Two-stream execution (SlowFast only): the fast pathway depends on the slow one only through the lateral
        # convs, so it runs on a side stream next to the slow pathway; each stream sizes its persistent tap-GEMM
        # grids for its share of the SMs so that the two sets of CTAs are resident side by side.
        self.two_stream = model.num_pathways == 2 and os.environ.get("ASF_TWO_STREAM", "1") != "0"
        split = os.environ.get("ASF_SM_SPLIT", "0,0").split(",")  # 0 = no cap: CTAs of both streams share all SMs
        self.sm_split = (int(split[0]), int(split[1]))
        self._side = None
        self._slow = None
        self._wg = {}  # weight-gradient streams (wg_per_path per pathway, round robin), off the dz -> dgrad chain
        self.wg_per_path = max(1, int(os.environ.get("ASF_WG_PER_PATH", "2")))
        self._wg_rr = [0, 0]
        self.wgrad_streams = os.environ.get("ASF_WGRAD_STREAMS", "1") != "0"
        self.defer_unpack = os.environ.get("ASF_DEFER_UNPACK", "1") != "0"
        self._pending_unpack = []

    # ------------------------------------------------------------------------------------------ plan
    def invalidate_weights(self, folds=True):
        """Marks every derived weight copy (packed bf16 layouts and, with ``folds``, the folded BN affine) of every
        cached plan stale.  Needed after the fp32 masters / BN buffers changed behind autograd's version counters --
        e.g. a CUDA-graph replay of a captured optimiser step, which never runs the Python that bumps them."""
        for plan in self._plans.values():
            for u in plan.stems + plan.units:
                u.w_key = None
                if folds:
                    u.fold_key = None

    def prepare_weights(self, training):
        """Refreshes the stale derived copies of the current plan now, on the current stream (TrainStep calls this
        before capturing: the folded affine of FROZEN BatchNorm layers is constant during training and must not be
        re-derived by 100 tiny launches inside every replay; the weight re-pack must)."""
        if self.plan is not None:
            self._prepare_all(training)

    def _build_plan(self, B, in_shapes):
        m = self.model
        dev, dt = self.device, self.dtype
        npath = m.num_pathways
        plan = type("Plan", (), {})()
        plan.B, plan.in_shapes = B, in_shapes
        plan.units, plan.stems, plan.blocks, plan.fuses = [], [], [], []
        named = dict(m.named_modules())

        # ---- stems: conv (SIMT) -> raw ; BN+ReLU+maxpool -> pooled
        x = []
        stem_out_c = []
        for p in range(npath):
            st = getattr(m.s1, "pathway%d_stem" % p)
            stem_out_c.append(st.conv.out_channels)
        # slow pathway buffers carry the lateral channels in the same allocation (no torch.cat)
        def lateral_c(stage_idx):
            if npath == 1:
                return 0
            f = getattr(m, "s%d_fuse" % stage_idx)
            return f.conv_f2s.out_channels

        for p in range(npath):
            st = getattr(m.s1, "pathway%d_stem" % p)
            _, _, T, F = in_shapes[p]
            conv = st.conv
            To = (T + 2 * conv.padding[0] - conv.kernel_size[0]) // conv.stride[0] + 1
            Fo = (F + 2 * conv.padding[1] - conv.kernel_size[1]) // conv.stride[1] + 1
            xin = Act(B, T, F, 1, torch.float32, dev, name="in%d" % p)
            raw = Act(B, To, Fo, conv.out_channels, dt, dev, name="stem%d.raw" % p)
            Tp, Fp = (To + 2 - 3) // 2 + 1, (Fo + 2 - 3) // 2 + 1
            if p == 0 and npath == 2:
                full = Act(B, Tp, Fp, conv.out_channels + lateral_c(1), dt, dev, name="s1.cat")
                pooled = full.slice(0, conv.out_channels, name="s1.p0")
                pooled.full = full
            else:
                pooled = Act(B, Tp, Fp, conv.out_channels, dt, dev, name="s1.p%d" % p)
                pooled.full = pooled
            u = ConvUnit("s1.pathway%d_stem" % p, conv, st.bn, xin, raw, relu=False)
            u.bn_name = "s1.pathway%d_stem.bn" % p
            u.pooled = pooled
            # arg-max codes of the stem max-pool (written by the forward, read by the backward): bf16 vector path only
            u.pool_argmax = (torch.empty((B, Tp, Fp, conv.out_channels // 8), dtype=torch.int32, device=dev)
                             if dt == torch.bfloat16 and conv.out_channels % 8 == 0 else None)
            u.is_stem = True
            u.path = p
            plan.stems.append(u)
            x.append(pooled)

        def add_unit(name, conv, bn, xa, ya, relu, res=None, bn_name=None):
            u = ConvUnit(name, conv, bn, xa, ya, relu, res)
            u.bn_name = bn_name
            u.is_stem = False
            plan.units.append(u)
            return u

        depths = [len([1 for n in named if n.startswith("s%d.pathway0_res" % s) and n.count(".") == 1]) for s in (2, 3, 4, 5)]
        order = []  # execution order entries: ("fuse", unit) | ("block", dict)
        for si, sname in enumerate(("s2", "s3", "s4", "s5")):
            stage = getattr(m, sname)
            if npath == 2:
                f = getattr(m, "s%d_fuse" % (si + 1))
                xs_full = x[0].full
                fuse_out = xs_full.slice(x[0].C, f.conv_f2s.out_channels, name="s%d_fuse.out" % (si + 1))
                fu = add_unit("s%d_fuse" % (si + 1), f.conv_f2s, f.bn, x[1], fuse_out, relu=True,
                              bn_name="s%d_fuse.bn" % (si + 1))
                fu.path = 1
                order.append(("fuse", fu))
                x[0] = xs_full  # the stage reads slow + lateral channels
            for p in range(npath):
                h = x[p]
                for i in range(depths[si]):
                    blk = getattr(stage, "pathway%d_res%d" % (p, i))
                    pre = "%s.pathway%d_res%d" % (sname, p, i)
                    b2 = blk.branch2
                    sf = b2.b.stride[1]
                    cout = b2.c.out_channels
                    T, F = h.T, h.F
                    Fo = (F + 2 * b2.b.padding[1] - 3) // sf + 1
                    last = i == depths[si] - 1
                    # the last slow block of s2..s4 writes into the next concat buffer
                    if last and p == 0 and npath == 2 and si < 3:
                        full = Act(B, T, Fo, cout + lateral_c(si + 2), dt, dev, name=pre + ".cat")
                        out = full.slice(0, cout, name=pre + ".out")
                        out.full = full
                    else:
                        out = Act(B, T, Fo, cout, dt, dev, name=pre + ".out")
                        out.full = out
                    entry = {"name": pre, "x": h, "out": out, "path": p}
                    if hasattr(blk, "branch1"):
                        sc = Act(B, T, Fo, cout, dt, dev, name=pre + ".sc")
                        entry["b1"] = add_unit(pre + ".branch1", blk.branch1, blk.branch1_bn, h, sc, relu=False,
                                               bn_name=pre + ".branch1_bn")
                        res = sc
                    else:
                        res = h
                    ha = Act(B, T, F, b2.a.out_channels, dt, dev, name=pre + ".a")
                    hb = Act(B, T, Fo, b2.b.out_channels, dt, dev, name=pre + ".b")
                    entry["a"] = add_unit(pre + ".branch2.a", b2.a, b2.a_bn, h, ha, relu=True, bn_name=pre + ".branch2.a_bn")
                    entry["b"] = add_unit(pre + ".branch2.b", b2.b, b2.b_bn, ha, hb, relu=True, bn_name=pre + ".branch2.b_bn")
                    entry["c"] = add_unit(pre + ".branch2.c", b2.c, b2.c_bn, hb, out, relu=True, res=res,
                                          bn_name=pre + ".branch2.c_bn")
                    for k_ in ("b1", "a", "b", "c"):
                        if k_ in entry:
                            entry[k_].path = p
                    order.append(("block", entry))
                    h = out
                x[p] = h
        plan.order = order
        plan.final = x
        # head
        head = m.head
        feat_dim = sum(head.dim_in)
        plan.feat = torch.empty((B, feat_dim), dtype=torch.float32, device=dev)
        plan.feat_drop = torch.empty_like(plan.feat)
        plan.dfeat = torch.empty_like(plan.feat)
        for p in range(npath):
            ps = head.pool_size[p]
            assert (x[p].T, x[p].F) == (ps[0], ps[1]), (
                "head pooling %s needs a %dx%d map, got %dx%d: the clip length does not match "
                "AUDIO_DATA.NUM_FRAMES" % (ps, ps[0], ps[1], x[p].T, x[p].F))
        plan.projs = head.projections()
        plan.logits = [torch.empty((B, pr.out_features), dtype=torch.float32, device=dev) for pr in plan.projs]
        # per-unit BN scratch
        # (one flat allocation each: the per-step zeroing of all statistics is ONE memset, not 220 fill kernels)
        allu = plan.stems + plan.units
        tot = sum(u.Cout for u in allu)
        flat = torch.empty(4 * tot, dtype=torch.float32, device=dev)
        # batch statistics are DOUBLES written (not accumulated) by the producers' ordered two-stage reductions
        # (csrc/stats_reduce.cuh): nothing to zero per pass, nothing order-dependent
        plan.stats_flat = torch.zeros(2 * tot, dtype=torch.float64, device=dev)
        plan.red_flat = torch.zeros(2 * tot, dtype=torch.float64, device=dev)
        plan.red_local = torch.zeros(2 * tot, dtype=torch.float64, device=dev)  # per-rank sums under SyncBN
        o = 0
        for u in allu:
            Cn = u.Cout
            u.scale, u.shift = flat[4 * o:4 * o + Cn], flat[4 * o + Cn:4 * o + 2 * Cn]
            u.mean, u.invstd = flat[4 * o + 2 * Cn:4 * o + 3 * Cn], flat[4 * o + 3 * Cn:4 * o + 4 * Cn]
            u.stats = plan.stats_flat[2 * o:2 * o + 2 * Cn]
            u.red = plan.red_flat[2 * o:2 * o + 2 * Cn]
            u.red_local = plan.red_local[2 * o:2 * o + 2 * Cn]
            u.stats_ws = u.stats_ticket = None  # allocated when the unit first runs with batch statistics
            o += Cn
        return plan

    def _sync_sum(self, u, which, t):
        """SyncBatchNorm: ``t`` (doubles) <- sum over the replicas.  Over NVLink peer memory (parallel.PeerExchange, one
        single-CTA kernel, ~5 us) when the symmetric-memory rendezvous works on this system, otherwise the process
        group's all-reduce (~20 us of latency per call; there are 220 of them per from-scratch step).
        ASF_PEER_SYNCBN=0 forces the all-reduce."""
        group, world = self.sync_bn
        plan = self.plan
        px = getattr(plan, "peer", None)
        if px is None and not getattr(plan, "peer_failed", False):
            if os.environ.get("ASF_PEER_SYNCBN", "1") == "0" or not self._plain_units_keys_ok():
                plan.peer_failed = True
            else:
                try:
                    from .parallel import PeerExchange
                    layout = [((v.name, w), 2 * v.Cout) for v in plan.stems + plan.units for w in (0, 1)]
                    px = plan.peer = PeerExchange(group, self.device, layout)
                except Exception as e:  # no peer mapping on this system: keep the library collective
                    plan.peer_failed = True
                    if dist.get_rank(group) == 0:
                        print("asf_b200: NVLink peer exchange unavailable (%s: %s); SyncBN statistics use all_reduce"
                              % (type(e).__name__, str(e)[:120]))
        if px is not None:
            px.allreduce((u.name, which), t)
        else:
            dist.all_reduce(t, group=group)

    def _plain_units_keys_ok(self):
        return not torch.cuda.is_current_stream_capturing()  # the rendezvous is a collective setup: never inside a capture

    def _stat_bufs(self, u):
        """Workspace (per-CTA partial sums) and self-resetting tickets of a unit's statistics reductions; the forward
        statistics and the backward reduction of one unit are stream-ordered, so they share both."""
        if getattr(u, "stats_ws", None) is None:
            u.stats_ws = torch.empty(L.stats_ws_floats(u.Cout), dtype=torch.float32, device=self.device)
            u.stats_ticket = torch.zeros(L.STATS_TICKETS, dtype=torch.int32, device=self.device)
        return u.stats_ws, u.stats_ticket

    # ------------------------------------------------------------------------------------------ weights
    def _prepare_unit(self, u, training):
        """Refreshes (when stale) the folded BN affine and the packed weight copies of a unit."""
        bn, conv = u.bn, u.conv
        u.batch_mode = bool(training and bn.training)
        key = (conv.weight._version, conv.weight.data_ptr(), bn.weight._version, bn.bias._version,
               bn.running_mean._version, bn.running_var._version, u.batch_mode, self.force_simt)
        if key == u.w_key:
            return
        st = L.stream_ptr()
        w32 = conv.weight.detach()
        if w32.dtype != torch.float32 or not w32.is_contiguous():
            raise RuntimeError("asf_b200 keeps fp32 contiguous master weights (got %s)" % w32.dtype)
        wcode = self.code if not u.is_stem else L.F32
        wdt = self.dtype if not u.is_stem else torch.float32
        if u.w_pack is None:
            u.w_pack = torch.empty((u.Cout, u.taps, u.Cin), dtype=wdt, device=self.device)
            if not u.is_stem:
                u.w_pack_t = torch.empty((u.Cin, u.taps, u.Cout), dtype=wdt, device=self.device)
        rs = None
        if not u.batch_mode:
            L.call("asf_bn_fold", L.ptr(bn.weight), L.ptr(bn.bias), L.ptr(bn.running_mean), L.ptr(bn.running_var),
                   C.c_float(bn.eps), u.Cout, L.ptr(u.scale), L.ptr(u.shift), st)
            if not u.is_stem:
                rs = u.scale  # fold the BN scale into the packed weights; the epilogue only adds the shift
        L.call("asf_pack_weight", L.ptr(w32), L.ptr(u.w_pack), wcode, u.Cout, u.Cin, u.taps, L.ptr(rs), st)
        if not u.is_stem:
            L.call("asf_pack_weight_t", L.ptr(w32), L.ptr(u.w_pack_t), wcode, u.Cout, u.Cin, u.taps, L.ptr(rs), st)
        u.w_key = key

    def _prepare_all(self, training):
        """Refreshes every stale folded-BN affine and packed weight copy of the plan: one bn_fold launch per BN
        whose parameters changed (frozen / eval mode only) and ONE multi-tensor pack launch for all stale convs."""
        plan = self.plan
        st = L.stream_ptr()
        stale = []
        for u in plan.stems + plan.units:
            bn, conv = u.bn, u.conv
            u.batch_mode = bool(training and bn.training)
            fold_key = None
            if not u.batch_mode:
                fold_key = (bn.weight._version, bn.bias._version, bn.running_mean._version, bn.running_var._version,
                            bn.weight.data_ptr())
                if fold_key != getattr(u, "fold_key", None):
                    L.call("asf_bn_fold", L.ptr(bn.weight), L.ptr(bn.bias), L.ptr(bn.running_mean),
                           L.ptr(bn.running_var), C.c_float(bn.eps), u.Cout, L.ptr(u.scale), L.ptr(u.shift), st)
                    u.fold_key = fold_key
            key = (conv.weight._version, conv.weight.data_ptr(), fold_key, u.batch_mode, self.force_simt)
            if key != u.w_key:
                w32 = conv.weight
                if w32.dtype != torch.float32 or not w32.is_contiguous():
                    raise RuntimeError("asf_b200 keeps fp32 contiguous master weights (got %s)" % w32.dtype)
                if u.w_pack is None:
                    wdt = self.dtype if not u.is_stem else torch.float32
                    u.w_pack = torch.empty((u.Cout, u.taps, u.Cin), dtype=wdt, device=self.device)
                    if not u.is_stem:
                        u.w_pack_t = torch.empty((u.Cin, u.taps, u.Cout), dtype=wdt, device=self.device)
                stale.append(u)
                u.w_key = key
        if not stale:
            return
        # Two launches: everything the first millisecond of the forward needs (stems, laterals, the whole fast pathway,
        # slow s2 / s3: 6 % of the parameters) and the slow pathway's s4 / s5 (94 %).  When the refresh runs on a side
        # stream (refresh_weights) the forward joins the first before the stems and the second only before the first
        # slow s4 block, so the big re-layout hides beside the early layers.
        late = [u for u in stale if u.name.startswith(("s4.pathway0", "s5.pathway0"))]
        early = [u for u in stale if not u.name.startswith(("s4.pathway0", "s5.pathway0"))]
        if early:
            self._pack_launch(early, st)
        if self._pack_record:
            self._pack_join = torch.cuda.Event()
            self._pack_join.record(torch.cuda.current_stream(self.device))
        if late:
            self._pack_launch(late, st)

    def _pack_launch(self, stale, st):
        plan = self.plan
        tkey = tuple((u.name, u.conv.weight.data_ptr(), u.batch_mode) for u in stale)
        cache = getattr(plan, "_pack_tables", None)
        if cache is None:
            cache = plan._pack_tables = {}
        ent = cache.get(tkey)
        if ent is None:
            arr = (L.PackDesc * len(stale))()
            chunk = 0
            for i, u in enumerate(stale):
                d = arr[i]
                d.w, d.packed = u.conv.weight.data_ptr(), u.w_pack.data_ptr()
                d.packed_t = u.w_pack_t.data_ptr() if u.w_pack_t is not None else None
                fold = (not u.batch_mode) and (not u.is_stem)  # BN scale folded into the packed weights
                d.rscale = u.scale.data_ptr() if fold else None
                d.O, d.I, d.taps, d.dtype = u.Cout, u.Cin, u.taps, L.dtype_code(u.w_pack.dtype)
                d.chunk0 = chunk
                chunk += -(-(u.Cout * u.Cin * u.taps) // L.PACK_CHUNK)
            host = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8)
            ent = (host.to(self.device), len(stale), chunk)
            if len(cache) > 8:
                cache.clear()
            cache[tkey] = ent
        L.call("asf_pack_weights_multi", L.ptr(ent[0]), ent[1], ent[2], st)

    _pack_record = False
    _pack_join = None
    _pack_join_late = None

    def refresh_weights(self, training=True, stream=None, after=None):
        """Re-packs the stale bf16 weight copies now, on ``stream`` (forked from and joined back to the current stream),
        instead of at the start of the next forward: the training step calls this before the log-mel extraction, so
        the HBM-bound re-layout of the 26.9 M parameters that the optimiser just changed runs beside the tensor-core
        log-mel kernel instead of in front of the stems.  No-op until the first forward has built the plan."""
        if self.plan is None:
            return
        cur = torch.cuda.current_stream(self.device)
        if stream is None:
            if getattr(self, "_pack_stream", None) is None:
                self._pack_stream = torch.cuda.Stream(device=self.device)
            stream = self._pack_stream
        # ``after``: an event recorded on the current stream BEFORE the work this refresh is meant to run beside (the
        # caller launches that work first, so its CTAs are resident when the short pack blocks arrive)
        if after is not None:
            stream.wait_event(after)
        else:
            stream.wait_stream(cur)
        self._pack_join = None
        with torch.cuda.stream(stream):
            self._pack_record = True
            try:
                self._prepare_all(training)
            finally:
                self._pack_record = False
            done = torch.cuda.Event()
            done.record(stream)
        if self._pack_join is None:
            self._pack_join = done
        self._pack_join_late = done

    def join_refresh(self, late=False):
        """Makes the current stream wait for the side-stream weight refresh: the early group (before the first launch of
        the forward) or, ``late``, everything (before the first slow s4 block and at the end of the forward)."""
        if late:
            ev, self._pack_join_late = self._pack_join_late, None
        else:
            ev, self._pack_join = self._pack_join, None
        if ev is not None:
            torch.cuda.current_stream(self.device).wait_event(ev)

    # ------------------------------------------------------------------------------------------ conv launches
    # ------------------------------------------------------------------------------------------ streams
    def _ts(self):
        return self.two_stream and self.plan is not None and len(self.plan.stems) == 2

    def _side_stream(self):
        if self._side is None:
            # ASF_SIDE_PRIO: priority of the fast pathway's stream alone (its backward chain is what the step ends on)
            self._side = torch.cuda.Stream(device=self.device, priority=int(os.environ.get(
                "ASF_SIDE_PRIO", os.environ.get("ASF_PATH_PRIO", "0"))))
        return self._side

    def _slow_stream(self):
        """The slow pathway's stream.  Both pathway streams outrank the weight-gradient streams, so that when an SM
        frees up the next dgrad / forward CTA (which the following layer waits for) is placed before a wgrad CTA."""
        if self._slow is None:
            self._slow = torch.cuda.Stream(device=self.device, priority=int(os.environ.get("ASF_PATH_PRIO", "0")))
        return self._slow

    def _wg_stream(self, path, advance=False):
        """Weight-gradient stream of a pathway (round robin over wg_per_path streams when ``advance``)."""
        idx = self._wg_rr[path]
        if advance:
            self._wg_rr[path] = (idx + 1) % self.wg_per_path
        key = (path, idx)
        if key not in self._wg:
            self._wg[key] = torch.cuda.Stream(device=self.device, priority=int(os.environ.get("ASF_WG_PRIO", "0")))
        return self._wg[key]

    def _on(self, path):
        """Context that issues work of pathway ``path`` on its stream (fast pathway -> side stream)."""
        if self._ts():
            return torch.cuda.stream(self._side_stream() if path == 1 else self._slow_stream())
        return contextlib.nullcontext()

    def _cap(self, u):
        return self.sm_split[getattr(u, "path", 0)] if self._ts() else 0

    def _tc_ok(self, u):
        return self.use_tc and not self.force_simt and u.axis is not None and u.Cin % 8 == 0

    def _conv_fwd(self, u, out, epi_kwargs):
        st = L.stream_ptr()
        if self._tc_ok(u):
            tg = L.TapGemm()
            o_stride = tg_forward_geometry(u, tg, u.x)
            tg.bw, tg.N = u.w_pack.data_ptr(), u.Cout
            tg.max_ctas = self._cap(u)
            tg.out, tg.out_dtype, tg.ld_out = out.ptr, out.code, out.ld
            for j in range(3):
                tg.o_stride[j] = o_stride[j]
            tg.epi = _epilogue(**epi_kwargs)
            L.call("asf_tapgemm", C.byref(tg), 0, st)
        else:
            g = u.geom()
            g.ld_out = out.ld
            e = _epilogue(**epi_kwargs)
            L.call("asf_conv_fwd_simt", C.c_void_p(u.x.ptr), u.x.code, L.ptr(u.w_pack), L.dtype_code(u.w_pack.dtype),
                   C.c_void_p(out.ptr), out.code, C.byref(g), C.byref(e), st)

    def _conv_dgrad(self, u, dy, dx, accumulate, res=None, mask=None):
        """dx (+)= conv^T(dy) [+ res] then masked by (mask > 0).  Handles strided convs as one launch per
        input phase; the caller guarantees dx was zeroed if some phase has no tap and accumulate is False."""
        st = L.stream_ptr()
        if not self._tc_ok(u):
            g = u.geom()
            g.ld_in, g.ld_out = dx.ld, dy.ld
            e = _epilogue(res=res, mask=mask, accumulate=accumulate)
            L.call("asf_conv_dgrad_simt", C.c_void_p(dy.ptr), dy.code, L.ptr(u.w_pack), L.dtype_code(u.w_pack.dtype),
                   C.c_void_p(dx.ptr), dx.code, C.byref(g), C.byref(e), st)
            return
        B, k, s, p = dy.B, u.k, u.s, u.p
        # test hook (tests/test_gpu_model.py::test_forced_parity_detects_a_broken_dgrad): shift the first tap of every
        # 3-tap dgrad by one position, i.e. a deliberately wrong backward pass.  Never set outside that test.
        broken = k == 3 and os.environ.get("ASF_TEST_BREAK_DGRAD") == "1"
        for rho in range(s):
            taps = [(j, (rho + p - j) // s) for j in range(k) if (rho + p - j) % s == 0]
            if broken:
                taps = [(j, off + (1 if i == 0 else 0)) for i, (j, off) in enumerate(taps)]
            if not taps:
                assert accumulate or getattr(dx, "zeroed", False), "uncovered phase without zero fill: " + u.name
                continue
            tg = L.TapGemm()
            tg.K, tg.N = u.Cout, u.Cin
            tg.ntaps, tg.ntaps_w = len(taps), k
            T, F, ld = dy.T, dy.F, dy.ld
            row_off = 0
            if u.axis == "flat":
                _set_a(tg, dy, [B * T * F], [ld])
                _set_m(tg, (B * T * F, 1, 1), (128, 1, 1), (1, 0, 0))
                o_stride = (1, 0, 0)
            elif u.axis == "T":
                _set_a(tg, dy, [F, T, B], [ld, F * ld, T * F * ld])
                Ti = dx.T
                _set_m(tg, (F, Ti // s, B), _time_box(F, Ti // s, B), (1, 2, 3))
                for i, (j, off) in enumerate(taps):
                    tg.a_off[i][1] = off
                o_stride = (1, s * F, Ti * F)
                row_off = rho * F
            else:
                _set_a(tg, dy, [F, B * T], [ld, F * ld])
                Fi = dx.F
                _set_m(tg, (Fi // s, B * T, 1), _pow2_box(Fi // s), (1, 2, 0))
                for i, (j, off) in enumerate(taps):
                    tg.a_off[i][0] = off
                o_stride = (s, Fi, 0)
                row_off = rho
            for i, (j, off) in enumerate(taps):
                tg.b_tap[i] = j
            tg.bw = u.w_pack_t.data_ptr()
            tg.max_ctas = self._cap(u) or int(os.environ.get("ASF_DGRAD_CTAS", "0"))
            tg.out, tg.out_dtype, tg.ld_out = dx.row_ptr(row_off), dx.code, dx.ld
            for j in range(3):
                tg.o_stride[j] = o_stride[j]
            tg.epi = _epilogue(res=res, mask=mask, accumulate=accumulate, row_off=row_off)
            L.call("asf_tapgemm", C.byref(tg), 0, st)

    def _conv_wgrad(self, u, dy, rscale, gview, accumulate):
        st = L.stream_ptr()
        if u.is_stem:
            g = u.geom()
            g.ld_out = dy.ld
            if rscale is not None:
                raise RuntimeError("stem BN is never folded in training")
            n = u.Cout * u.taps
            rows = getattr(u, "wgf_rows", None)
            if rows is None:
                rows = 0
                if os.environ.get("ASF_STEM_FUSED", "1") != "0" and gview.is_contiguous() and gview.dtype == torch.float32:
                    rows = L.lib().asf_stem_wgrad_fused_rows(C.byref(g), dy.code)
                u.wgf_rows = rows
                if rows > 0:
                    u.wgf_ws = torch.empty(rows * n, dtype=torch.float32, device=self.device)
                    u.wgf_ticket = torch.zeros(1, dtype=torch.int32, device=self.device)
            if rows > 0:
                # input lines staged in shared memory, cross-block sum folded into the kernel: the last kernel of the
                # backward pass is not followed by a reduction launch
                L.call("asf_stem_wgrad_fused", C.c_void_p(u.x.ptr), C.c_void_p(dy.ptr), dy.code, L.ptr(u.wgf_ws), L.ptr(gview),
                       int(bool(accumulate)), L.ptr(u.wgf_ticket), C.byref(g), st)
                return
            ks = getattr(u, "wg_ksplit", None)
            if ks is None:
                ks = L.lib().asf_stem_wgrad_ksplit(C.byref(g), dy.code)
                if ks < 1:
                    raise RuntimeError("asf_stem_wgrad_ksplit failed (code %d): %s" % (ks, L.lib().asf_last_error().decode()))
                u.wg_ksplit = ks
            ws = getattr(u, "wg_ws", None)
            if ws is None or ws.numel() < ks * n:
                ws = u.wg_ws = torch.empty(ks * n, dtype=torch.float32, device=self.device)
            # row-range partial sums [ks][Cout][taps]; their fixed-order sum is a split-K reduction with O = taps = 1
            L.call("asf_stem_wgrad", C.c_void_p(u.x.ptr), C.c_void_p(dy.ptr), dy.code, L.ptr(ws), C.byref(g), st)
            if self._defer_unpack():
                self._pending_unpack.append((u.name, ws, gview, None, bool(accumulate), ks, 1, n, 1))
            else:
                L.call("asf_unpack_wgrad", L.ptr(ws), L.ptr(gview), 1, n, 1, None, int(accumulate), ks, st)
            return
        if not self._tc_ok(u):
            g = u.geom()
            g.ld_out = dy.ld
            ks = L.lib().asf_conv_wgrad_simt_ksplit(C.byref(g))
            if ks < 1:
                raise RuntimeError("asf_conv_wgrad_simt_ksplit failed (code %d): %s" % (ks, L.lib().asf_last_error().decode()))
            ws = self._wgrad_ws(ks * u.Cout * u.taps * u.Cin)
            L.call("asf_conv_wgrad_simt", C.c_void_p(u.x.ptr), u.x.code, C.c_void_p(dy.ptr), dy.code, L.ptr(gview),
                   C.byref(g), L.ptr(rscale), int(accumulate), L.ptr(ws), st)
            return
        tg = L.TapGemm()
        tg_forward_geometry(u, tg, u.x)
        tg.N = u.Cout
        tg.max_ctas = self._cap(u)
        dims, strides, m2a = tg_output_space(u, dy)
        tg.y, tg.y_nd = dy.ptr, len(dims)
        for i, (d, s_) in enumerate(zip(dims, strides)):
            tg.y_dim[i], tg.y_stride[i] = d, s_
        for j in range(3):
            tg.y_m_to_a[j] = m2a[j]
        ks = getattr(u, "wg_ksplit", None) if getattr(u, "wg_ksplit_cap", None) == tg.max_ctas else None
        if ks is None:  # the launcher's split-K choice depends on the geometry only: ask once per unit
            ks = L.lib().asf_tapgemm_wgrad_ksplit(C.byref(tg))
            if ks < 1:
                raise RuntimeError("asf_tapgemm_wgrad_ksplit failed (code %d): %s" % (ks, L.lib().asf_last_error().decode()))
            u.wg_ksplit, u.wg_ksplit_cap = ks, tg.max_ctas
        if self._defer_unpack():
            # own workspace per layer; the split-K reduction of a whole backward segment is ONE launch (_flush_unpacks)
            ws = getattr(u, "wg_ws", None)
            if ws is None or ws.numel() < ks * u.Cout * u.taps * u.Cin:
                ws = u.wg_ws = torch.empty(ks * u.Cout * u.taps * u.Cin, dtype=torch.float32, device=self.device)
            tg.dwp, tg.ksplit = ws.data_ptr(), ks
            L.call("asf_tapgemm", C.byref(tg), 1, st)
            self._pending_unpack.append((u.name, ws, gview, rscale, bool(accumulate), ks, u.Cout, u.Cin, u.taps))
            return
        ws = self._wgrad_ws(ks * u.Cout * u.taps * u.Cin)
        tg.dwp, tg.ksplit = ws.data_ptr(), ks
        L.call("asf_tapgemm", C.byref(tg), 1, st)
        L.call("asf_unpack_wgrad", L.ptr(ws), L.ptr(gview), u.Cout, u.Cin, u.taps, L.ptr(rscale), int(accumulate), ks, st)

    def _defer_unpack(self):
        return self._ts() and self.wgrad_streams and self.defer_unpack

    def _flush_unpacks(self):
        """Split-K reduction + OIHW re-layout of every weight gradient computed since the last flush, in one launch
        on the current stream (which the caller has ordered after the weight-gradient streams)."""
        pend = self._pending_unpack
        if not pend:
            return
        self._pending_unpack = []
        key = tuple((name, ws.data_ptr(), g.data_ptr(), rs.data_ptr() if rs is not None else 0, acc, ks)
                    for name, ws, g, rs, acc, ks, O, I, taps in pend)
        cache = self.__dict__.setdefault("_unpack_tables", {})
        ent = cache.get(key)
        if ent is None:
            arr = (L.UnpackDesc * len(pend))()
            chunk = 0
            for i, (name, ws, g, rs, acc, ks, O, I, taps) in enumerate(pend):
                d = arr[i]
                d.dwp, d.dw = ws.data_ptr(), g.data_ptr()
                d.rscale = rs.data_ptr() if rs is not None else None
                d.O, d.I, d.taps, d.ksplit, d.accumulate = O, I, taps, ks, int(acc)
                d.chunk0 = chunk
                chunk += -(-(O * I * taps) // (1024 if ks < 32 else 256))  # ASF_UNPACK_CHUNK / ASF_UNPACK_CHUNK_DEEP
            host = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8)
            if len(cache) > 64:
                cache.clear()
            ent = cache[key] = (host.to(self.device), len(pend), chunk)
        L.call("asf_unpack_wgrad_multi", L.ptr(ent[0]), ent[1], ent[2], L.stream_ptr())

    # ------------------------------------------------------------------------------------------ grads
    def _ensure_grad_buffers(self):
        if self._flat_grad is not None:
            return
        params = [(n, p) for n, p in self.model.named_parameters()]
        # reverse registration order ~ order in which backward produces them: head, s5, s4, ... stems
        params = params[::-1]
        # every view starts on a 16-byte boundary (the 97-element verb bias would otherwise knock every later tensor off
        # it): the split-K reduction and the optimiser then use 16-byte accesses on all of them
        total = sum((p.numel() + 3) // 4 * 4 for _, p in params)
        self._flat_grad = torch.zeros(total, dtype=torch.float32, device=self.device)
        off = 0
        self._grad_offsets = {}
        for n, p in params:
            self._grad_views[n] = self._flat_grad[off:off + p.numel()].view_as(p)
            self._grad_offsets[n] = (off, p.numel())
            off += (p.numel() + 3) // 4 * 4
        self._param_by_name = dict(params)
        self._name_of = {id(p): n for n, p in params}

    def grad_prefix_end(self, names):
        """End offset (elements) of the flat gradient buffer covered by the parameters of the given finished
        units ("head", "s5.pathway1_res2", "s4_fuse", ..., "stems").  The buffer is laid out in backward order, so
        everything before the returned offset is final once those units are."""
        cache = self.__dict__.setdefault("_prefix_end", {})
        end = 0
        for nm in names:
            pre = "s1." if nm == "stems" else nm + "."
            key = (pre, id(self._flat_grad))
            e = cache.get(key)
            if e is None:
                e = max([o + n for k, (o, n) in self._grad_offsets.items() if k.startswith(pre)] or [0])
                cache[key] = e
            end = max(end, e)
        return end

    def _wgrad_ws(self, nfloats):
        """Split-K workspace of the tcgen05 wgrad ([ksplit][Cout][taps][Cin] fp32 partial tiles, written then summed
        by asf_unpack_wgrad on the same stream): one grow-only buffer per stream, shared by the layers issued on it."""
        # one workspace per issuing stream, keyed by the stream's handle (an index into the list of streams would shift
        # while the weight-gradient streams are being created during the first backward pass and hand a buffer that
        # another stream's kernels are still using to a new owner)
        key = torch.cuda.current_stream(self.device).cuda_stream
        ws = self._ws.get(key)
        if ws is None or ws.numel() < nfloats:
            ws = self._ws[key] = torch.empty(max(nfloats, 1 << 22), dtype=torch.float32, device=self.device)
        return ws

    def _grad_target(self, p):
        """Returns (grad_view, accumulate) for a parameter following autograd semantics: a missing .grad is
        (re)attached as a view of the flat buffer and overwritten; an existing one is accumulated into."""
        if not p.requires_grad:
            return None, False
        n = self._name_of[id(p)]
        view = self._grad_views[n]
        if p.grad is None:
            p.grad = view
            return view, False
        if p.grad.data_ptr() != view.data_ptr():
            # user-owned grad tensor: accumulate there
            return p.grad, True
        return view, True

    # ------------------------------------------------------------------------------------------ forward
    def run(self, xs):
        m = self.model
        B = xs[0].shape[0]
        shapes = tuple(tuple(t.shape) for t in xs)
        for t in xs:
            if t.dim() != 4 or t.shape[1] != 1:
                raise ValueError("expected inputs of shape (B, 1, T, F), got %s" % (tuple(t.shape),))
            if t.device != self.device:
                raise RuntimeError("input on %s but the model is on %s" % (t.device, self.device))
        if self.plan is None or self.plan.B != B or self.plan.in_shapes != shapes:
            plan = self._plans.get((B, shapes))
            if plan is None:
                plan = self._plans[(B, shapes)] = self._build_plan(B, shapes)
            self.plan = plan
        training = m.training
        needs_grad = torch.is_grad_enabled() and training and any(p.requires_grad for p in m.parameters())
        if needs_grad:
            self._ensure_grad_buffers()
            params = [p for p in m.parameters() if p.requires_grad]
            outs = _NetFn.apply(self, len(xs), *xs, *params)
        else:
            outs = self._forward(xs, training, None)
        outs = tuple(outs)
        return outs if len(outs) > 1 else outs[0]

    def _bn_forward(self, u, raw, out, res, relu, training, st):
        """Training-mode BN on a raw conv output whose per-channel stats are already in u.stats."""
        bn = u.bn
        rows = raw.rows
        mom = bn.momentum if bn.momentum is not None else 0.1
        track = bn.track_running_stats
        count = float(rows)
        if self.sync_bn is not None:
            # SyncBatchNorm forward: sum(x), sum(x^2) and the element count are summed over the replicas; mean, biased
            # variance and the unbiased running-variance update then follow from the global count
            group, world = self.sync_bn
            self._sync_sum(u, 0, u.stats)
            count *= world
        # statistics -> scale / shift / saved mean, invstd / running stats, and the normalise(+res)(+ReLU) pass: one launch
        L.call("asf_bn_train_fwd", C.c_void_p(raw.ptr), raw.ld, C.c_void_p(res.ptr) if res is not None else None,
               res.ld if res is not None else 0, C.c_void_p(out.ptr) if out is not None else None,
               out.ld if out is not None else 0, raw.code, rows, u.Cout, L.ptr(u.stats), C.c_double(count),
               L.ptr(bn.weight), L.ptr(bn.bias), C.c_float(bn.eps), C.c_float(mom),
               L.ptr(bn.running_mean) if track else None, L.ptr(bn.running_var) if track else None,
               L.ptr(bn.num_batches_tracked) if track else None, L.ptr(u.scale), L.ptr(u.shift), L.ptr(u.mean),
               L.ptr(u.invstd), int(relu), st)

    def _unit_forward(self, u, training, prepared=False):
        st = L.stream_ptr()
        if not prepared:
            self._prepare_unit(u, training)
        if u.batch_mode:
            if u.raw is None:
                u.raw = Act(u.y.B, u.y.T, u.y.F, u.Cout, self.dtype, self.device, name=u.name + ".raw")
            ws, tk = self._stat_bufs(u)
            if self._tc_ok(u):
                self._conv_fwd(u, u.raw, dict(stats=u.stats, stats_stride=u.Cout, stats_ws=ws, stats_ticket=tk))
            else:
                self._conv_fwd(u, u.raw, dict())
                L.call("asf_channel_stats", C.c_void_p(u.raw.ptr), u.raw.code, u.raw.rows, u.Cout, u.raw.ld,
                       L.ptr(u.stats), L.ptr(ws), L.ptr(tk), st)
            self._bn_forward(u, u.raw, u.y, u.res, u.relu, training, st)
        else:
            self._conv_fwd(u, u.y, dict(shift=u.shift, res=u.res, relu=u.relu))

    def _forward(self, xs, training, ctx):
        plan, m = self.plan, self.model
        st = L.stream_ptr()
        self.join_refresh()
        self._prepare_all(training)
        ts = self._ts()
        cur = torch.cuda.current_stream(self.device)
        main = self._slow_stream() if ts else cur   # the slow pathway's stream
        side = self._side_stream() if ts else None  # the fast pathway's stream
        if ts:
            main.wait_stream(cur)  # fork: inputs, zeroed statistics and refreshed weights are ready on `cur`
            side.wait_stream(cur)
        # stems
        for p, u in enumerate(plan.stems):
            with self._on(p):
                st = L.stream_ptr()
                xin = xs[p]
                if xin.dtype != torch.float32:
                    xin = xin.float()
                u.x.buf.copy_(xin.reshape(u.x.buf.shape))  # (B,1,T,F) fp32 is already channels-last for C = 1
                g = u.geom()
                e = _epilogue()
                L.call("asf_conv_fwd_simt", C.c_void_p(u.x.ptr), L.F32, L.ptr(u.w_pack), L.F32, C.c_void_p(u.y.ptr),
                       u.y.code, C.byref(g), C.byref(e), st)
                if u.batch_mode:
                    ws, tk = self._stat_bufs(u)
                    L.call("asf_channel_stats", C.c_void_p(u.y.ptr), u.y.code, u.y.rows, u.Cout, u.y.ld,
                           L.ptr(u.stats), L.ptr(ws), L.ptr(tk), st)
                    self._bn_forward(u, u.y, None, None, False, training, st)
                L.call("asf_bn_relu_maxpool_fwd", C.c_void_p(u.y.ptr), C.c_void_p(u.pooled.ptr), u.pooled.ld,
                       u.y.code, u.y.B, u.y.T, u.y.F, u.Cout, L.ptr(u.scale), L.ptr(u.shift), L.ptr(u.pool_argmax), st)
        # stages.  Two-stream mode: laterals and fast blocks run ahead on the side stream; the slow blocks of a stage
        # wait for that stage's lateral (which writes the lateral channels of the buffer they read).
        ev_fuse = None
        for kind, item in plan.order:
            if kind == "fuse":
                with self._on(1):
                    self._unit_forward(item, training, True)
                    if ts:
                        ev_fuse = torch.cuda.Event()
                        ev_fuse.record(side)
            else:
                if ts and item["path"] == 0 and ev_fuse is not None:
                    main.wait_event(ev_fuse)
                    ev_fuse = None
                with self._on(item["path"]):
                    if self._pack_join_late is not None and item["name"].startswith(("s4.pathway0", "s5.pathway0")):
                        self.join_refresh(late=True)
                    if "b1" in item:
                        self._unit_forward(item["b1"], training, True)
                    self._unit_forward(item["a"], training, True)
                    self._unit_forward(item["b"], training, True)
                    self._unit_forward(item["c"], training, True)
        if ts:
            cur.wait_stream(main)  # join
            cur.wait_stream(side)
        if self._pack_join_late is not None:
            self.join_refresh(late=True)
        st = L.stream_ptr()
        # head
        head = m.head
        c_off = 0
        for p, xa in enumerate(plan.final):
            L.call("asf_avgpool_fwd", C.c_void_p(xa.ptr), xa.code, xa.B, xa.T * xa.F, xa.C, L.ptr(plan.feat),
                   plan.feat.shape[1], c_off, st)
            c_off += xa.C
        feat = plan.feat
        plan.drop_mask = None
        if training and head.dropout_rate > 0.0:
            keep = 1.0 - head.dropout_rate
            if getattr(plan, "drop_mask_buf", None) is None:
                plan.drop_mask_buf = torch.empty_like(plan.feat)
            plan.drop_mask = plan.drop_mask_buf
            draws = torch.rand_like(plan.feat)
            L.call("asf_dropout_apply", L.ptr(plan.feat), L.ptr(draws), L.ptr(plan.feat_drop),
                   L.ptr(plan.drop_mask), plan.feat.numel(), C.c_float(keep), st)
            feat = plan.feat_drop
        plan.feat_used = feat
        # every task head in ONE launch, straight into fresh output tensors (csrc/head.cu)
        projs = plan.projs
        outs = [torch.empty((plan.B, pr.out_features), dtype=torch.float32, device=self.device) for pr in projs]
        L.call("asf_head_fwd", L.ptr(feat), feat.shape[1], plan.B, projs[0].in_features, len(projs),
               L.ptr_array([pr.weight for pr in projs]), L.ptr_array([pr.bias for pr in projs]),
               L.int_array([pr.out_features for pr in projs]), L.ptr_array(outs),
               L.i64_array([pr.out_features for pr in projs]), st)
        if not training:
            for pr, lg in zip(projs, outs):
                L.call("asf_softmax_rows", L.ptr(lg), plan.B, pr.out_features, pr.out_features, st)
        return outs

    # ------------------------------------------------------------------------------------------ backward
    def _unit_backward(self, u, dz, need_dx, dx_accumulate, dx_res=None, dx_mask=None):
        """dz: Act holding dL/d(BN output [+res]) of unit u (already ReLU-masked).  Produces the parameter
        gradients of u and, when need_dx, u.x.grad (+)= dgrad."""
        st = L.stream_ptr()
        bn, conv = u.bn, u.conv
        if u.batch_mode:
            rows = dz.rows
            ws, tk = self._stat_bufs(u)
            L.call("asf_bn_bwd_reduce", C.c_void_p(dz.ptr), dz.ld, C.c_void_p(u.raw.ptr), u.raw.ld, dz.code, rows,
                   u.Cout, L.ptr(u.mean), L.ptr(u.invstd), L.ptr(u.red), L.ptr(ws), L.ptr(tk), st)
            gw, acc_w = self._grad_target(bn.weight)
            gb, acc_b = self._grad_target(bn.bias)
            if (gw is not None) and (gb is not None) and acc_w != acc_b:
                raise RuntimeError("%s: BN weight and bias gradients must both exist or both be unset" % u.name)
            dconv = u.raw.grad
            count = float(rows)
            pg_w, pg_b = gw, gb
            if self.sync_bn is not None:
                # SyncBatchNorm backward: dx uses sum(dz), sum(dz * xhat) of the GLOBAL batch; dgamma / dbeta stay the
                # per-rank sums (the gradient all-reduce averages them like every other parameter gradient)
                group, world = self.sync_bn
                u.red_local.copy_(u.red)
                self._sync_sum(u, 1, u.red)
                count *= world
                pg_w = pg_b = None
            L.call("asf_bn_bwd_apply", C.c_void_p(dz.ptr), dz.ld, C.c_void_p(u.raw.ptr), u.raw.ld,
                   C.c_void_p(dconv.ptr), dconv.ld, dz.code, rows, u.Cout, L.ptr(bn.weight), L.ptr(u.mean),
                   L.ptr(u.invstd), L.ptr(u.red), C.c_double(count), L.ptr(pg_w), L.ptr(pg_b), int(acc_w or acc_b), st)
            if self.sync_bn is not None:
                for g_, acc_, src in ((gb, acc_b, u.red_local[:u.Cout]), (gw, acc_w, u.red_local[u.Cout:])):
                    if g_ is not None:
                        g_.add_(src) if acc_ else g_.copy_(src)
            rscale = None
        else:
            dconv = dz
            rscale = None if u.is_stem else u.scale
            if bn.weight.requires_grad or bn.bias.requires_grad:
                raise RuntimeError(
                    "%s: BN in eval mode with trainable gamma/beta is not supported by the fused backward "
                    "(use freeze_fn('bn_parameters') together with freeze_fn('bn_statistics'))" % u.name)
        gview, acc = self._grad_target(conv.weight)
        if gview is not None:
            if self._ts() and self.wgrad_streams and not u.is_stem:
                # dW needs dconv and the saved input only; nothing downstream in this pass reads it
                wg = self._wg_stream(u.path, advance=True)
                ev = torch.cuda.Event()
                ev.record(torch.cuda.current_stream(self.device))
                wg.wait_event(ev)
                with torch.cuda.stream(wg):
                    self._conv_wgrad(u, dconv, rscale, gview, acc)
            else:
                self._conv_wgrad(u, dconv, rscale, gview, acc)
        if need_dx:
            self._conv_dgrad(u, dconv, u.x.grad, dx_accumulate, res=dx_res, mask=dx_mask)

    def _backward(self, dlogits):
        plan, m = self.plan, self.model
        st = L.stream_ptr()
        B = plan.B
        self._pending_unpack = []
        # ---- head: dW, db of every task head and d feat (dropout mask applied) in one launch (csrc/head.cu)
        feat = plan.feat_used
        projs = plan.projs
        live = [(pr, dl) for pr, dl in zip(projs, dlogits) if dl is not None]
        if not live:
            raise RuntimeError("asf_b200: backward called without a gradient for any head output")
        dls = []
        for pr, dl in live:
            if dl.dtype != torch.float32 or dl.stride(-1) != 1:
                dl = dl.contiguous().float()
            dls.append(dl)
        gws = [self._grad_target(pr.weight) for pr, _ in live]
        gbs = [self._grad_target(pr.bias) for pr, _ in live]
        keep = 1.0 - m.head.dropout_rate
        L.call("asf_head_bwd", len(live), L.ptr_array(dls), L.i64_array([dl.stride(0) for dl in dls]),
               L.int_array([pr.out_features for pr, _ in live]), L.ptr_array([pr.weight for pr, _ in live]),
               L.ptr(feat), feat.shape[1], L.ptr(plan.drop_mask), C.c_float(1.0 / keep if plan.drop_mask is not None else 1.0),
               B, projs[0].in_features, L.ptr_array([g for g, _ in gws]), L.int_array([a for _, a in gws]),
               L.ptr_array([g for g, _ in gbs]), L.int_array([a for _, a in gbs]), L.ptr(plan.dfeat),
               plan.dfeat.shape[1], st)
        self._notify(["head"])
        c_off = 0
        for xa in plan.final:
            # gradient w.r.t. the pre-ReLU block output: avg-pool backward masked by (out > 0)
            L.call("asf_avgpool_bwd", L.ptr(plan.dfeat), plan.dfeat.shape[1], c_off, C.c_void_p(xa.ptr),
                   C.c_void_p(xa.grad.ptr), xa.code, xa.B, xa.T * xa.F, xa.C, st)
            c_off += xa.C
        # which activations already hold a (partial) gradient in this pass
        written = set(id(xa.full.grad.buf) for xa in plan.final)

        def first_write(act):
            key = id(act.grad.buf)
            fresh = key not in written
            written.add(key)
            return fresh

        ts = self._ts()
        cur = torch.cuda.current_stream(self.device)
        main = self._slow_stream() if ts else cur
        side = self._side_stream() if ts else None
        if ts:
            main.wait_stream(cur)  # fork: head gradients / pooled gradients of both pathways are on `cur`
            side.wait_stream(cur)
        pending = []  # two-stream mode: units finished since the last point where both streams were ordered
        for kind, item in reversed(plan.order):
            if kind == "fuse":
                u = item
                if ts:
                    # the lateral reads its slice of the slow concat gradient (written by the first slow block of the
                    # stage above, on `main`) and is the last writer of the fast activation's gradient (side stream)
                    ev = torch.cuda.Event()
                    ev.record(main)
                    side.wait_event(ev)
                with self._on(1):
                    # dz of the lateral = its channel slice of the slow concat gradient (already masked by the
                    # consumer block's last dgrad); dx accumulates into the fast activation's gradient.
                    fresh = first_write(u.x)
                    u.x.grad.zeroed = False
                    # the lateral is always the LAST writer of its (fast) input's gradient: it applies the ReLU mask
                    self._unit_backward(u, u.y.grad, True, not fresh, dx_mask=u.x)
                    # on the side stream, after the wait above, everything issued so far on either stream is ordered
                    self._notify_joined(pending + [u.name])
                    pending = []
            else:
                blk = item
                with self._on(blk["path"]):
                    xin, out = blk["x"], blk["out"]
                    dz_out = out.grad
                    c, b_, a = blk["c"], blk["b"], blk["a"]
                    # c: dgrad -> d(h_b), masked by h_b > 0
                    self._unit_backward(c, dz_out, True, False, dx_mask=c.x)
                    self._unit_backward(b_, b_.y.grad, True, False, dx_mask=b_.x)
                    fresh = first_write(xin)
                    xg = xin.grad
                    # is this block input also consumed elsewhere later in the backward (lateral of the fast path)?
                    # the mask is applied by the LAST writer; a stage's first fast block is followed (in backward
                    # order) by the lateral conv that reads the same activation, so the mask must wait for it.
                    later = blk.get("mask_later", False)
                    if "b1" in blk:
                        b1 = blk["b1"]
                        if fresh and b1.s > 1:
                            xg.buf.zero_() if xg.parent is None else xg.view().zero_()
                            xg.zeroed = True
                            fresh = False
                        self._unit_backward(b1, dz_out, True, not fresh)
                        self._unit_backward(a, a.y.grad, True, True, dx_mask=None if later else xin)
                    else:
                        self._unit_backward(a, a.y.grad, True, not fresh, dx_res=dz_out, dx_mask=None if later else xin)
                    if ts:
                        pending.append(blk["name"])
                    else:
                        self._notify([blk["name"]])
        # ---- stems
        for u in plan.stems:
            with self._on(u.path):
                st = L.stream_ptr()
                pooled = u.pooled
                dz = u.y.grad
                L.call("asf_bn_relu_maxpool_bwd", C.c_void_p(u.y.ptr), C.c_void_p(pooled.grad.ptr), pooled.grad.ld,
                       C.c_void_p(dz.ptr), dz.code, u.y.B, u.y.T, u.y.F, u.Cout, L.ptr(u.scale), L.ptr(u.shift),
                       L.ptr(u.pool_argmax), st)
                if u.batch_mode:
                    u.raw = u.y  # the stem keeps its raw conv output in u.y
                self._unit_backward(u, dz, False, False)
        if ts:
            cur.wait_stream(main)  # join
            cur.wait_stream(side)
            for wg in self._wg.values():
                cur.wait_stream(wg)
        self._flush_unpacks()
        self._notify(pending + ["stems"])

    def _notify_joined(self, names):
        """Two-stream mode: gradient-ready point, issued where the current (side) stream is ordered after everything
        both pathway streams have been given so far.  The weight gradients live on the wgrad streams, so the deferred
        split-K reductions and the hook run on one of those after it has been ordered behind the other and behind the
        current stream -- nothing on the dz -> dgrad chain waits."""
        if self.stage_hook is not None:
            self.stage_hook(self, names[-1])
        if not self.wgrad_streams:
            self._hooks(names)
            return
        if self.grad_hook is None and self.update_hook is None and not self._pending_unpack:
            return
        cur = torch.cuda.current_stream(self.device)
        j = self._wg_stream(1)
        ev = torch.cuda.Event()
        ev.record(cur)
        j.wait_event(ev)
        for wg in self._wg.values():
            if wg is not j:
                j.wait_stream(wg)
        with torch.cuda.stream(j):
            self._flush_unpacks()
            self._hooks(names)

    def _notify(self, names):
        self._hooks(names)

    def _hooks(self, names):
        if self.grad_hook is not None:
            self.grad_hook(self, names)  # data parallel: the bucketer forwards to update_hook once a range is reduced
        elif self.update_hook is not None:
            if names and names[0] == "head":
                self._upd_done = 0
            end = self._flat_grad.numel() if names and names[-1] == "stems" else self.grad_prefix_end(names)
            if end > self._upd_done:
                self.update_hook(self, self._upd_done, end)
                self._upd_done = end


class _NetFn(torch.autograd.Function):
    """One autograd node for the whole network.  Parameter gradients are written straight into p.grad
    (views of the engine's flat buffer), so the node returns None for them."""

    @staticmethod
    def forward(ctx, engine, n_in, *tensors):
        xs = tensors[:n_in]
        ctx.engine = engine
        ctx.n_in = n_in
        ctx.n_params = len(tensors) - n_in
        outs = engine._forward(xs, True, ctx)
        return tuple(outs)

    @staticmethod
    def backward(ctx, *dlogits):
        eng = ctx.engine
        # lateral-aware mask ordering (computed once per plan)
        eng._mark_mask_later()
        eng._backward(list(dlogits))
        return (None, None) + (None,) * (ctx.n_in + ctx.n_params)


def _mark_mask_later(self):
    plan = self.plan
    if getattr(plan, "_marked", False):
        return
    fuse_inputs = set(id(u.x.buf) for kind, u in plan.order if kind == "fuse")
    # a block whose input is also read by a lateral conv that runs EARLIER in forward order (i.e. LATER in
    # backward order) must leave the ReLU mask to that lateral's dgrad... but the lateral's dgrad is strided
    # in 4 phases that together cover every row, so it can apply the mask.
    for kind, item in plan.order:
        if kind == "block" and id(item["x"].buf) in fuse_inputs and item["x"].parent is None:
            item["mask_later"] = True
    plan._marked = True


Engine._mark_mask_later = _mark_mask_later

"""GPU (-m gpu): batched SpecAugment kernels (csrc/specaug.cu) through the C ABI against the oracle
(oracle/spec_augment_np.py) on explicit parameters, against outputs of the real reference (tests/golden/specaug.npz),
and the on-device parameter generator against the distributions of the reference's draws."""
import ctypes as C
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from asf_b200 import _lib as L  # noqa: E402
from asf_b200.datasets import spec_augment as GA  # noqa: E402
from oracle import logmel_np as OL  # noqa: E402
from oracle import spec_augment_np as SA  # noqa: E402


def test_apply_matches_reference_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "specaug.npz"))
    n = int(g["count"])
    recs = [{"coef": g["coef%d" % i], "fmask": [tuple(r) for r in g["fmask%d" % i]], "tmask": [tuple(r) for r in g["tmask%d" % i]]}
            for i in range(n)]
    x = torch.from_numpy(np.stack([g["in%d" % i] for i in range(n)])).unsqueeze(1).cuda()
    out = GA.apply_params(x, GA.pack_params(recs, x.device))[:, 0].cpu().numpy()
    for i in range(n):
        assert np.abs(out[i] - g["ref%d" % i]).max() < 2e-3, i                      # the real reference's output
        assert np.abs(out[i] - SA.spec_augment(g["in%d" % i], recs[i])).max() < 1e-4, i  # the oracle, same parameters


def test_apply_full_size_edge_cases():
    """400 x 128 clips: no warp, large flows that clamp at both borders, negative flows, empty / full-height / adjacent
    masks, and bit-reproducibility."""
    wav = OL.synthetic_waveforms(6, 47975, 11)
    spec = np.ascontiguousarray(np.stack([OL.extract_sound_feature(w, 400)[0] for w in wav]))
    recs = [
        {"coef": (0.0, 0.0, 0.0), "fmask": [], "tmask": []},
        {"coef": (0.03, -0.008, 2.5), "fmask": [(98, 100), (50, 61)], "tmask": [(259, 271), (272, 272)]},
        {"coef": (-0.4, 0.9, -30.0), "fmask": [(0, 26)], "tmask": [(0, 24), (376, 400)]},
        {"coef": (0.5, -1.2, 55.0), "fmask": [(102, 128), (0, 1)], "tmask": []},
        {"coef": (0.0, 0.0, -3.25), "fmask": [], "tmask": [(10, 11), (11, 12)]},
        {"coef": (1e-3, 1e-3, 0.5), "fmask": [(64, 65)], "tmask": [(399, 400)]},
    ]
    x = torch.from_numpy(spec).unsqueeze(1).cuda()
    params = GA.pack_params(recs, x.device)
    out = GA.apply_params(x, params)
    again = GA.apply_params(x, params)
    assert torch.equal(out, again)
    o = out[:, 0].cpu().numpy()
    assert np.array_equal(o[0], spec[0])
    for i, r in enumerate(recs):
        ref = SA.spec_augment(spec[i], r)
        # positions are computed in fp32 on both sides; a fused multiply-add can move one by an ulp
        assert np.abs(o[i] - ref).max() < 2e-4, (i, np.abs(o[i] - ref).max())
    with pytest.raises(RuntimeError):
        L.call("asf_specaug_apply", L.ptr(x), L.ptr(x), L.ptr(params), 6, 400, 128, L.stream_ptr())


def test_parameter_generator_distributions():
    B, T, F = 4096, 400, 128
    g = torch.Generator().manual_seed(0)
    x = (torch.randn(B, 1, T, F, generator=g) - 3.0).cuda()
    aug = GA.SpecAugmenter(seed=1234)
    raw = aug.draw(x).cpu().numpy().tobytes()
    arr = (L.SpecAugParams * B).from_buffer_copy(raw)
    coef = np.array([[p.coef[j] for j in range(3)] for p in arr])
    fm = np.array([[[p.fmask[m][0], p.fmask[m][1]] for m in range(2)] for p in arr])
    tm = np.array([[[p.tmask[m][0], p.tmask[m][1]] for m in range(2)] for p in arr])
    assert np.isfinite(coef).all()
    # d = randrange(-5, 5) is 0 with probability 1/10: no warp at all
    nowarp = (np.abs(coef).sum(1) == 0).mean()
    assert 0.07 < nowarp < 0.13, nowarp
    # the flow at the control point (f = 64, t = x_c + d ~ small) is ~ d: |a_f 64 + a_0| stays within W + 1 for the bulk
    for masks, size, mx in ((fm, F, 27), (tm, T, 25)):
        z, e = masks[..., 0], masks[..., 1]
        assert (z >= 0).all() and (e <= size).all() and ((e - z) < mx).all()
        first_nonempty = (e[:, 0] > z[:, 0]).mean()
        # width w uniform in 0..mx-1, end uniform in [z, z + w): E[non-empty] = 1 - (1/mx) sum_{w>=1} 1/w - 1/mx
        expect = 1.0 - (1.0 + sum(1.0 / w for w in range(1, mx))) / mx
        assert abs(first_nonempty - expect) < 0.03, (first_nonempty, expect)
        starts = z[:, 0][e[:, 0] > z[:, 0]]
        assert starts.min() < size * 0.1 and starts.max() > size * 0.75
    # a second call draws different numbers (device counter), the same seed replays the same sequence
    raw2 = aug.draw(x).cpu().numpy().tobytes()
    assert raw2 != raw
    aug_b = GA.SpecAugmenter(seed=1234)
    assert aug_b.draw(x).cpu().numpy().tobytes() == raw and aug_b.draw(x).cpu().numpy().tobytes() == raw2
    assert GA.SpecAugmenter(seed=99).draw(x).cpu().numpy().tobytes() != raw


def test_augmenter_end_to_end_and_reference_signature():
    wav = OL.synthetic_waveforms(4, 47975, 21)
    spec = torch.from_numpy(np.ascontiguousarray(np.stack([OL.extract_sound_feature(w, 400) for w in wav]))).cuda()  # (4, 1, 400, 128)
    aug = GA.SpecAugmenter(seed=7)
    out = aug(spec)
    assert out.shape == spec.shape and torch.isfinite(out).all()
    raw = aug._state(spec)[0].cpu().numpy().tobytes()
    arr = (L.SpecAugParams * 4).from_buffer_copy(raw)
    for b in range(4):
        rec = {"coef": [arr[b].coef[j] for j in range(3)],
               "fmask": [(arr[b].fmask[m][0], arr[b].fmask[m][1]) for m in range(2)],
               "tmask": [(arr[b].tmask[m][0], arr[b].tmask[m][1]) for m in range(2)]}
        ref = SA.spec_augment(spec[b, 0].cpu().numpy(), rec)
        assert np.abs(out[b, 0].cpu().numpy() - ref).max() < 2e-4
    y = GA.spec_augment(spec[0].permute(0, 2, 1))  # reference signature: (1, F, T) in, (1, F, T) out
    assert y.shape == (1, 128, 400) and y.is_cuda

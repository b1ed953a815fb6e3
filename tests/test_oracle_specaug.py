"""CPU: the SpecAugment restatement (oracle/spec_augment_np.py) against OUTPUTS OF THE REAL REFERENCE -- the committed
vectors tests/golden/specaug.npz (made by tests/golden/make_golden.py running datasets/spec_augment.py itself) and, when
/root/reference is present, the reference functions run live with their Python / torch random streams replayed."""
import os
import random

import numpy as np
import pytest
import torch

from oracle import ref_shim
from oracle import spec_augment_np as SA


def _params(g, i):
    return {"coef": g["coef%d" % i], "fmask": [tuple(r) for r in g["fmask%d" % i]], "tmask": [tuple(r) for r in g["tmask%d" % i]]}


def test_oracle_matches_reference_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "specaug.npz"))
    n = int(g["count"])
    assert n >= 8
    kinds = set()
    for i in range(n):
        out = SA.spec_augment(g["in%d" % i], _params(g, i))
        # the reference solves its 4 x 4 spline system in fp32 with entries of 1e-10 next to entries of 1e2: its flow
        # carries ~1e-4 pixels of noise, i.e. ~1e-4 on values with unit slope
        assert np.abs(out - g["ref%d" % i]).max() < 2e-3, i
        kinds.add((int(g["draw%d" % i][1]) != 0, len(g["fmask%d" % i]), len(g["tmask%d" % i])))
    assert len(kinds) >= 3, "the fixture should cover warps, early-returning and full mask sets"


def test_draw_order_and_early_return():
    class Script:
        def __init__(self, vals):
            self.vals, self.calls = list(vals), []

        def randrange(self, a, b):
            v = self.vals.pop(0)
            assert a <= v < b, (a, b, v)
            self.calls.append((a, b))
            return v

    row = np.linspace(-5, 1, 400).astype(np.float32)
    # i, d | f, f0, end | f = 0 -> f0 drawn, then RETURN (no second freq mask) | t, t0, end | t, t0, end
    s = Script([100, -3, 5, 10, 12, 0, 7, 20, 30, 35, 1, 398, 398])
    p = SA.draw_params(s, lambda: np.eye(3), row, 128, 400)
    assert (p["i"], p["d"]) == (100, -3) and p["fmask"] == [(10, 12)] and p["tmask"] == [(30, 35), (398, 398)]
    assert s.calls[:2] == [(5, 395), (-5, 5)] and s.calls[2:5] == [(0, 27), (0, 123), (10, 15)]
    assert s.calls[5:7] == [(0, 27), (0, 128)] and s.calls[7] == (0, 25)
    # identity normal block: u = b, flow = d (b . q - 1e-10 K) / |b|^2
    b = np.array([64, row[100] - 3, 1.0])
    want = -3 * b / (b @ b)
    want[2] -= -3 * 1e-10 * SA.radial_constant(128, 400) / (b @ b)
    assert np.allclose(p["coef"], want, rtol=1e-12)


def test_masks_use_the_running_mean():
    x = np.arange(12, dtype=np.float32).reshape(3, 4)  # [F, T]
    out = SA.apply_masks(x, [(0, 1)], [(1, 2)])
    m1 = x.mean(dtype=np.float32)
    y = x.copy()
    y[0, :] = m1
    m2 = y.mean(dtype=np.float32)
    y[:, 1] = m2
    assert np.array_equal(out, y)


@pytest.mark.skipif(not ref_shim.reference_available(), reason="needs /root/reference (build container only)")
def test_oracle_matches_reference_live():
    from oracle import logmel_np as L

    sa = ref_shim.load_spec_augment()
    wav = L.synthetic_waveforms(2, 47975, 3)
    checked = 0
    for trial in range(6):
        spec = L.extract_sound_feature(wav[trial % 2], 400)[0]          # [T, F] = [400, 128], the real clip size
        x = torch.from_numpy(spec).unsqueeze(0).permute(0, 2, 1)         # (C, F, T) as at epickitchens.py:163
        random.seed(100 + trial)
        torch.manual_seed(200 + trial)
        ref = sa.spec_augment(x).permute(0, 2, 1)[0].numpy()
        random.seed(100 + trial)
        torch.manual_seed(200 + trial)
        p = SA.draw_params(random, lambda: torch.randn((1, 3, 3)).numpy(), spec[:, 64], 128, 400)
        if np.abs(p["coef"]).max() > 1.0:
            continue  # near-singular draw: the fp32 solve of the reference is dominated by rounding
        assert np.abs(SA.spec_augment(spec, p) - ref).max() < 2e-3, trial
        checked += 1
    assert checked >= 4

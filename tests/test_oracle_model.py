"""CPU: the functional torch oracle against the golden vectors produced by the real reference classes,
and (when /root/reference is present, i.e. in the build container) against the reference itself."""
import os

import numpy as np
import pytest
import torch

from oracle import model_torch as M
from oracle import ref_shim

CASES = {
    "slowfast_epic": ("slowfast", (97, 300), 2),
    "slowfast_vgg": ("slowfast", (309,), 2),
    "slow_r50": ("slow", (97, 300), 1),
}


def _tup(x):
    return x if isinstance(x, tuple) else (x,)


@pytest.mark.parametrize("tag", list(CASES))
def test_oracle_matches_reference_golden(golden_dir, tag):
    torch.set_num_threads(max(1, min(8, os.cpu_count() or 1)))
    arch, nc, B = CASES[tag]
    g = np.load(os.path.join(golden_dir, "model_%s.npz" % tag))
    cfg = M.epic_cfg(arch=arch, num_classes=nc, dropout=0.0)
    sd = M.calibrated_state(cfg)
    inputs = M.synthetic_inputs(cfg, B)
    labels = M.synthetic_labels(cfg, B)
    with torch.no_grad():
        probs = _tup(M.Net(cfg, sd).forward(inputs, training=False))
    for i, p in enumerate(probs):
        assert np.allclose(p.numpy(), g["eval_probs%d" % i], rtol=1e-4, atol=1e-7), (tag, i)
        assert np.allclose(p.sum(1).numpy(), 1.0, atol=1e-5)
    for key, freeze in (("frozen", True), ("train", False)):
        cfg.bn_freeze = freeze
        loss, preds, grads, sd2 = M.loss_and_grads(cfg, sd, inputs, labels)
        assert abs(float(loss) - float(g[key + "_loss"])) <= 1e-4 * max(1.0, abs(float(g[key + "_loss"])))
        for i, p in enumerate(_tup(preds)):
            assert np.allclose(p.detach().numpy(), g["%s_logits%d" % (key, i)], rtol=1e-3, atol=1e-4)
        names = list(g[key + "_grad_names"])
        assert sorted(grads) == names
        norms = np.array([float(grads[k].double().norm()) for k in names])
        ref = g[key + "_grad_norms"]
        assert np.allclose(norms, ref, rtol=2e-3, atol=1e-7), float(np.abs(norms / np.maximum(ref, 1e-12) - 1).max())
        for bn in M.UNFROZEN_BN:
            k = "%s_%s.running_mean" % (key, bn)
            if k in g.files:
                assert np.allclose(sd2[bn + ".running_mean"].numpy(), g[k], rtol=1e-4, atol=1e-6)
                assert np.allclose(sd2[bn + ".running_var"].numpy(), g[k.replace("mean", "var")], rtol=1e-4, atol=1e-6)
        last = str(g[key + "_last_bn_name"])
        assert np.allclose(sd2[last].numpy(), g[key + "_last_bn_running_var"], rtol=1e-4, atol=1e-6)


def test_parameter_counts():
    # 26 882 293 parameters / 664 state_dict entries; SLOW_R50 24 894 157; VGG 26 679 453
    for arch, nc, n_params, n_sd in (("slowfast", (97, 300), 26882293, 664), ("slow", (97, 300), 24894157, 322),
                                     ("slowfast", (309,), 26679453, 662)):
        sd = M.init_state_dict(M.epic_cfg(arch=arch, num_classes=nc), 0)
        assert len(sd) == n_sd
        assert sum(v.numel() for k, v in sd.items() if "running_" not in k and "num_batches" not in k) == n_params


@pytest.mark.skipif(not ref_shim.reference_available(), reason="reference tree only exists in the build container")
def test_oracle_bit_exact_against_reference_code():
    ref = ref_shim.load_reference()
    rcfg = ref_shim.reference_cfg()
    rcfg.MODEL.DROPOUT_RATE = 0.0
    torch.manual_seed(0)
    rm = ref.build_model(rcfg)
    cfg = M.epic_cfg(dropout=0.0)
    sd = M.init_state_dict(cfg, 0)
    rsd = rm.state_dict()
    assert list(sd) == list(rsd) and all(torch.equal(sd[k], rsd[k]) for k in sd)
    psd = M.calibrated_state(cfg)
    rm.load_state_dict(psd)
    inputs = M.synthetic_inputs(cfg, 1)
    rm.eval()
    with torch.no_grad():
        a = rm([t.clone() for t in inputs])
        b = M.Net(cfg, psd).forward(inputs, training=False)
    assert all(torch.equal(x, y) for x, y in zip(a, b))
    # the real pack_pathway_output against the restated indices
    spec = torch.arange(400, dtype=torch.float32).view(1, 400, 1).repeat(1, 1, 2)
    slow, fast = ref.pack_pathway_output(rcfg, spec)
    from oracle import logmel_np as L
    assert np.array_equal(slow[0, :, 0].long().numpy(), L.slow_indices(400, 4))

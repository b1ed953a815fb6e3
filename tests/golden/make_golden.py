#!/usr/bin/env python3
"""Generates the committed golden vectors under tests/golden/ by running the REFERENCE ITSELF.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden.py

* model_*.npz  -- outputs of the real ``audio_slowfast.models`` classes (imported unmodified through
  oracle/ref_shim.py) on seeded synthetic inputs with the deterministic "calibrated" state of
  oracle/model_torch.py loaded through ``load_state_dict``: eval probabilities, training logits,
  loss, per-parameter gradient norms + leading values, running-stat updates.
* pack_pathway.npz -- ``pack_pathway_output`` of the real reference (datasets/utils.py:43-72):
  gather indices for several (T, alpha).
* logmel.npz -- the reference's ``_log_specgram`` cannot run here (librosa is not installed and
  is not vendored), so these vectors come from the numpy restatement oracle/logmel_np.py,
  cross-checked against torchaudio at generation time ("parity unpinned" against real librosa).
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import logmel_np as L  # noqa: E402
from oracle import model_torch as M  # noqa: E402
from oracle import ref_shim  # noqa: E402


def grads_summary(named_grads):
    names = sorted(named_grads)
    norms = np.array([float(named_grads[k].double().norm()) for k in names], dtype=np.float64)
    heads = np.stack([np.pad(named_grads[k].reshape(-1)[:8].numpy(), (0, max(0, 8 - named_grads[k].numel())))
                      for k in names]).astype(np.float32)
    return names, norms, heads


def model_case(tag, arch, num_classes, batch):
    ref = ref_shim.load_reference()
    cfg = M.epic_cfg(arch=arch, num_classes=num_classes, dropout=0.0)
    sd = M.calibrated_state(cfg)
    inputs = M.synthetic_inputs(cfg, batch)
    labels = M.synthetic_labels(cfg, batch)
    out = {}
    rcfg = ref_shim.reference_cfg(arch=arch, num_classes=num_classes)
    rcfg.MODEL.DROPOUT_RATE = 0.0
    model = ref.build_model(rcfg)
    model.load_state_dict(sd)
    model.eval()
    with torch.no_grad():
        probs = model([t.clone() for t in inputs])
    probs = probs if isinstance(probs, tuple) else (probs,)
    for i, p in enumerate(probs):
        out["eval_probs%d" % i] = p.numpy()
    for freeze in (True, False):
        model = ref.build_model(rcfg)
        model.load_state_dict(sd)
        model.train()
        if freeze:  # tools/train_net.py:709-710 and :390-392
            model.freeze_fn("bn_parameters")
            model.freeze_fn("bn_statistics")
        preds = model([t.clone() for t in inputs])
        loss, _, _ = M.compute_loss(preds, labels)
        loss.backward()
        preds = preds if isinstance(preds, tuple) else (preds,)
        key = "frozen" if freeze else "train"
        for i, p in enumerate(preds):
            out["%s_logits%d" % (key, i)] = p.detach().numpy()
        out["%s_loss" % key] = np.float64(loss.item())
        g = {k: p.grad.detach() for k, p in model.named_parameters() if p.grad is not None}
        names, norms, heads = grads_summary(g)
        out["%s_grad_names" % key] = np.array(names)
        out["%s_grad_norms" % key] = norms
        out["%s_grad_heads" % key] = heads
        rsd = model.state_dict()
        for bn in M.UNFROZEN_BN:
            if bn + ".running_mean" in rsd:
                out["%s_%s.running_mean" % (key, bn)] = rsd[bn + ".running_mean"].numpy()
                out["%s_%s.running_var" % (key, bn)] = rsd[bn + ".running_var"].numpy()
        last_bn = [k for k in rsd if k.endswith("c_bn.running_var")][-1]
        out["%s_last_bn_name" % key] = np.array(last_bn)
        out["%s_last_bn_running_var" % key] = rsd[last_bn].numpy()
    np.savez_compressed(os.path.join(HERE, "model_%s.npz" % tag), **out)
    print(tag, {k: (v.shape if hasattr(v, "shape") else v) for k, v in list(out.items())[:4]})


def pack_case():
    ref = ref_shim.load_reference()
    rcfg = ref_shim.reference_cfg()
    out = {}
    for T, alpha in [(400, 4), (401, 4), (200, 4), (800, 4), (2000, 4), (256, 8), (100, 2)]:
        rcfg.SLOWFAST.ALPHA = alpha
        spec = torch.arange(T, dtype=torch.float32).view(1, T, 1).expand(1, T, 3).contiguous()
        slow, fast = ref.pack_pathway_output(rcfg, spec)
        assert fast is spec
        out["T%d_a%d" % (T, alpha)] = slow[0, :, 0].long().numpy()
    np.savez_compressed(os.path.join(HERE, "pack_pathway.npz"), **out)
    print("pack", {k: v.shape for k, v in out.items()})


def logmel_case():
    import torchaudio

    out = {}
    wav = L.synthetic_waveforms(2, 47975, 1234)
    t = np.arange(47975) / 24000.0
    cases = {
        "noise0": wav[0],
        "noise1": wav[1],
        "sine1k": (0.5 * np.sin(2 * np.pi * 1000.0 * t)).astype(np.float32),
        "zeros": np.zeros(47975, np.float32),
        "len48000": L.synthetic_waveforms(1, 48000, 99)[0],
        "len23975": L.synthetic_waveforms(1, 23975, 98)[0],
        "len5000": L.synthetic_waveforms(1, 5000, 97)[0],
        "len119": L.synthetic_waveforms(1, 119, 96)[0],
    }
    ms = torchaudio.transforms.MelSpectrogram(
        sample_rate=24000, n_fft=2048, win_length=240, hop_length=120, f_min=0, f_max=12000, n_mels=128,
        power=1.0, center=True, pad_mode="constant", norm=None, mel_scale="htk")
    for k, w in cases.items():
        a = L.log_specgram(w)
        b = torch.log(ms(torch.from_numpy(w)) + 1e-6).T.numpy()
        assert a.shape == b.shape, (k, a.shape, b.shape)
        rel_log, rel_lin, frac = L.logmel_parity(a, b)
        assert rel_log < 1e-4 and rel_lin < 1e-5, (k, rel_log, rel_lin)
        out[k] = a.astype(np.float32)
        print("logmel", k, a.shape, "vs torchaudio: rel_log(<=80dB) %.2e  rel_lin %.2e  well-conditioned %.3f"
              % (rel_log, rel_lin, frac))
    out["mel_basis_rowsum"] = L.mel_filterbank().sum(1)
    out["mel_basis_nnz"] = (L.mel_filterbank() > 0).sum(1)
    np.savez_compressed(os.path.join(HERE, "logmel.npz"), **out)


def specaug_case():
    """specaug.npz -- the REAL reference spec_augment (datasets/spec_augment.py:144-191) on small seeded clips, with the
    random draws it made (replayed from the same Python / torch seeds in the order the reference consumes them, the
    3 x 3 normal block included), so that the oracle and the GPU kernel can be checked against reference OUTPUTS."""
    import random

    from oracle import spec_augment_np as SA

    sa = ref_shim.load_spec_augment()
    out = {}
    F, T = 32, 64
    rng = np.random.default_rng(5)
    n = 0
    for trial in range(40):
        spec = (rng.standard_normal((T, F)) * 1.5 - 3.0).astype(np.float32)      # [T, F], log-mel-like values
        x = torch.from_numpy(spec).unsqueeze(0).permute(0, 2, 1)                  # (C, F, T) as in epickitchens.py:163
        random.seed(1000 + trial)
        torch.manual_seed(2000 + trial)
        ref = sa.spec_augment(x, F=9, T=12).permute(0, 2, 1)[0].numpy()
        random.seed(1000 + trial)
        torch.manual_seed(2000 + trial)
        p = SA.draw_params(random, lambda: torch.randn((1, 3, 3)).numpy(), spec[:, F // 2], F, T, Fmax=9, Tmax=12)
        if np.abs(p["coef"]).max() > 2.0:   # a near-singular draw: the reference's fp32 solve is noise there
            continue
        out["in%d" % n], out["ref%d" % n] = spec, ref
        out["coef%d" % n] = np.asarray(p["coef"], dtype=np.float64)
        out["fmask%d" % n] = np.asarray(p["fmask"], dtype=np.int32).reshape(-1, 2)
        out["tmask%d" % n] = np.asarray(p["tmask"], dtype=np.int32).reshape(-1, 2)
        out["draw%d" % n] = np.array([p["i"], p["d"]], dtype=np.int32)
        n += 1
        if n == 12:
            break
    out["count"] = np.array(n)
    np.savez_compressed(os.path.join(HERE, "specaug.npz"), **out)
    print("specaug.npz: %d cases" % n)


if __name__ == "__main__":
    torch.set_num_threads(8)
    logmel_case()
    pack_case()
    model_case("slowfast_epic", "slowfast", (97, 300), 2)
    model_case("slowfast_vgg", "slowfast", (309,), 2)
    model_case("slow_r50", "slow", (97, 300), 1)
    specaug_case()

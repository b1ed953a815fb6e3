"""GPU (-m gpu): the rows next to the hot path through the C ABI -- device metrics and the multi-view
test ensemble against the reference's formulas restated with plain torch ops, clip slicing against numpy slicing, the
pinned-host ring, precise BatchNorm against the oracle network, and the asynchronous checkpoint writer."""
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from asf_b200 import _lib as L  # noqa: E402
from asf_b200.datasets import staging  # noqa: E402
from asf_b200.utils import bn_helper, checkpoint as CK, meters, metrics  # noqa: E402
from oracle import model_torch as M  # noqa: E402


# ---- the reference's metric formulas (utils/metrics.py:11-41, :80-108), restated with torch ops on the CPU
def ref_topks_correct(preds, labels, ks):
    _, idx = torch.topk(preds, max(ks), dim=1, largest=True, sorted=True)
    hit = idx.t().eq(labels.view(1, -1).expand_as(idx.t()))
    return [float(hit[:k, :].float().sum()) for k in ks]


def ref_multitask_topks_correct(preds, labels, ks):
    max_k = max(ks)
    all_correct = torch.zeros(max_k, labels[0].size(0), dtype=torch.uint8)
    for output, label in zip(preds, labels):
        _, idx = output.topk(max_k, dim=1, largest=True, sorted=True)
        all_correct.add_(idx.t().eq(label.view(1, -1).expand_as(idx.t())).to(torch.uint8))
    return [float(torch.ge(all_correct[:k].float().sum(0), len(preds)).float().sum(0)) for k in ks]


def test_topk_metrics_match_reference_formulas():
    g = torch.Generator().manual_seed(0)
    for B in (1, 7, 64, 1000):
        pv, pn = torch.randn(B, 97, generator=g), torch.randn(B, 300, generator=g)
        lv, ln = torch.randint(0, 97, (B,), generator=g), torch.randint(0, 300, (B,), generator=g)
        # make a good share of the labels actually rank high
        pv[torch.arange(B), lv] += 2.5 * torch.rand(B, generator=g)
        pn[torch.arange(B), ln] += 3.5 * torch.rand(B, generator=g)
        got_v = [float(x) for x in metrics.topks_correct(pv.cuda(), lv.cuda(), (1, 5))]
        got_n = [float(x) for x in metrics.topks_correct(pn.cuda(), ln.cuda(), (1, 5))]
        got_a = [float(x) for x in metrics.multitask_topks_correct((pv.cuda(), pn.cuda()), (lv.cuda(), ln.cuda()), (1, 5))]
        assert got_v == ref_topks_correct(pv, lv, (1, 5)) and got_n == ref_topks_correct(pn, ln, (1, 5))
        assert got_a == ref_multitask_topks_correct((pv, pn), (lv, ln), (1, 5))
        acc = metrics.topk_accuracies(pv.cuda(), lv.cuda(), (1, 5))
        assert acc[0].is_cuda and abs(float(acc[1]) - 100.0 * got_v[1] / B) < 1e-4
        err = metrics.topk_errors(pv.cuda(), lv.cuda(), (1, 5))
        assert abs(float(err[0]) + float(acc[0]) - 100.0) < 1e-4
    # ties: the lower class index ranks first (a fixed rule); a label tied with 5 lower-index classes is not in the top 5
    p = torch.zeros(2, 10)
    assert [float(x) for x in metrics.topks_correct(p.cuda(), torch.tensor([0, 5]).cuda(), (1, 5))] == [1.0, 1.0]
    assert [float(x) for x in metrics.topks_correct(p.cuda(), torch.tensor([4, 9]).cuda(), (1, 5))] == [0.0, 1.0]
    # bf16 / strided predictions and int32 labels are accepted
    pv = torch.randn(9, 97).cuda()
    lv = torch.randint(0, 97, (9,)).cuda()
    a = metrics.topk_counts(pv.bfloat16(), lv.int(), (1, 5))
    b = metrics.topk_counts(pv.bfloat16().float(), lv, (1, 5))
    assert torch.equal(a, b)


def test_train_metrics_accumulate_without_sync():
    g = torch.Generator().manual_seed(1)
    tm = metrics.TrainMetrics(torch.device("cuda"))
    tot, hits, loss_sum = 0, np.zeros(6), 0.0
    for it in range(5):
        B = 8 + it
        pv, pn = torch.randn(B, 97, generator=g), torch.randn(B, 300, generator=g)
        lv, ln = torch.randint(0, 97, (B,), generator=g), torch.randint(0, 300, (B,), generator=g)
        pv[torch.arange(B), lv] += 3 * torch.rand(B, generator=g)
        pn[torch.arange(B), ln] += 4 * torch.rand(B, generator=g)
        loss = torch.tensor(1.0 + it)
        tm.update((pv.cuda(), pn.cuda()), {"verb": lv.cuda(), "noun": ln.cuda()}, loss.cuda(), loss.cuda() * 2, loss.cuda() * 3)
        hits += np.array(ref_topks_correct(pv, lv, (1, 5)) + ref_topks_correct(pn, ln, (1, 5))
                         + ref_multitask_topks_correct((pv, pn), (lv, ln), (1, 5)))
        tot += B
        loss_sum += float(loss) * B
    s = tm.summary()
    assert s["samples"] == tot and abs(s["loss"] - loss_sum / tot) < 1e-9 and abs(s["noun_loss"] - 3 * loss_sum / tot) < 1e-9
    for key, h in zip(("verb_top1_acc", "verb_top5_acc", "noun_top1_acc", "noun_top5_acc", "top1_acc", "top5_acc"), hits):
        assert abs(s[key] - 100.0 * h / tot) < 1e-9, key


@pytest.mark.parametrize("method", ["sum", "max"])
def test_epic_test_meter_ensemble(method):
    """10-view ensemble: clips arrive shuffled over ragged batches; ensembled scores, labels, per-view scores and the
    verb / noun / action accuracies against the reference's per-sample Python loop (meters.py:1641-1755) restated."""
    V, K, Cv, Cn = 13, 10, 97, 300
    g = torch.Generator().manual_seed(3)
    lv, ln = torch.randint(0, Cv, (V,), generator=g), torch.randint(0, Cn, (V,), generator=g)
    pv = torch.softmax(torch.randn(V * K, Cv, generator=g) * 2, 1)
    pn = torch.softmax(torch.randn(V * K, Cn, generator=g) * 2, 1)
    ids = torch.randperm(V * K, generator=g)
    for v in range(V):  # a classifier that is often right
        pv[v * K:(v + 1) * K, lv[v]] += 0.05 * torch.rand(K, generator=g)
        pn[v * K:(v + 1) * K, ln[v]] += 0.03 * torch.rand(K, generator=g)
    meter = meters.EPICTestMeter(V, K, (Cv, Cn), overall_iters=0, ensemble_method=method)
    ref_v, ref_n = torch.zeros(V, Cv), torch.zeros(V, Cn)
    pos = 0
    for bs in (7, 32, 1, 50, 40):
        sel = ids[pos:pos + bs]
        pos += bs
        vid = sel // K
        meter.update_stats((pv[sel].cuda(), pn[sel].cuda()), (lv[vid].cuda(), ln[vid].cuda()),
                           {"narration_id": ["n%d" % int(i) for i in vid]}, sel.cuda())
        for j, cid in enumerate(sel.tolist()):  # the reference's loop
            if method == "sum":
                ref_v[cid // K] += pv[cid]
                ref_n[cid // K] += pn[cid]
            else:
                ref_v[cid // K] = torch.max(ref_v[cid // K], pv[cid])
                ref_n[cid // K] = torch.max(ref_n[cid // K], pn[cid])
    assert pos == V * K
    (ev, en), (cv, cn), (gl_v, gl_n), meta = meter.finalize_metrics((1, 5))
    assert meter.complete and list(meta) == ["n%d" % v for v in range(V)]
    assert np.allclose(ev, ref_v.numpy(), atol=1e-6) and np.allclose(en, ref_n.numpy(), atol=1e-6)
    assert np.array_equal(cv.reshape(V * K, Cv), pv.numpy()) and np.array_equal(gl_v, lv.numpy()) and np.array_equal(gl_n, ln.numpy())
    want = {"verb": ref_topks_correct(ref_v, lv, (1, 5)), "noun": ref_topks_correct(ref_n, ln, (1, 5)),
            "action": ref_multitask_topks_correct((ref_v, ref_n), (lv, ln), (1, 5))}
    for name, w in want.items():
        for k, c in zip((1, 5), w):
            assert meter.stats["%s_top%d_acc" % (name, k)] == "{:.2f}".format(100.0 * c / V)


def test_resident_audio_crops_and_views():
    rng = np.random.default_rng(0)
    audio = {"P01_01": rng.standard_normal(300000).astype(np.float32), "P02_07": rng.standard_normal(47000).astype(np.float32),
             "P03_03": rng.standard_normal(100001).astype(np.float32)}
    ra = staging.ResidentAudio(audio)
    clip = 47976
    names = ["P01_01", "P03_03", "P01_01", "P02_07", "P03_03"]
    starts = [0, 52025, 252025, 10, 99990]
    valid = torch.zeros(5, dtype=torch.int32, device="cuda")
    out = ra.crop(names, starts, clip - 1, valid=valid).cpu().numpy()
    for b, (nm, s) in enumerate(zip(names, starts)):
        ref = audio[nm][s:s + clip - 1]          # Python slicing clamps at the end of the array
        assert int(valid[b]) == len(ref)
        assert np.array_equal(out[b, :len(ref)], ref) and not out[b, len(ref):].any()
    views = ra.views("P01_01", clip, 10).cpu().numpy()
    st = np.linspace(0, 300000 - clip, 10)
    for v in range(10):
        assert np.array_equal(views[v], audio["P01_01"][int(st[v]):int(st[v]) + clip - 1])
    # a record inside a longer video (start_audio_sample, num_audio_samples of the reference's AudioRecord)
    sub = ra.views("P01_01", clip, 10, num_samples=60000, start_sample=1000).cpu().numpy()
    st = 1000 + np.linspace(0, 60000 - clip, 10)
    assert np.array_equal(sub[9], audio["P01_01"][int(st[9]):int(st[9]) + clip - 1])


def test_waveform_ring_overlapped_copies():
    B, n = 4, 47975
    ring = staging.WaveformRing(B, n, slots=2)
    rng = np.random.default_rng(1)
    batches = [(rng.standard_normal((B, n)).astype(np.float32), {"verb": rng.integers(0, 97, B), "noun": rng.integers(0, 300, B)})
               for _ in range(5)]
    ring.put(0, *batches[0])
    for i in range(5):
        if i + 1 < 5:
            ring.put((i + 1) % 2, *batches[i + 1])   # next copy in flight while batch i is consumed
        w, lab = ring.get(i % 2)
        got = (w.clone(), lab["verb"].clone(), lab["noun"].clone())
        ring.release(i % 2)
        torch.cuda.synchronize()
        assert np.array_equal(got[0].cpu().numpy(), batches[i][0])
        assert np.array_equal(got[1].cpu().numpy(), batches[i][1]["verb"]) and np.array_equal(got[2].cpu().numpy(), batches[i][1]["noun"])


def _model(dtype, freeze, sd):
    from asf_b200.config.defaults import epic_slowfast_cfg
    from asf_b200.models import build_model
    cfg = epic_slowfast_cfg((97, 300), freeze)
    cfg.MODEL.DROPOUT_RATE = 0.0
    cfg.B200.COMPUTE_DTYPE = dtype
    cfg.NUM_GPUS = 1
    model = build_model(cfg)
    model.load_state_dict(sd)
    return cfg, model


def test_precise_bn_matches_oracle():
    """compute_and_update_bn_stats on the engine model (all BatchNorm in training mode) against the same algorithm run
    on the fp32 oracle network: every running_mean / running_var after 3 batches."""
    ocfg = M.epic_cfg(dropout=0.0, bn_freeze=False)
    sd = M.calibrated_state(ocfg)
    cfg, model = _model("float32", False, sd)
    model.train()
    batches = [M.synthetic_inputs(ocfg, 2, seed=700 + i) for i in range(3)]
    n = bn_helper.compute_and_update_bn_stats(model, (([t.clone() for t in b], None, None) for b in batches), num_batches=3)
    assert n == 3
    # oracle: momentum-1 forwards, running averages of E[x] and E[x^2]
    names = [k[:-len(".running_mean")] for k in sd if k.endswith("running_mean")]
    mean = {k: torch.zeros_like(sd[k + ".running_mean"]) for k in names}
    sq = {k: torch.zeros_like(sd[k + ".running_mean"]) for k in names}
    for i, b in enumerate(batches):
        s2 = {k: v.clone() for k, v in sd.items()}
        with torch.no_grad():
            M.Net(ocfg, s2, bn_momentum=1.0).forward(b, training=True)
        for k in names:
            mean[k] += (s2[k + ".running_mean"] - mean[k]) / (i + 1)
            sq[k] += (s2[k + ".running_var"] + s2[k + ".running_mean"] ** 2 - sq[k]) / (i + 1)
    msd = model.state_dict()
    worst = 0.0
    for k in names:
        var = sq[k] - mean[k] ** 2
        em = float((msd[k + ".running_mean"].cpu() - mean[k]).abs().max() / (mean[k].abs().max() + 1e-6))
        ev = float((msd[k + ".running_var"].cpu() - var).abs().max() / (var.abs().max() + 1e-6))
        worst = max(worst, em, ev)
    assert worst < 2e-3, worst
    assert all(m.momentum == 0.1 for m in model.modules() if isinstance(m, torch.nn.BatchNorm2d))


def test_async_checkpoint_writer(tmp_path):
    from asf_b200.optim import FusedSGD
    ocfg = M.epic_cfg(dropout=0.0)
    sd = M.calibrated_state(ocfg)
    cfg, model = _model("bfloat16", True, sd)
    model.train()
    model.freeze_fn("bn_parameters")
    model.freeze_fn("bn_statistics")
    opt = FusedSGD([p for p in model.parameters() if p.requires_grad], lr=0.01, momentum=0.9, nesterov=True)
    x = [t.cuda() for t in M.synthetic_inputs(ocfg, 2)]
    y = {k: v.cuda() for k, v in M.synthetic_labels(ocfg, 2).items()}
    loss, _, _ = M.compute_loss(model(x), y)
    loss.backward()
    opt.step()
    snap = {k: v.detach().clone() for k, v in model.state_dict().items()}
    w = CK.AsyncCheckpointWriter()
    path = w.save(str(tmp_path), model, opt, 4, cfg)
    w.fence()
    # the training stream goes on and changes the weights while the file is being written
    opt.zero_grad(set_to_none=True)
    loss, _, _ = M.compute_loss(model(x), y)
    loss.backward()
    opt.step()
    w.wait()
    assert path.endswith("checkpoint_epoch_00005.pyth") and os.path.exists(path) and not os.path.exists(path + ".tmp")
    ck = torch.load(path, map_location="cpu")
    assert set(ck) == {"epoch", "model_state", "optimizer_state", "cfg"} and ck["epoch"] == 4 and len(ck["model_state"]) == 664
    for k, v in snap.items():
        assert torch.equal(ck["model_state"][k], v.cpu()), k   # the snapshot, not the later weights
    assert not torch.equal(model.state_dict()["head.projection_verb.weight"].cpu(), ck["model_state"]["head.projection_verb.weight"])
    assert len(ck["optimizer_state"]["state"]) > 0
    cfg2, model2 = _model("bfloat16", True, M.init_state_dict(ocfg, 1))
    assert CK.load_checkpoint(path, model2, data_parallel=False) == 4
    model.load_state_dict(snap)
    model.eval()
    model2.eval()
    with torch.no_grad():
        a, b = model(x), model2(x)
    assert all(torch.equal(p, q) for p, q in zip(a, b))


def test_train_step_with_specaugment_in_the_graph():
    """SpecAugment inside the captured training step: every replay draws new masks (device-side counter), the loss stays
    finite, and the augmented spectrogram differs from the plain one exactly where the masks / warp say."""
    from asf_b200.datasets.spec_augment import SpecAugmenter
    from asf_b200.optim import FusedSGD
    from asf_b200.train_step import TrainStep
    from oracle import logmel_np as OL
    B, N = 4, 47975
    ocfg = M.epic_cfg(dropout=0.0)
    sd = M.calibrated_state(ocfg)
    cfg, model = _model("bfloat16", True, sd)
    model.train()
    model.freeze_fn("bn_parameters")
    model.freeze_fn("bn_statistics")
    opt = FusedSGD([p for p in model.parameters() if p.requires_grad], lr=0.01, momentum=0.9, nesterov=True)
    aug = SpecAugmenter(seed=11)
    step = TrainStep(cfg, model, opt, B, N, graph=True, warmup=3, augment=aug)
    wav = torch.from_numpy(OL.synthetic_waveforms(B, N, 5)).cuda()
    lab = {k: v.cuda() for k, v in M.synthetic_labels(ocfg, B).items()}
    seen = []
    for it in range(3):
        loss = float(step(wav, lab))
        assert np.isfinite(loss)
        key = (B, 1, 400, 128)
        params = [v for k, v in aug._bufs.items() if k[1] == key][0][0]
        seen.append(params.cpu().numpy().tobytes())
    assert len(set(seen)) == 3, "every replay must draw fresh augmentation parameters"
    counter = list(aug._counter.values())[0]
    assert int(counter) >= 3


# ---- the classifier head and the loss (csrc/head.cu) against torch ops
@pytest.mark.parametrize("B,K,Ns", [(64, 2304, (97, 300)), (5, 2304, (309,)), (130, 2048, (97, 300)), (2, 96, (3, 5))])
def test_head_forward_loss_backward_match_torch(B, K, Ns):
    """asf_dropout_apply / asf_head_fwd / asf_softmax_ce / asf_head_bwd = dropout + nn.Linear per task +
    F.cross_entropy per task, mean over tasks (head_helper.py:70-109, train_utils.py:76-101) and their autograd
    gradients, for the EPIC (two tasks), VGG-Sound (one task) and ragged shapes, with and without accumulation."""
    import ctypes as C
    g = torch.Generator(device="cuda").manual_seed(B * 7 + K)
    dev = "cuda"
    feat = torch.randn(B, K, device=dev, generator=g)
    u = torch.rand(B, K, device=dev, generator=g)
    keep = 0.5
    Ws = [torch.randn(n, K, device=dev, generator=g) * 0.05 for n in Ns]
    bs = [torch.randn(n, device=dev, generator=g) for n in Ns]
    labels = [torch.randint(0, n, (B,), device=dev, generator=g) for n in Ns]
    st = L.stream_ptr()
    # forward
    y, mask = torch.empty_like(feat), torch.empty_like(feat)
    L.call("asf_dropout_apply", L.ptr(feat), L.ptr(u), L.ptr(y), L.ptr(mask), feat.numel(), C.c_float(keep), st)
    ref_mask = (u < keep).float()
    assert torch.equal(mask, ref_mask) and torch.equal(y, feat * ref_mask * (1.0 / keep))
    outs = [torch.empty(B, n, device=dev) for n in Ns]
    L.call("asf_head_fwd", L.ptr(y), K, B, K, len(Ns), L.ptr_array(Ws), L.ptr_array(bs), L.int_array(Ns),
           L.ptr_array(outs), L.i64_array(Ns), st)
    # torch reference with autograd
    f_ref = feat.clone().requires_grad_(True)
    Wr = [w.clone().requires_grad_(True) for w in Ws]
    br = [b.clone().requires_grad_(True) for b in bs]
    yr = f_ref * ref_mask * (1.0 / keep)
    lr = [F.linear(yr, w, b) for w, b in zip(Wr, br)]
    for o, r in zip(outs, lr):
        assert torch.allclose(o, r, rtol=1e-5, atol=1e-5 * float(r.abs().max())), float((o - r).abs().max())
    ces = [F.cross_entropy(r, l) for r, l in zip(lr, labels)]
    loss_ref = sum(ces) / len(ces)
    loss_ref.backward()
    # fused loss + gradient w.r.t. the logits (computed from OUR logits)
    loss = torch.empty((), device=dev)
    parts = torch.empty(2, device=dev)
    dls = [torch.empty(B, n + 3, device=dev)[:, :n] for n in Ns]  # pitched rows
    scratch, ticket = torch.empty(len(Ns) * B, device=dev), torch.zeros(1, dtype=torch.int32, device=dev)
    L.call("asf_softmax_ce", len(Ns), L.ptr_array(outs), L.i64_array(Ns), L.int_array(Ns), L.ptr_array(labels), B,
           L.ptr(loss), L.ptr(parts), L.ptr_array(dls), L.i64_array([d.stride(0) for d in dls]),
           L.ptr(scratch), L.ptr(ticket), st)
    assert int(ticket) == 0  # self-resetting
    assert abs(float(loss) - float(loss_ref)) <= 1e-5 * abs(float(loss_ref))
    for h, ce in enumerate(ces):
        assert abs(float(parts[h]) - float(ce)) <= 1e-5 * abs(float(ce))
    # backward of the head: plain and accumulating into existing gradients
    for acc in (0, 1):
        dW = [torch.full_like(w, 0.25) for w in Ws]
        db = [torch.full_like(b, -0.5) for b in bs]
        dfeat = torch.full_like(feat, 7.0)
        L.call("asf_head_bwd", len(Ns), L.ptr_array(dls), L.i64_array([d.stride(0) for d in dls]), L.int_array(Ns),
               L.ptr_array(Ws), L.ptr(y), K, L.ptr(mask), C.c_float(1.0 / keep), B, K, L.ptr_array(dW),
               L.int_array([acc] * len(Ns)), L.ptr_array(db), L.int_array([acc] * len(Ns)), L.ptr(dfeat), K, st)
        torch.cuda.synchronize()
        for h in range(len(Ns)):
            rw, rb = Wr[h].grad + (0.25 if acc else 0.0), br[h].grad + (-0.5 if acc else 0.0)
            assert float((dW[h] - rw).abs().max()) <= 2e-5 * float(Wr[h].grad.abs().max()) + 1e-7, (acc, h)
            assert float((db[h] - rb).abs().max()) <= 2e-5 * float(br[h].grad.abs().max()) + 1e-7, (acc, h)
        assert float((dfeat - f_ref.grad).abs().max()) <= 2e-5 * float(f_ref.grad.abs().max()) + 1e-9
    # one head without a weight gradient (frozen parameter) and no dfeat
    dW = [None] + [torch.zeros_like(w) for w in Ws[1:]]
    L.call("asf_head_bwd", len(Ns), L.ptr_array(dls), L.i64_array([d.stride(0) for d in dls]), L.int_array(Ns),
           L.ptr_array(Ws), L.ptr(y), K, None, C.c_float(1.0), B, K, L.ptr_array(dW), L.int_array([0] * len(Ns)),
           L.ptr_array([None] * len(Ns)), L.int_array([0] * len(Ns)), None, 0, st)
    torch.cuda.synchronize()


def test_fused_cross_entropy_autograd_matches_torch():
    """train_step.compute_loss on CUDA logits = tools/train_utils.py:76-101 (value, per-task values, gradients through
    autograd, a scaled backward, and a head whose logits need no gradient)."""
    from asf_b200.train_step import compute_loss
    g = torch.Generator(device="cuda").manual_seed(5)
    B = 16
    lv = torch.randn(B, 97, device="cuda", generator=g, requires_grad=True)
    ln = torch.randn(B, 300, device="cuda", generator=g, requires_grad=True)
    labels = {"verb": torch.randint(0, 97, (B,), device="cuda", generator=g),
              "noun": torch.randint(0, 300, (B,), device="cuda", generator=g)}
    loss, cv, cn = compute_loss((lv, ln), labels)
    (3.0 * loss).backward()
    rv, rn = lv.detach().clone().requires_grad_(True), ln.detach().clone().requires_grad_(True)
    rcv, rcn = F.cross_entropy(rv, labels["verb"]), F.cross_entropy(rn, labels["noun"])
    rl = 0.5 * (rcv + rcn)
    (3.0 * rl).backward()
    assert abs(float(loss) - float(rl)) < 1e-5 and abs(float(cv) - float(rcv)) < 1e-5 and abs(float(cn) - float(rcn)) < 1e-5
    assert torch.allclose(lv.grad, rv.grad, rtol=1e-5, atol=1e-7) and torch.allclose(ln.grad, rn.grad, rtol=1e-5, atol=1e-7)
    # single task, logits that are a non-leaf expression, and a detached second call
    x = torch.randn(B, 309, device="cuda", generator=g, requires_grad=True)
    y = torch.randint(0, 309, (B,), device="cuda", generator=g)
    l1, _, _ = compute_loss(x * 2.0, y)
    l1.backward()
    xr = x.detach().clone().requires_grad_(True)
    F.cross_entropy(xr * 2.0, y).backward()
    assert torch.allclose(x.grad, xr.grad, rtol=1e-5, atol=1e-7)
    with torch.no_grad():
        l2, _, _ = compute_loss(x, y)
    assert abs(float(l2) - float(F.cross_entropy(x, y))) < 1e-5

"""Debug aid (GPU): per-tensor gradient errors of the engine vs the oracle, in backward order."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "auditory-slow-fast_b200")):
    sys.path.insert(0, p)
import torch

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
from oracle import model_torch as M
from asf_b200.config.defaults import epic_slowfast_cfg
from asf_b200.models import build_model


def rel(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    return float((a - b).norm() / (b.norm() + 1e-20))


def main():
    freeze = (sys.argv[1] if len(sys.argv) > 1 else "frozen") == "frozen"
    B = int(sys.argv[2]) if len(sys.argv) > 2 else 2
    ocfg = M.epic_cfg(dropout=0.0, bn_freeze=freeze)
    kind = sys.argv[3] if len(sys.argv) > 3 else "calibrated"
    sd = M.calibrated_state(M.epic_cfg(dropout=0.0)) if kind == "calibrated" else M.init_state_dict(M.epic_cfg(dropout=0.0), 0)
    inputs = M.synthetic_inputs(ocfg, B)
    labels = M.synthetic_labels(ocfg, B)
    rloss, rpreds, rgrads, _ = M.loss_and_grads(ocfg, sd, inputs, labels)
    # the oracle on the GPU (fp32, no TF32)
    sdg = {k: v.cuda() for k, v in sd.items()}
    gloss, gpreds, ggrads, _ = M.loss_and_grads(ocfg, sdg, [t.cuda() for t in inputs], {k: v.cuda() for k, v in labels.items()})
    cols = {"oracle_gpu": ggrads}
    with torch.autocast("cuda", dtype=torch.bfloat16):
        aloss, apreds, agrads, _ = M.loss_and_grads(ocfg, sdg, [t.cuda() for t in inputs], {k: v.cuda() for k, v in labels.items()})
    print("autocast loss", float(aloss), "logit err", [rel(a, b) for a, b in zip(apreds, rpreds)])
    cols["autocast"] = agrads
    for mode in ("float32", "bf16_simt", "bf16_tc"):
        cfg = epic_slowfast_cfg((97, 300), freeze)
        cfg.BN.FREEZE = freeze
        cfg.MODEL.DROPOUT_RATE = 0.0
        cfg.B200.COMPUTE_DTYPE = "float32" if mode == "float32" else "bfloat16"
        cfg.NUM_GPUS = 1
        model = build_model(cfg)
        model.load_state_dict(sd)
        model.train()
        if freeze:
            model.freeze_fn("bn_parameters")
            model.freeze_fn("bn_statistics")
        model.engine.force_simt = mode == "bf16_simt"
        preds = model([t.cuda() for t in inputs])
        loss, _, _ = M.compute_loss(preds, {k: v.cuda() for k, v in labels.items()})
        loss.backward()
        torch.cuda.synchronize()
        print(mode, "loss", float(loss), "ref", float(rloss), "logit err", [rel(a, b) for a, b in zip(preds, rpreds)])
        cols[mode] = {k: p.grad.clone() for k, p in model.named_parameters() if p.grad is not None}
    names = list(rgrads.keys())[::-1]
    print("%-50s %s" % ("param", " ".join("%11s" % c for c in cols)))
    for k in names:
        print("%-50s %s" % (k, " ".join("%11.2e" % rel(cols[c][k], rgrads[k]) if k in cols[c] else "       none" for c in cols)))


main()

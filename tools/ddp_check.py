"""GPU aid (2+ GPUs): gradients of asf_b200.parallel.DataParallel after backward == mean over ranks of the gradients
each rank's batch gives on its own, with the two-stream / weight-gradient-stream engine, eager and inside a CUDA graph.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/ddp_check.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "auditory-slow-fast_b200")):
    sys.path.insert(0, p)
import torch
import torch.distributed as dist

from asf_b200.config.defaults import epic_slowfast_cfg
from asf_b200.models import build_model
from asf_b200.train_step import compute_loss

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
B = 4


def batch(r):
    g = torch.Generator().manual_seed(100 + r)
    x = [torch.randn(B, 1, 100, 128, generator=g).cuda(), torch.randn(B, 1, 400, 128, generator=g).cuda()]
    lab = {"verb": torch.randint(0, 97, (B,), generator=g).cuda(), "noun": torch.randint(0, 300, (B,), generator=g).cuda()}
    return x, lab


def grads_of(model, x, lab):
    net = model.module if hasattr(model, "module") else model
    for p in net.parameters():
        p.grad = None
    loss, _, _ = compute_loss(model(x), lab)
    loss.backward()
    torch.cuda.synchronize()
    return {n: p.grad.detach().clone() for n, p in net.named_parameters() if p.grad is not None}


worst = 0.0
for freeze in (True, False):
    cfg = epic_slowfast_cfg((97, 300), freeze)
    cfg.BN.FREEZE = freeze
    cfg.MODEL.DROPOUT_RATE = 0.0
    cfg.NUM_GPUS = 1
    torch.manual_seed(0)
    single = build_model(cfg)        # no wrapper: per-rank gradients
    single.train()
    cfg.NUM_GPUS = world
    torch.manual_seed(0)
    ddp = build_model(cfg)           # wrapped in asf_b200.parallel.DataParallel (models/build.py)
    ddp.train()
    if freeze:
        for m in (single, ddp.module):
            m.freeze_fn("bn_parameters")
            m.freeze_fn("bn_statistics")
    ddp.module.load_state_dict(single.state_dict())
    ref = None
    for r in range(world):
        x, lab = batch(r)
        g = grads_of(single, x, lab)
        # BN running statistics must not drift between the per-rank passes of the reference model
        single.load_state_dict(ddp.module.state_dict())
        ref = g if ref is None else {k: ref[k] + g[k] for k in g}
    ref = {k: v / world for k, v in ref.items()}
    x, lab = batch(rank)
    got = grads_of(ddp, x, lab)
    errs = sorted(((float((got[k] - ref[k]).norm() / (ref[k].norm() + 1e-12)), k) for k in ref), reverse=True)
    worst, med = errs[0][0], errs[len(errs) // 2][0]
    # frozen BN: deterministic up to bf16 rounding.  Training-mode BN at B = 4 from random init is chaotic (fp32 atomics of
    # the batch statistics reorder between runs; the same model twice on the same GPU differs by a few percent on the
    # stems), so there the bound is on the bulk of the tensors.
    if freeze:
        assert worst < 2e-2, errs[:5]
    else:
        assert med < 1e-2 and worst < 0.15, (med, errs[:5])
    if rank == 0:
        print("freeze=%s: %d gradient tensors vs the mean over %d ranks: median rel err %.2e, worst %.2e (%s)"
              % (freeze, len(ref), world, med, worst, errs[0][1]))
dist.barrier()
dist.destroy_process_group()

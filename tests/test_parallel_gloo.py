"""CPU, world_size 2 over gloo: the bucketed gradient all-reduce of asf_b200.parallel (the N > 1 path of the
training step) averages the flat gradient buffer exactly and splits it into the expected buckets; and the flat
buffer's prefix order equals the order in which the fused backward finishes its units."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from asf_b200.parallel import GradBucketer

        n = 1000
        flat = (rank + 1) * torch.arange(n, dtype=torch.float32)
        b = GradBucketer(flat, bucket_bytes=4 * 300)
        b.begin()
        for done in (100, 250, 320, 650, 700, 990):   # prefixes becoming final during "backward"
            b.ready(done)
        b.finish()
        want = torch.arange(n, dtype=torch.float32) * (sum(range(1, world + 1)) / world)
        ok = torch.allclose(flat, want)
        q.put((rank, ok, b.buckets))
    finally:
        dist.destroy_process_group()


def test_bucketed_allreduce_world2():
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    os.environ["PYTHONPATH"] = os.pathsep.join([root, os.path.join(root, "auditory-slow-fast_b200"), os.environ.get("PYTHONPATH", "")])
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, ok, buckets in res:
        assert ok, "rank %d: averaged gradients differ" % rank
        # a bucket is launched as soon as >= 300 elements are final: [0,320) [320,650) [650,990) + the tail
        assert buckets == [(0, 320), (320, 650), (650, 990), (990, 1000)], buckets


def test_flat_gradient_layout_follows_backward_order():
    """reverse registration order == the order in which Engine._backward notifies finished units"""
    from asf_b200.config.defaults import epic_slowfast_cfg
    from asf_b200.models.audio_model_builder import AudioSlowFast

    cfg = epic_slowfast_cfg()
    cfg.NUM_GPUS = 0
    m = AudioSlowFast(cfg)
    names = [n for n, _ in m.named_parameters()][::-1]
    order = ["head"]
    for s, depth in ((5, 3), (4, 6), (3, 4), (2, 3)):
        for p in (1, 0):
            order += ["s%d.pathway%d_res%d" % (s, p, i) for i in reversed(range(depth))]
        order.append("s%d_fuse" % (s - 1))
    order.append("s1")
    pos = 0
    for unit in order:
        pre = unit + "."
        assert names[pos].startswith(pre), (unit, names[pos])
        while pos < len(names) and names[pos].startswith(pre):
            pos += 1
    assert pos == len(names)

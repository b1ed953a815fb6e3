"""CPU: host-side logic of the rows next to the hot path: clip-start arithmetic of the staging layer
against the reference's get_start_end_idx, the checkpoint file format / naming, and the VGG-Sound loader's framing."""
import os
import random

import numpy as np
import pytest
import torch

from asf_b200.datasets import staging
from asf_b200.datasets import utils as U
from asf_b200.utils import checkpoint as CK
from oracle import ref_shim


def test_clip_starts_follow_get_start_end_idx():
    clip = int(round(24000 * 1.999))
    ns = np.array([47976, 60000, 1234567, 30000])
    ss = np.array([0, 100, 5000, 7])
    for view in range(10):
        got = staging.clip_starts(ns, clip, view, 10, ss)
        for i in range(len(ns)):
            s, e = U.get_start_end_idx(int(ns[i]), clip, view, 10, start_sample=int(ss[i]))
            assert got[i] == int(s) and int(e) - int(s) in (clip - 1, clip - 2)
    mv = staging.multi_view_starts(ns, clip, 10, ss)
    assert mv.shape == (4, 10) and (np.diff(mv, axis=1) >= 0).all()
    assert (mv[:, 0] == ss).all() and mv[2, 9] == 5000 + 1234567 - clip
    assert (mv[3] == 7).all()  # shorter than a clip: delta = 0, every view starts at the record's first sample
    random.seed(5)
    a = staging.clip_starts(ns, clip, -1, 10, ss)
    random.seed(5)
    b = [int(U.get_start_end_idx(int(n), clip, -1, 10, start_sample=int(s))[0]) for n, s in zip(ns, ss)]
    assert list(a) == b  # the same draws from Python's RNG, in the same order


@pytest.mark.skipif(not ref_shim.reference_available(), reason="needs /root/reference (build container only)")
def test_clip_starts_match_the_reference_function():
    ref = ref_shim.load_reference()
    clip = 47976
    for n, s0 in ((47976, 0), (100000, 12), (3000000, 99999)):
        for view in range(10):
            want = int(ref.get_start_end_idx(n, clip, view, 10, start_sample=s0)[0])
            assert staging.clip_starts([n], clip, view, 10, [s0])[0] == want


def test_checkpoint_naming_and_roundtrip(tmp_path):
    job = str(tmp_path)
    assert CK.get_path_to_checkpoint(job, 7).endswith(os.path.join("checkpoints", "checkpoint_epoch_00007.pyth"))
    assert CK.get_path_to_checkpoint(job, 7, True).endswith("checkpoint_best.pyth")
    assert not CK.has_checkpoint(job)

    class Cfg:
        class SOLVER:
            MAX_EPOCH = 30

        class TRAIN:
            CHECKPOINT_PERIOD = 5

        def dump(self):
            return "cfg-dump"

    assert CK.is_checkpoint_epoch(Cfg, 4) and not CK.is_checkpoint_epoch(Cfg, 5) and CK.is_checkpoint_epoch(Cfg, 29)
    net = torch.nn.Sequential(torch.nn.Conv2d(1, 4, 3), torch.nn.BatchNorm2d(4))
    opt = torch.optim.SGD(net.parameters(), lr=0.1, momentum=0.9)
    net(torch.randn(2, 1, 8, 8)).sum().backward()
    opt.step()
    path = CK.save_checkpoint(job, net, opt, 2, Cfg())
    assert path == CK.get_path_to_checkpoint(job, 3) and CK.has_checkpoint(job) and CK.get_last_checkpoint(job) == path
    ck = torch.load(path, map_location="cpu")
    assert set(ck) == {"epoch", "model_state", "optimizer_state", "cfg"} and ck["epoch"] == 2 and ck["cfg"] == "cfg-dump"
    net2 = torch.nn.Sequential(torch.nn.Conv2d(1, 4, 3), torch.nn.BatchNorm2d(4), torch.nn.Conv2d(4, 2, 1))
    opt2 = torch.optim.SGD(net.parameters(), lr=0.1, momentum=0.9)
    assert CK.load_checkpoint(path, net2, data_parallel=False, optimizer=None) == 2
    assert torch.equal(net2[0].weight, net[0].weight) and torch.equal(net2[1].running_mean, net[1].running_mean)
    assert sorted(CK.load_checkpoint.not_loaded) == ["2.bias", "2.weight"]  # name + shape matching, the rest untouched
    assert CK.load_checkpoint(path, net2, data_parallel=False, epoch_reset=True) == -1
    # exclusion and renaming (fine-tuning from a checkpoint with another head / a wrapper prefix)
    before = net2[0].weight.clone()
    with torch.no_grad():
        net2[0].weight.add_(1.0)
    CK.load_checkpoint(path, net2, data_parallel=False, exclude_layers=("0.",))
    assert torch.equal(net2[0].weight, before + 1.0)

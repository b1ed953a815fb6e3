"""CPU: the drop-in boundary -- package import, config tree, registry semantics, state_dict layout / init
parity with the reference (through the oracle), pack_pathway_output, and the C-ABI library exports."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

from oracle import logmel_np as OL
from oracle import model_torch as M

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_cfg_defaults_and_merge():
    from asf_b200.config.defaults import epic_slowfast_cfg, get_cfg

    cfg = get_cfg()
    assert cfg.AUDIO_DATA.SAMPLING_RATE == 24000 and cfg.AUDIO_DATA.N_FFT == 2048
    assert cfg.MODEL.MODEL_NAME == "SlowFast" and cfg.SLOWFAST.ALPHA == 8  # reference defaults
    e = epic_slowfast_cfg()
    assert (e.SLOWFAST.ALPHA, e.SLOWFAST.BETA_INV, e.SLOWFAST.FUSION_KERNEL_SZ) == (4, 8, 7)
    assert e.MODEL.NUM_CLASSES == [97, 300] and e.AUDIO_DATA.NUM_FRAMES == 400 and e.BN.FREEZE is True
    e.merge_from_list(["MODEL.DROPOUT_RATE", 0.0])
    assert e.MODEL.DROPOUT_RATE == 0.0
    assert int(round(e.AUDIO_DATA.SAMPLING_RATE * e.AUDIO_DATA.CLIP_SECS)) == 47976
    # the reference converts every BatchNorm to SyncBatchNorm when NUM_GPUS > 1 (models/build.py:45-46): same default
    # here, as a switch of the data-parallel wrapper rather than a module swap
    import inspect

    from asf_b200.parallel import DataParallel
    assert e.B200.SYNC_BN is True
    assert inspect.signature(DataParallel.__init__).parameters["sync_bn"].default is True


def test_registry_and_build_model_contract():
    from asf_b200.config.defaults import epic_slowfast_cfg
    from asf_b200.models import MODEL_REGISTRY, build_model

    assert "AudioSlowFast" in MODEL_REGISTRY and "ResNet" in MODEL_REGISTRY
    cfg = epic_slowfast_cfg()
    cfg.MODEL.MODEL_NAME = "SlowFast"  # the reference's (unregistered) default name -> KeyError
    cfg.NUM_GPUS = 0
    with pytest.raises(KeyError):
        build_model(cfg)
    if not torch.cuda.is_available():
        cfg.MODEL.MODEL_NAME = "AudioSlowFast"
        cfg.NUM_GPUS = 1
        with pytest.raises(AssertionError):
            build_model(cfg)


@pytest.mark.parametrize("arch,nc", [("slowfast", (97, 300)), ("slow", (97, 300)), ("slowfast", (309,))])
def test_state_dict_layout_and_init_match_reference(arch, nc):
    from asf_b200.config.defaults import epic_slowfast_cfg, slow_r50_cfg
    from asf_b200.models import build_model

    cfg = epic_slowfast_cfg(nc) if arch == "slowfast" else slow_r50_cfg(nc)
    cfg.NUM_GPUS = 0
    torch.manual_seed(0)
    model = build_model(cfg)
    sd = model.state_dict()
    ref = M.init_state_dict(M.epic_cfg(arch=arch, num_classes=nc), 0)  # bit-exact restatement of the reference init
    assert list(sd.keys()) == list(ref.keys())
    for k in sd:
        assert sd[k].shape == ref[k].shape and sd[k].dtype == ref[k].dtype and torch.equal(sd[k], ref[k]), k
    # optimizer grouping contract (models/optimizer.py:31-46): BN params are those with "bn" in the name
    n_bn = sum(1 for n, _ in model.named_parameters() if "bn" in n)
    assert n_bn == 2 * sum(1 for m in model.modules() if isinstance(m, torch.nn.BatchNorm2d))


def test_freeze_fn_semantics():
    from asf_b200.config.defaults import epic_slowfast_cfg
    from asf_b200.models import build_model

    cfg = epic_slowfast_cfg()
    cfg.NUM_GPUS = 0
    model = build_model(cfg)
    model.train()
    model.freeze_fn("bn_parameters")
    model.freeze_fn("bn_statistics")
    bns = {n: m for n, m in model.named_modules() if isinstance(m, torch.nn.BatchNorm2d)}
    assert len(bns) == 110
    unfrozen = [n for n, m in bns.items() if m.training]
    assert sorted(unfrozen) == ["s1.pathway0_stem.bn", "s1.pathway1_stem.bn", "s1_fuse.bn"]
    assert sum(p.requires_grad for p in model.parameters()) == 120  # 120 tensors with BN frozen
    model.train()
    assert all(m.training for m in bns.values())  # model.train() re-enables, as in the reference


def test_forward_without_cuda_fails_loudly():
    from asf_b200.config.defaults import epic_slowfast_cfg
    from asf_b200.models import build_model

    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    cfg = epic_slowfast_cfg()
    cfg.NUM_GPUS = 0
    model = build_model(cfg).eval()
    with pytest.raises(RuntimeError):
        model([torch.zeros(1, 1, 100, 128), torch.zeros(1, 1, 400, 128)])
    with pytest.raises(AssertionError):
        model([torch.zeros(1, 1, 400, 128)])  # wrong pathway count asserts first, like the reference


def test_pack_pathway_output(golden_dir):
    from asf_b200.config.defaults import epic_slowfast_cfg, slow_r50_cfg
    from asf_b200.datasets.utils import get_start_end_idx, pack_pathway_output

    g = np.load(os.path.join(golden_dir, "pack_pathway.npz"))
    cfg = epic_slowfast_cfg()
    spec = torch.arange(400, dtype=torch.float32).view(1, 400, 1).repeat(1, 1, 4)
    slow, fast = pack_pathway_output(cfg, spec)
    assert fast is spec and slow.shape == (1, 100, 4)
    assert np.array_equal(slow[0, :, 0].long().numpy(), g["T400_a4"])
    assert len(pack_pathway_output(slow_r50_cfg(), spec)) == 1
    cfg.MODEL.ARCH = "bogus"
    with pytest.raises(NotImplementedError):
        pack_pathway_output(cfg, spec)
    s, e = get_start_end_idx(100000, 47976, 3, 10, start_sample=5)
    assert int(e) - int(s) == 47975 and int(s) == 5 + int(np.linspace(0, 100000 - 47976, 10)[3])


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "asf_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(asf_[a-z0-9_]+)\s*\(", src)))


def test_c_abi_library_exports_every_declared_symbol():
    from asf_b200 import _lib

    assert os.path.exists(_lib.LIB_PATH), "libasf_b200.so missing: run auditory-slow-fast_b200/build.sh"
    lib = ctypes.CDLL(_lib.LIB_PATH)  # loads without a GPU (static cudart, driver entry points resolved lazily)
    names = _declared_symbols()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), "symbol %s declared in include/asf_b200.h but not exported" % n
    # and every bound signature exists in the header
    for n in _lib.SIGNATURES:
        assert n in names, n
    assert lib.asf_version() >= 100


def test_no_oracle_import_in_product_path():
    pkg = os.path.join(ROOT, "auditory-slow-fast_b200", "asf_b200")
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                txt = open(os.path.join(d, f)).read()
                assert "import oracle" not in txt and "from oracle" not in txt, os.path.join(d, f)

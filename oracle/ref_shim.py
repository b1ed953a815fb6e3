"""TEST INFRASTRUCTURE ONLY -- never imported by the product path.

Import shim that lets the *unmodified* reference model code under
``/root/reference/audio_slowfast`` run in this container, where ``fvcore`` is not
installed and ``import audio_slowfast`` itself fails (its ``__init__`` drags in
wandb / h5py / matplotlib through ``tools``).

Recipe:
  * register an empty package object called ``audio_slowfast`` whose ``__path__``
    points at the reference tree, so sub-modules import normally but the
    package ``__init__`` is never executed;
  * provide a ~60 line stand-in for the three ``fvcore`` symbols the model path
    touches (``Registry``, ``CfgNode``, ``c2_msra_fill``).

The arithmetic that then executes is 100 % the reference's own files
(audio_model_builder.py, stem_helper.py, resnet_helper.py, head_helper.py,
batchnorm_helper.py, weight_init_helper.py) on stock torch.

``/root/reference`` does not exist on the GPU box: only
``tests/golden/make_golden.py`` and the ``-m "not gpu"`` cross-check tests (which
skip when the tree is absent) may call :func:`load_reference`.
"""
from __future__ import annotations

import copy
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("ASF_REFERENCE_ROOT", "/root/reference")


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "audio_slowfast", "models"))


# --------------------------------------------------------------------------- fvcore stand-in
class _Registry:
    """fvcore.common.registry.Registry: name -> object, ``register()`` decorator, ``get``."""

    def __init__(self, name):
        self._name = name
        self._obj_map = {}

    def register(self, obj=None):
        if obj is None:
            def deco(o):
                self._obj_map[o.__name__] = o
                return o
            return deco
        self._obj_map[obj.__name__] = obj
        return obj

    def get(self, name):
        if name not in self._obj_map:
            raise KeyError("No object named '{}' found in '{}' registry!".format(name, self._name))
        return self._obj_map[name]

    def __repr__(self):
        return "Registry({}: {})".format(self._name, sorted(self._obj_map))


class _CfgNode(dict):
    """Minimal yacs/fvcore CfgNode: attribute access, clone, merge_from_file, merge_from_list."""

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        self[k] = v

    def clone(self):
        return copy.deepcopy(self)

    def _merge(self, other):
        for k, v in other.items():
            if isinstance(v, dict) and isinstance(self.get(k), dict):
                self[k]._merge(v)
            elif isinstance(v, dict):
                node = _CfgNode()
                node._merge(v)
                self[k] = node
            else:
                self[k] = v

    def merge_from_file(self, path):
        import yaml

        with open(path) as f:
            self._merge(yaml.safe_load(f) or {})

    def merge_from_list(self, lst):
        for k, v in zip(lst[0::2], lst[1::2]):
            node = self
            parts = k.split(".")
            for p in parts[:-1]:
                node = node[p]
            node[parts[-1]] = v

    def dump(self):
        import yaml

        return yaml.safe_dump(_plain(self))


def _plain(n):
    return {k: _plain(v) if isinstance(v, dict) else v for k, v in n.items()}


def _c2_msra_fill(module):
    import torch.nn as nn

    nn.init.kaiming_normal_(module.weight, mode="fan_out", nonlinearity="relu")
    if module.bias is not None:
        nn.init.constant_(module.bias, 0)


def _install_fvcore_stub():
    if "fvcore" in sys.modules:
        return

    def mod(name):
        m = types.ModuleType(name)
        m.__path__ = []
        sys.modules[name] = m
        return m

    fv = mod("fvcore")
    common = mod("fvcore.common")
    registry = mod("fvcore.common.registry")
    config = mod("fvcore.common.config")
    nn_ = mod("fvcore.nn")
    winit = mod("fvcore.nn.weight_init")
    registry.Registry = _Registry
    config.CfgNode = _CfgNode
    winit.c2_msra_fill = _c2_msra_fill
    fv.common, fv.nn = common, nn_
    common.registry, common.config = registry, config
    nn_.weight_init = winit


def _install_loguru_stub():
    try:
        import loguru  # noqa: F401
    except Exception:
        m = types.ModuleType("loguru")

        class _L:
            def __getattr__(self, k):
                return lambda *a, **kw: None

        m.logger = _L()
        sys.modules["loguru"] = m


_REF = None


def load_reference():
    """Returns a namespace with the reference's get_cfg, build_model, MODEL_REGISTRY,
    pack_pathway_output, get_start_end_idx and the config directory."""
    global _REF
    if _REF is not None:
        return _REF
    if not reference_available():
        raise RuntimeError("reference tree not present at " + REFERENCE_ROOT)
    _install_fvcore_stub()
    _install_loguru_stub()
    if "audio_slowfast" not in sys.modules:
        pkg = types.ModuleType("audio_slowfast")
        pkg.__path__ = [os.path.join(REFERENCE_ROOT, "audio_slowfast")]
        sys.modules["audio_slowfast"] = pkg
        for sub in ("datasets", "utils", "config"):
            m = types.ModuleType("audio_slowfast." + sub)
            m.__path__ = [os.path.join(REFERENCE_ROOT, "audio_slowfast", sub)]
            sys.modules["audio_slowfast." + sub] = m
            setattr(pkg, sub, m)
    from audio_slowfast.config.defaults import get_cfg  # type: ignore
    import audio_slowfast.models as models  # type: ignore  # registers AudioSlowFast / ResNet
    from audio_slowfast.models.build import MODEL_REGISTRY, build_model  # type: ignore
    from audio_slowfast.datasets.utils import get_start_end_idx, pack_pathway_output  # type: ignore

    ns = types.SimpleNamespace(
        get_cfg=get_cfg,
        build_model=build_model,
        MODEL_REGISTRY=MODEL_REGISTRY,
        models=models,
        pack_pathway_output=pack_pathway_output,
        get_start_end_idx=get_start_end_idx,
        config_dir=os.path.join(REFERENCE_ROOT, "models", "asf", "config"),
    )
    _REF = ns
    return ns


def load_spec_augment():
    """The reference's own datasets/spec_augment.py module (spec_augment, time_warp, freq_mask, time_mask; it imports
    datasets/sparse_image_warp.py), unmodified."""
    load_reference()
    import audio_slowfast.datasets.spec_augment as sa  # type: ignore

    return sa


def reference_cfg(arch="slowfast", num_classes=(97, 300), bn_freeze=None):
    """cfg of the shipped EPIC recipe (models/asf/config/asf-original-augment.yaml) on CPU,
    optionally switched to a single pathway (BASELINE config 1 = SLOW_R50)."""
    ref = load_reference()
    cfg = ref.get_cfg()
    cfg.merge_from_file(os.path.join(ref.config_dir, "asf-original-augment.yaml"))
    cfg.NUM_GPUS = 0
    cfg.MODEL.NUM_CLASSES = list(num_classes)
    if arch != "slowfast":
        cfg.MODEL.ARCH = arch
        cfg.MODEL.MODEL_NAME = "ResNet"
        cfg.DATA.INPUT_CHANNEL_NUM = [1]
        cfg.RESNET.NUM_BLOCK_TEMP_KERNEL = [[3], [4], [6], [3]]
        cfg.RESNET.FREQUENCY_STRIDES = [[1], [2], [2], [2]]
        cfg.RESNET.FREQUENCY_DILATIONS = [[1], [1], [1], [1]]
    if bn_freeze is not None:
        cfg.BN.FREEZE = bn_freeze
    return cfg

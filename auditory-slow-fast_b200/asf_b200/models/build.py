"""Model construction (mirror of /root/reference/audio_slowfast/models/build.py:10-58)."""
import torch


class Registry:
    """name -> class registry with the fvcore semantics the reference relies on
    (``@MODEL_REGISTRY.register()``, ``get(name)`` raising KeyError for unknown names)."""

    def __init__(self, name):
        self._name = name
        self._obj_map = {}

    def register(self, obj=None):
        if obj is None:
            def deco(o):
                self._do_register(o.__name__, o)
                return o
            return deco
        self._do_register(obj.__name__, obj)
        return obj

    def _do_register(self, name, obj):
        assert name not in self._obj_map, "An object named '{}' was already registered in '{}' registry!".format(
            name, self._name)
        self._obj_map[name] = obj

    def get(self, name):
        ret = self._obj_map.get(name)
        if ret is None:
            raise KeyError("No object named '{}' found in '{}' registry!".format(name, self._name))
        return ret

    def __contains__(self, name):
        return name in self._obj_map

    def __repr__(self):
        return "Registry({}: {})".format(self._name, sorted(self._obj_map))


MODEL_REGISTRY = Registry("MODEL")
MODEL_REGISTRY.__doc__ = """Registry for audio models: the registered object is called as ``obj(cfg)``."""


def build_model(cfg, gpu_id=None):
    """Builds the audio model named by ``cfg.MODEL.MODEL_NAME`` on the current (or given) GPU.

    Same contract as the reference: asserts on NUM_GPUS vs. the CUDA devices present, registry lookup
    (KeyError if unknown), device placement, and -- when NUM_GPUS > 1 -- a data-parallel wrapper whose
    ``.module`` is the model (callers reach through it: train_net.py:392, :710; checkpoint.py:111).
    Differences, both deliberate: the wrapper is asf_b200.parallel.DataParallel (bucketed NCCL all-reduce
    issued from inside the fused backward, overlapping the remaining wgrads) instead of torch DDP, and the
    SyncBatchNorm conversion (build.py:45-46) is a switch of the engine (global-batch statistics by all-reduce of
    the per-channel sums, cfg.B200.SYNC_BN, default True) rather than a module swap, so module names stay put.
    """
    if torch.cuda.is_available():
        assert cfg.NUM_GPUS <= torch.cuda.device_count(), "Cannot use more GPU devices than available"
    else:
        assert cfg.NUM_GPUS == 0, "Cuda is not available. Please set `NUM_GPUS: 0 for running on CPUs."
    model = MODEL_REGISTRY.get(cfg.MODEL.MODEL_NAME)(cfg)
    if cfg.NUM_GPUS:
        cur_device = torch.cuda.current_device() if gpu_id is None else gpu_id
        model = model.cuda(device=cur_device)
    if cfg.NUM_GPUS > 1:
        from ..parallel import DataParallel

        model = DataParallel(model, sync_bn=bool(getattr(getattr(cfg, "B200", None), "SYNC_BN", True)))
    return model

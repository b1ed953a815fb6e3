"""Configuration tree for the hot path.

Mirrors the fields that the reference reads on the path
(/root/reference/audio_slowfast/config/defaults.py:17-39, :114-145, :151-194, :200-215, :221-261, :324),
with the same names and default values, so a reference yaml (e.g. models/asf/config/asf-original-augment.yaml)
merges unchanged.  ``CfgNode`` is a small attribute-dict (the reference uses fvcore/yacs, which is not a
dependency here); the model builders only need attribute access, so a real fvcore CfgNode works as well.
"""
from __future__ import annotations

import copy


class CfgNode(dict):
    def __init__(self, init=None):
        super().__init__()
        for k, v in (init or {}).items():
            self[k] = CfgNode(v) if isinstance(v, dict) and not isinstance(v, CfgNode) else v

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        self[k] = v

    def clone(self):
        return copy.deepcopy(self)

    def merge_from_dict(self, other):
        for k, v in other.items():
            if isinstance(v, dict):
                if not isinstance(self.get(k), dict):
                    self[k] = CfgNode()
                self[k].merge_from_dict(v)
            else:
                self[k] = v

    def merge_from_file(self, path):
        import yaml

        with open(path) as f:
            self.merge_from_dict(yaml.safe_load(f) or {})

    def merge_from_list(self, lst):
        assert len(lst) % 2 == 0, "Override list has odd length: {}".format(lst)
        for k, v in zip(lst[0::2], lst[1::2]):
            node = self
            parts = k.split(".")
            for p in parts[:-1]:
                node = node[p]
            node[parts[-1]] = v

    def dump(self):
        import yaml

        def plain(n):
            return {k: plain(v) if isinstance(v, dict) else v for k, v in n.items()}

        return yaml.safe_dump(plain(self))


_C = CfgNode()

_C.BN = CfgNode()
_C.BN.FREEZE = False
_C.BN.USE_PRECISE_STATS = False
_C.BN.NUM_BATCHES_PRECISE = 200
_C.BN.WEIGHT_DECAY = 0.0
_C.BN.NORM_TYPE = "batchnorm"
_C.BN.NUM_SPLITS = 1
_C.BN.NUM_SYNC_DEVICES = 1

_C.TRAIN = CfgNode()
_C.TRAIN.ENABLE = True
_C.TRAIN.DATASET = "vggsound"
_C.TRAIN.BATCH_SIZE = 64
_C.TRAIN.EVAL_PERIOD = 10
_C.TRAIN.CHECKPOINT_PERIOD = 10
_C.TRAIN.AUTO_RESUME = True
_C.TRAIN.CHECKPOINT_FILE_PATH = ""
_C.TRAIN.CHECKPOINT_EPOCH_RESET = False
_C.TRAIN.CHECKPOINT_CLEAR_NAME_PATTERN = ()

_C.TEST = CfgNode()
_C.TEST.ENABLE = True
_C.TEST.DATASET = "vggsound"
_C.TEST.BATCH_SIZE = 8
_C.TEST.CHECKPOINT_FILE_PATH = ""
_C.TEST.NUM_ENSEMBLE_VIEWS = 10

_C.RESNET = CfgNode()
_C.RESNET.TRANS_FUNC = "bottleneck_transform"
_C.RESNET.NUM_GROUPS = 1
_C.RESNET.WIDTH_PER_GROUP = 64
_C.RESNET.INPLACE_RELU = True
_C.RESNET.STRIDE_1X1 = False
_C.RESNET.ZERO_INIT_FINAL_BN = False
_C.RESNET.DEPTH = 50
_C.RESNET.NUM_BLOCK_TEMP_KERNEL = [[3], [4], [6], [3]]
_C.RESNET.FREQUENCY_STRIDES = [[1], [2], [2], [2]]
_C.RESNET.FREQUENCY_DILATIONS = [[1], [1], [1], [1]]

_C.MODEL = CfgNode()
_C.MODEL.ARCH = "slowfast"
_C.MODEL.MODEL_NAME = "SlowFast"
_C.MODEL.NUM_CLASSES = [400]
_C.MODEL.LOSS_FUNC = "cross_entropy"
_C.MODEL.SINGLE_PATHWAY_ARCH = ["slow", "fast"]
_C.MODEL.MULTI_PATHWAY_ARCH = ["slowfast"]
_C.MODEL.DROPOUT_RATE = 0.5
_C.MODEL.FC_INIT_STD = 0.01
_C.MODEL.HEAD_ACT = "softmax"

_C.SLOWFAST = CfgNode()
_C.SLOWFAST.BETA_INV = 8
_C.SLOWFAST.ALPHA = 8
_C.SLOWFAST.FUSION_CONV_CHANNEL_RATIO = 2
_C.SLOWFAST.FUSION_KERNEL_SZ = 5

_C.DATA = CfgNode()
_C.DATA.INPUT_CHANNEL_NUM = [1, 1]

_C.AUDIO_DATA = CfgNode()
_C.AUDIO_DATA.SAMPLING_RATE = 24000
_C.AUDIO_DATA.N_FFT = 2048
_C.AUDIO_DATA.CLIP_SECS = 1.279
_C.AUDIO_DATA.WINDOW_LENGTH = 10.0
_C.AUDIO_DATA.HOP_LENGTH = 5.0
_C.AUDIO_DATA.NUM_FRAMES = 256
_C.AUDIO_DATA.NUM_FREQUENCIES = 128

_C.SOLVER = CfgNode()
_C.SOLVER.BASE_LR = 0.1
_C.SOLVER.MOMENTUM = 0.9
_C.SOLVER.DAMPENING = 0.0
_C.SOLVER.NESTEROV = True
_C.SOLVER.WEIGHT_DECAY = 1e-4
_C.SOLVER.OPTIMIZING_METHOD = "sgd"

_C.NUM_GPUS = 1
_C.NUM_SHARDS = 1
_C.SHARD_ID = 0
_C.RNG_SEED = 1
_C.DIST_BACKEND = "nccl"

# --- additive (not in the reference): knobs of the B200 engine -------------------------------------------
_C.B200 = CfgNode()
_C.B200.COMPUTE_DTYPE = "bfloat16"   # "bfloat16" (tcgen05 path) or "float32" (CUDA-core parity mode)
# NUM_GPUS > 1: train-mode BatchNorm uses the statistics of the global batch, as the torch.nn.SyncBatchNorm the
# reference converts every BatchNorm to (models/build.py:45-46).  False = per-replica statistics (no BN collectives).
_C.B200.SYNC_BN = True


def _assert_and_infer_cfg(cfg):
    # defaults.py:495-516
    if cfg.BN.USE_PRECISE_STATS:
        assert cfg.BN.NUM_BATCHES_PRECISE >= 0
    assert cfg.TRAIN.BATCH_SIZE % max(cfg.NUM_GPUS, 1) == 0
    assert cfg.TEST.BATCH_SIZE % max(cfg.NUM_GPUS, 1) == 0
    assert cfg.RESNET.NUM_GROUPS > 0
    assert cfg.RESNET.WIDTH_PER_GROUP > 0
    assert cfg.RESNET.WIDTH_PER_GROUP % cfg.RESNET.NUM_GROUPS == 0
    assert cfg.SHARD_ID < cfg.NUM_SHARDS
    return cfg


def get_cfg():
    """Get a copy of the default config (defaults.py:519-523)."""
    return _assert_and_infer_cfg(_C.clone())


def epic_slowfast_cfg(num_classes=(97, 300), bn_freeze=True):
    """The shipped EPIC-KITCHENS recipe (models/asf/config/asf-original-augment.yaml:14-58) as a cfg."""
    cfg = get_cfg()
    cfg.merge_from_dict({
        "TRAIN": {"DATASET": "EpicKitchens", "BATCH_SIZE": 128},
        "DATA": {"INPUT_CHANNEL_NUM": [1, 1]},
        "AUDIO_DATA": {"CLIP_SECS": 1.999, "NUM_FRAMES": 400},
        "SLOWFAST": {"ALPHA": 4, "BETA_INV": 8, "FUSION_CONV_CHANNEL_RATIO": 2, "FUSION_KERNEL_SZ": 7},
        "RESNET": {"ZERO_INIT_FINAL_BN": True, "WIDTH_PER_GROUP": 64, "NUM_GROUPS": 1, "DEPTH": 50,
                   "TRANS_FUNC": "bottleneck_transform", "STRIDE_1X1": False,
                   "NUM_BLOCK_TEMP_KERNEL": [[3, 3], [4, 4], [6, 6], [3, 3]],
                   "FREQUENCY_STRIDES": [[1, 1], [2, 2], [2, 2], [2, 2]],
                   "FREQUENCY_DILATIONS": [[1, 1], [1, 1], [1, 1], [1, 1]]},
        "BN": {"USE_PRECISE_STATS": True, "FREEZE": bool(bn_freeze), "NUM_BATCHES_PRECISE": 200},
        "SOLVER": {"BASE_LR": 0.001, "MOMENTUM": 0.9, "WEIGHT_DECAY": 1e-4, "OPTIMIZING_METHOD": "sgd"},
        "MODEL": {"NUM_CLASSES": list(num_classes), "ARCH": "slowfast", "MODEL_NAME": "AudioSlowFast",
                  "LOSS_FUNC": "cross_entropy", "DROPOUT_RATE": 0.5},
    })
    return cfg


def slow_r50_cfg(num_classes=(97, 300)):
    """Single-pathway SLOW_R50 (BASELINE config 1): ResNet / ARCH slow on the full 400-frame spectrogram."""
    cfg = epic_slowfast_cfg(num_classes)
    cfg.merge_from_dict({
        "MODEL": {"ARCH": "slow", "MODEL_NAME": "ResNet"},
        "DATA": {"INPUT_CHANNEL_NUM": [1]},
        "RESNET": {"NUM_BLOCK_TEMP_KERNEL": [[3], [4], [6], [3]], "FREQUENCY_STRIDES": [[1], [2], [2], [2]],
                   "FREQUENCY_DILATIONS": [[1], [1], [1], [1]]},
    })
    return cfg

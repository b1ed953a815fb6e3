"""Audio SlowFast / ResNet model builders on the B200 engine.

Mirror of /root/reference/audio_slowfast/models/audio_model_builder.py (AudioSlowFast :115-345, ResNet
:624-796, FuseFastToSlow :54-111): same constructor (``Model(cfg)``), same sub-module names and parameter
shapes (=> same 664-entry state_dict, same RNG consumption at construction), same
``forward(x: list[Tensor], bboxes=None)`` contract -- raw logits in training, softmax probabilities in
eval, a ``(verb, noun)`` tuple for two heads -- and same ``freeze_fn`` modes.  The difference is what runs:
``forward`` hands the whole network to asf_b200.engine, which executes it with hand-written sm_100a kernels
(and registers ONE autograd node whose backward runs the fused backward pass).
"""
import torch
import torch.nn as nn

from ..utils import weight_init_helper as init_helper
from . import head_helper, resnet_helper, stem_helper
from .batchnorm_helper import get_norm
from .build import MODEL_REGISTRY

# Number of residual blocks per stage for each supported depth.
_MODEL_STAGE_DEPTH = {50: (3, 4, 6, 3), 101: (3, 4, 23, 3)}

# Temporal kernel size of conv1 and res2..res5, per pathway.
_TEMPORAL_KERNEL_BASIS = {
    "slow": [[[1]], [[1]], [[1]], [[3]], [[3]]],
    "fast": [[[5]], [[3]], [[3]], [[3]], [[3]]],
    "slowfast": [[[1], [5]], [[1], [3]], [[1], [3]], [[3], [3]], [[3], [3]]],
}

_POOL1 = {"slow": [[1, 1]], "fast": [[1, 1]], "slowfast": [[1, 1], [1, 1]]}

_UNFROZEN_BN = ("s1.pathway0_stem.bn", "s1.pathway1_stem.bn", "s1_fuse.bn")


class FuseFastToSlow(stem_helper._EngineOnly):
    """Lateral connection: time-strided conv (k x 1, stride alpha) on the Fast pathway -> BN -> ReLU ->
    channel-concatenated onto the Slow pathway.  In the engine the conv writes straight into the channel
    slice of a pre-allocated Slow buffer, so no ``torch.cat`` copy exists."""

    def __init__(self, dim_in, fusion_conv_channel_ratio, fusion_kernel, alpha, eps=1e-5, bn_mmt=0.1,
                 inplace_relu=True, norm_module=nn.BatchNorm2d):
        super().__init__()
        self.conv_f2s = nn.Conv2d(dim_in, dim_in * fusion_conv_channel_ratio, kernel_size=[fusion_kernel, 1],
                                  stride=[alpha, 1], padding=[fusion_kernel // 2, 0], bias=False)
        self.bn = norm_module(num_features=dim_in * fusion_conv_channel_ratio, eps=eps, momentum=bn_mmt)
        self.relu = nn.ReLU(inplace_relu)


class _EngineModel(nn.Module):
    """Shared behaviour of the two registered models."""

    num_pathways = 1

    def _finish(self, cfg):
        init_helper.init_weights(self, cfg.MODEL.FC_INIT_STD, cfg.RESNET.ZERO_INIT_FINAL_BN)
        self._engine = None
        dt = getattr(getattr(cfg, "B200", None), "COMPUTE_DTYPE", "bfloat16")
        self.compute_dtype = {"bfloat16": torch.bfloat16, "float32": torch.float32}[dt]
        self._cfg_arch = cfg.MODEL.ARCH
        self._alpha = cfg.SLOWFAST.ALPHA if self.num_pathways == 2 else 1

    @property
    def engine(self):
        if self._engine is None:
            from ..engine import Engine

            self._engine = Engine(self)
        return self._engine

    def forward(self, x, bboxes=None):
        assert len(x) == self.num_pathways, "Input tensor does not contain {} pathway".format(self.num_pathways)
        return self.engine.run(x)

    def load_state_dict(self, *a, **k):
        r = super().load_state_dict(*a, **k)
        if self._engine is not None:
            self._engine.invalidate_weights()
        return r

    def freeze_fn(self, freeze_mode):
        """audio_model_builder.py:327-344: 'bn_parameters' stops gamma/beta updates and 'bn_statistics' puts
        BN layers in eval mode -- for every BN except the two stems' and the first lateral's."""
        for n, m in self.named_modules():
            if not isinstance(m, (nn.BatchNorm2d, nn.SyncBatchNorm)) or any(u in n for u in _UNFROZEN_BN):
                continue
            if freeze_mode == "bn_parameters":
                m.weight.requires_grad_(False)
                m.bias.requires_grad_(False)
            elif freeze_mode == "bn_statistics":
                m.eval()


def _stage_kwargs(cfg, si, npath):
    return dict(
        stride=cfg.RESNET.FREQUENCY_STRIDES[si],
        num_groups=[cfg.RESNET.NUM_GROUPS] * npath,
        num_block_temp_kernel=cfg.RESNET.NUM_BLOCK_TEMP_KERNEL[si],
        trans_func_name=cfg.RESNET.TRANS_FUNC,
        dilation=cfg.RESNET.FREQUENCY_DILATIONS[si],
    )


@MODEL_REGISTRY.register()
class AudioSlowFast(_EngineModel):
    """Two-pathway Auditory SlowFast network (Kazakos et al., ICASSP 2021) for cfg.MODEL.ARCH 'slowfast'."""

    def __init__(self, cfg):
        super().__init__()
        self.norm_module = get_norm(cfg)
        self.num_pathways = 2
        self._construct_network(cfg)
        self._finish(cfg)

    def _construct_network(self, cfg):
        assert cfg.MODEL.ARCH in _POOL1.keys()
        pool_size = _POOL1[cfg.MODEL.ARCH]
        assert len({len(pool_size), self.num_pathways}) == 1
        assert cfg.RESNET.DEPTH in _MODEL_STAGE_DEPTH.keys()
        depths = _MODEL_STAGE_DEPTH[cfg.RESNET.DEPTH]
        groups, wpg = cfg.RESNET.NUM_GROUPS, cfg.RESNET.WIDTH_PER_GROUP
        inner = groups * wpg
        binv, ratio = cfg.SLOWFAST.BETA_INV, cfg.SLOWFAST.FUSION_CONV_CHANNEL_RATIO
        out_dim_ratio = binv // ratio
        tk = _TEMPORAL_KERNEL_BASIS[cfg.MODEL.ARCH]
        norm = self.norm_module

        def fuse(width):
            return FuseFastToSlow(width // binv, ratio, cfg.SLOWFAST.FUSION_KERNEL_SZ, cfg.SLOWFAST.ALPHA,
                                  norm_module=norm)

        self.s1 = stem_helper.AudioModelStem(
            dim_in=cfg.DATA.INPUT_CHANNEL_NUM, dim_out=[wpg, wpg // binv],
            kernel=[tk[0][0] + [7], tk[0][1] + [7]], stride=[[2, 2]] * 2,
            padding=[[tk[0][0][0] // 2, 3], [tk[0][1][0] // 2, 3]], norm_module=norm)
        self.s1_fuse = fuse(wpg)
        w_in = wpg  # slow width entering the stage (before the lateral concat)
        for si, name in enumerate(("s2", "s3", "s4", "s5")):
            w_out = wpg * 4 * (2 ** si)
            stage = resnet_helper.ResStage(
                dim_in=[w_in + w_in // out_dim_ratio, w_in // binv],
                dim_out=[w_out, w_out // binv],
                dim_inner=[inner * (2 ** si), inner * (2 ** si) // binv],
                temp_kernel_sizes=tk[si + 1], num_blocks=[depths[si]] * 2, norm_module=norm,
                **_stage_kwargs(cfg, si, 2))
            setattr(self, name, stage)
            if si < 3:
                setattr(self, name + "_fuse", fuse(w_out))
            if si == 0:
                for p in range(self.num_pathways):
                    self.add_module("pathway{}_pool".format(p),
                                    nn.MaxPool2d(kernel_size=pool_size[p], stride=pool_size[p], padding=[0, 0]))
            w_in = w_out
        nf, nq = cfg.AUDIO_DATA.NUM_FRAMES, cfg.AUDIO_DATA.NUM_FREQUENCIES
        self.head = head_helper.ResNetBasicHead(
            dim_in=[wpg * 32, wpg * 32 // binv],
            num_classes=cfg.MODEL.NUM_CLASSES if len(cfg.MODEL.NUM_CLASSES) > 1 else cfg.MODEL.NUM_CLASSES[0],
            pool_size=[[nf // cfg.SLOWFAST.ALPHA // 4 // pool_size[0][0], nq // 32 // pool_size[0][1]],
                       [nf // 4 // pool_size[1][0], nq // 32 // pool_size[1][1]]],
            dropout_rate=cfg.MODEL.DROPOUT_RATE, act_func=cfg.MODEL.HEAD_ACT)


@MODEL_REGISTRY.register()
class ResNet(_EngineModel):
    """Single-pathway ResNet (cfg.MODEL.ARCH 'slow' or 'fast'), no lateral connections."""

    def __init__(self, cfg):
        super().__init__()
        self.norm_module = get_norm(cfg)
        self.num_pathways = 1
        self._construct_network(cfg)
        self._finish(cfg)

    def _construct_network(self, cfg):
        assert cfg.MODEL.ARCH in _POOL1.keys()
        pool_size = _POOL1[cfg.MODEL.ARCH]
        assert len({len(pool_size), self.num_pathways}) == 1
        assert cfg.RESNET.DEPTH in _MODEL_STAGE_DEPTH.keys()
        depths = _MODEL_STAGE_DEPTH[cfg.RESNET.DEPTH]
        groups, wpg = cfg.RESNET.NUM_GROUPS, cfg.RESNET.WIDTH_PER_GROUP
        inner = groups * wpg
        tk = _TEMPORAL_KERNEL_BASIS[cfg.MODEL.ARCH]
        norm = self.norm_module
        self.s1 = stem_helper.AudioModelStem(
            dim_in=cfg.DATA.INPUT_CHANNEL_NUM, dim_out=[wpg], kernel=[tk[0][0] + [7]], stride=[[2, 2]],
            padding=[[tk[0][0][0] // 2, 3]], norm_module=norm)
        w_in = wpg
        for si, name in enumerate(("s2", "s3", "s4", "s5")):
            w_out = wpg * 4 * (2 ** si)
            stage = resnet_helper.ResStage(
                dim_in=[w_in], dim_out=[w_out], dim_inner=[inner * (2 ** si)], temp_kernel_sizes=tk[si + 1],
                num_blocks=[depths[si]], stride_1x1=cfg.RESNET.STRIDE_1X1, inplace_relu=cfg.RESNET.INPLACE_RELU,
                norm_module=norm, **_stage_kwargs(cfg, si, 1))
            setattr(self, name, stage)
            if si == 0:
                self.add_module("pathway0_pool",
                                nn.MaxPool2d(kernel_size=pool_size[0], stride=pool_size[0], padding=[0, 0]))
            w_in = w_out
        nf, nq = cfg.AUDIO_DATA.NUM_FRAMES, cfg.AUDIO_DATA.NUM_FREQUENCIES
        self.head = head_helper.ResNetBasicHead(
            dim_in=[wpg * 32],
            num_classes=cfg.MODEL.NUM_CLASSES if len(cfg.MODEL.NUM_CLASSES) > 1 else cfg.MODEL.NUM_CLASSES[0],
            pool_size=[[nf // 4 // pool_size[0][0], nq // 32 // pool_size[0][1]]],
            dropout_rate=cfg.MODEL.DROPOUT_RATE, act_func=cfg.MODEL.HEAD_ACT)

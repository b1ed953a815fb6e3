"""ResNet-style initialisation (mirror of /root/reference/audio_slowfast/utils/weight_init_helper.py:10-39).

Conv: Kaiming normal, fan_out, relu gain (what fvcore's c2_msra_fill does); BN: gamma 1 (0 on each block's
final BN when zero_init_final_bn), beta 0; Linear: N(0, fc_init_std), bias 0.  Iterating ``modules()`` in
registration order reproduces the reference's RNG consumption, so ``torch.manual_seed(s); build_model(cfg)``
yields bit-identical parameters in both code bases (tests/test_boundary.py).
"""
import torch.nn as nn


def init_weights(model, fc_init_std=0.01, zero_init_final_bn=True):
    for m in model.modules():
        if isinstance(m, nn.Conv2d):
            nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
            if m.bias is not None:
                nn.init.constant_(m.bias, 0)
        elif isinstance(m, nn.BatchNorm2d):
            final = bool(getattr(m, "transform_final_bn", False))
            if m.weight is not None:
                m.weight.data.fill_(0.0 if (final and zero_init_final_bn) else 1.0)
            if m.bias is not None:
                m.bias.data.zero_()
        if isinstance(m, nn.Linear):
            m.weight.data.normal_(mean=0.0, std=fc_init_std)
            m.bias.data.zero_()

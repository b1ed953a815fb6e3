"""Stem parameter containers (mirror of /root/reference/audio_slowfast/models/stem_helper.py).

Same attribute names / parameter shapes as the reference (``pathway{p}_stem.conv.weight``,
``pathway{p}_stem.bn.*``) so state_dicts interchange.  The arithmetic -- conv -> BN -> ReLU ->
MaxPool(3x3, s2, p1), stem_helper.py:163-168 -- is executed by asf_b200.engine with CUDA kernels; these
modules only own parameters and hyper-parameters.
"""
import torch.nn as nn


class _EngineOnly(nn.Module):
    def forward(self, *a, **k):
        raise RuntimeError(
            "%s is a parameter container: its arithmetic runs inside the fused B200 engine "
            "(call the top-level model)" % type(self).__name__)


class ResNetBasicStem(_EngineOnly):
    def __init__(self, dim_in, dim_out, kernel, stride, padding, inplace_relu=True, eps=1e-5, bn_mmt=0.1,
                 norm_module=nn.BatchNorm2d):
        super().__init__()
        self.kernel, self.stride, self.padding = list(kernel), list(stride), list(padding)
        self.inplace_relu, self.eps, self.bn_mmt = inplace_relu, eps, bn_mmt
        self.conv = nn.Conv2d(dim_in, dim_out, self.kernel, stride=self.stride, padding=self.padding, bias=False)
        self.bn = norm_module(num_features=dim_out, eps=eps, momentum=bn_mmt)
        self.relu = nn.ReLU(inplace_relu)
        self.pool_layer = nn.MaxPool2d(kernel_size=[3, 3], stride=[2, 2], padding=[1, 1])


class AudioModelStem(_EngineOnly):
    def __init__(self, dim_in, dim_out, kernel, stride, padding, inplace_relu=True, eps=1e-5, bn_mmt=0.1,
                 norm_module=nn.BatchNorm2d):
        super().__init__()
        sizes = {len(dim_in), len(dim_out), len(kernel), len(stride), len(padding)}
        assert len(sizes) == 1, "Input pathway dimensions are not consistent."
        self.num_pathways = len(dim_in)
        self.kernel, self.stride, self.padding = kernel, stride, padding
        for p in range(self.num_pathways):
            self.add_module("pathway{}_stem".format(p),
                            ResNetBasicStem(dim_in[p], dim_out[p], kernel[p], stride[p], padding[p], inplace_relu,
                                            eps, bn_mmt, norm_module))

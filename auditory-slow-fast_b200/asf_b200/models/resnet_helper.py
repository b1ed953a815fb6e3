"""Residual-stage parameter containers (mirror of /root/reference/audio_slowfast/models/resnet_helper.py).

``BottleneckTransform`` = [Tk x 1] -> [1 x 3 (freq stride)] -> [1 x 1], each followed by BN (+ReLU on a, b)
(resnet_helper.py:170-225); ``ResBlock`` adds the (projection) shortcut and the final ReLU (:310-341);
``ResStage`` lays blocks out per pathway (:456-485).  Names, shapes and registration order match the
reference so ``state_dict`` keys and RNG consumption at construction are identical.  The arithmetic runs
in asf_b200.engine (tcgen05 implicit GEMM with the BN/ReLU/residual fused into the epilogue).
"""
import torch.nn as nn

from .stem_helper import _EngineOnly


def get_trans_func(name):
    if name == "bottleneck_transform":
        return BottleneckTransform
    if name == "basic_transform":
        raise NotImplementedError("basic_transform is never selected by a shipped config and is not on the B200 path")
    raise AssertionError("Transformation function '{}' not supported".format(name))


class BottleneckTransform(_EngineOnly):
    def __init__(self, dim_in, dim_out, temp_kernel_size, stride, dim_inner, num_groups, stride_1x1=False,
                 inplace_relu=True, eps=1e-5, bn_mmt=0.1, dilation=1, norm_module=nn.BatchNorm2d):
        super().__init__()
        if num_groups != 1 or dilation != 1:
            raise NotImplementedError("groups > 1 / dilation > 1 are outside the B200 hot path (every config uses 1)")
        self.temp_kernel_size = temp_kernel_size
        s_a, s_b = (stride, 1) if stride_1x1 else (1, stride)
        tk = temp_kernel_size
        self.a = nn.Conv2d(dim_in, dim_inner, kernel_size=[tk, 1], stride=[1, s_a], padding=[tk // 2, 0], bias=False)
        self.a_bn = norm_module(num_features=dim_inner, eps=eps, momentum=bn_mmt)
        self.a_relu = nn.ReLU(inplace=inplace_relu)
        self.b = nn.Conv2d(dim_inner, dim_inner, [1, 3], stride=[1, s_b], padding=[0, dilation], groups=num_groups,
                           bias=False, dilation=[1, dilation])
        self.b_bn = norm_module(num_features=dim_inner, eps=eps, momentum=bn_mmt)
        self.b_relu = nn.ReLU(inplace=inplace_relu)
        self.c = nn.Conv2d(dim_inner, dim_out, kernel_size=[1, 1], stride=[1, 1], padding=[0, 0], bias=False)
        self.c_bn = norm_module(num_features=dim_out, eps=eps, momentum=bn_mmt)
        self.c_bn.transform_final_bn = True


class ResBlock(_EngineOnly):
    def __init__(self, dim_in, dim_out, temp_kernel_size, stride, trans_func, dim_inner, num_groups=1,
                 stride_1x1=False, inplace_relu=True, eps=1e-5, bn_mmt=0.1, dilation=1, norm_module=nn.BatchNorm2d):
        super().__init__()
        if dim_in != dim_out or stride != 1:
            self.branch1 = nn.Conv2d(dim_in, dim_out, kernel_size=1, stride=[1, stride], padding=0, bias=False,
                                     dilation=1)
            self.branch1_bn = norm_module(num_features=dim_out, eps=eps, momentum=bn_mmt)
        self.branch2 = trans_func(dim_in, dim_out, temp_kernel_size, stride, dim_inner, num_groups,
                                  stride_1x1=stride_1x1, inplace_relu=inplace_relu, dilation=dilation,
                                  norm_module=norm_module)
        self.relu = nn.ReLU(inplace_relu)


class ResStage(_EngineOnly):
    def __init__(self, dim_in, dim_out, stride, temp_kernel_sizes, num_blocks, dim_inner, num_groups,
                 num_block_temp_kernel, dilation, trans_func_name="bottleneck_transform", stride_1x1=False,
                 inplace_relu=True, norm_module=nn.BatchNorm2d):
        super().__init__()
        npath = len(num_blocks)
        assert all(num_block_temp_kernel[i] <= num_blocks[i] for i in range(len(temp_kernel_sizes)))
        lens = {len(dim_in), len(dim_out), len(temp_kernel_sizes), len(stride), len(num_blocks), len(dim_inner),
                len(num_groups), len(num_block_temp_kernel)}
        assert len(lens) == 1
        self.num_blocks = num_blocks
        self.num_pathways = npath
        # temporal kernel of block i: the stage's kernel for the first num_block_temp_kernel blocks, 1 after
        self.temp_kernel_sizes = [
            (temp_kernel_sizes[p] * num_blocks[p])[: num_block_temp_kernel[p]]
            + [1] * (num_blocks[p] - num_block_temp_kernel[p])
            for p in range(npath)
        ]
        trans = get_trans_func(trans_func_name)
        for p in range(npath):
            for i in range(num_blocks[p]):
                blk = ResBlock(dim_in[p] if i == 0 else dim_out[p], dim_out[p], self.temp_kernel_sizes[p][i],
                               stride[p] if i == 0 else 1, trans, dim_inner[p], num_groups[p], stride_1x1=stride_1x1,
                               inplace_relu=inplace_relu, dilation=dilation[p], norm_module=norm_module)
                self.add_module("pathway{}_res{}".format(p, i), blk)

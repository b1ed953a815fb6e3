"""Head parameter container (mirror of /root/reference/audio_slowfast/models/head_helper.py:10-109).

AvgPool per pathway -> concat -> dropout -> Linear verb / noun (or a single ``projection``);
training returns raw logits, eval returns softmax probabilities averaged over the (1x1) pooled grid.
Executed by asf_b200.engine (warp-reduction pooling kernel + CUDA-core GEMMs).
"""
import torch.nn as nn

from .stem_helper import _EngineOnly


class ResNetBasicHead(_EngineOnly):
    def __init__(self, dim_in, num_classes, pool_size, dropout_rate=0.0, act_func="softmax"):
        super().__init__()
        assert len({len(pool_size), len(dim_in)}) == 1, "pathway dimensions are not consistent."
        self.num_pathways = len(pool_size)
        self.dim_in, self.pool_size, self.dropout_rate = list(dim_in), pool_size, dropout_rate
        for p in range(self.num_pathways):
            self.add_module("pathway{}_avgpool".format(p), nn.AvgPool2d(pool_size[p], stride=1))
        if dropout_rate > 0.0:
            self.dropout = nn.Dropout(dropout_rate)
        if isinstance(num_classes, (list, tuple)):
            self.projection_verb = nn.Linear(sum(dim_in), num_classes[0], bias=True)
            self.projection_noun = nn.Linear(sum(dim_in), num_classes[1], bias=True)
        else:
            self.projection = nn.Linear(sum(dim_in), num_classes, bias=True)
        self.num_classes = num_classes
        if act_func == "softmax":
            self.act = nn.Softmax(dim=3)
        elif act_func == "sigmoid":
            raise NotImplementedError("sigmoid head activation is not used by any shipped config (HEAD_ACT softmax)")
        else:
            raise NotImplementedError("{} is not supported as an activation function.".format(act_func))

    def projections(self):
        if isinstance(self.num_classes, (list, tuple)):
            return [self.projection_verb, self.projection_noun]
        return [self.projection]

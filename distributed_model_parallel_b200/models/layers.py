"""Small glue modules shared by the model zoo.

``Reshape`` / ``Reshape1`` exist in the reference so that the functional tail of ``MobileNetV2.forward``
can live inside an ``nn.Sequential`` pipeline stage
(/root/reference/code/distributed_training/distributed_layers.py:64-69, model/mobilenetv2.py:150-158).
The versions here route through the fused op layer.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from ..ops import functional as OF


class Reshape(nn.Module):
    """``x.view(*shape)``."""

    def __init__(self, *shape):
        super().__init__()
        self.shape = tuple(shape)

    def forward(self, x):
        return x.reshape(self.shape)

    def extra_repr(self):
        return "shape=%s" % (self.shape,)


class Reshape1(nn.Module):
    """relu -> global average pool -> flatten: the head glue of the CIFAR MobileNetV2."""

    def forward(self, x):
        return OF.global_avgpool_flatten(x, relu=True)


class ConvBNAct(nn.Module):
    """Adapter that runs ``conv -> bn -> (+res) -> relu`` through the fused op while the parameters stay in
    ordinary ``nn.Conv2d`` / ``nn.BatchNorm2d`` children owned by the parent (so state_dict keys are unchanged)."""

    def __init__(self, conv: nn.Conv2d, bn=None, relu: bool = True):
        super().__init__()
        self.conv, self.bn, self.relu = conv, bn, relu

    def forward(self, x, residual=None):
        return OF.conv_bn_act(x, self.conv, self.bn, self.relu, residual)


class FusedLinear(nn.Linear):
    """``nn.Linear`` whose forward runs the tcgen05 / skinny-linear kernels on CUDA."""

    def forward(self, x):
        return OF.linear(x, self)


class StemReLU(nn.Module):
    """Applies the conv1->bn1->relu stem as one fused op inside an ``nn.Sequential`` stage."""

    def __init__(self, conv, bn, relu=True):
        super().__init__()
        self.conv1, self.bn1, self.relu = conv, bn, relu

    def forward(self, x):
        return OF.conv_bn_act(x, self.conv1, self.bn1, self.relu)

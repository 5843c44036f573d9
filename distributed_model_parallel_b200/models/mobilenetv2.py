"""MobileNetV2 for 32x32 inputs (the only model the reference trains) and its no-BatchNorm variant.

Architecture parity with /root/reference/code/distributed_training/model/mobilenetv2.py:10-76 (Block,
MobileNetV2) and :84-148 (Block_nobn, MobileNetV2_nobn): 17 inverted-residual blocks from the
(expansion, out_planes, num_blocks, stride) table, stride-1 3x3 stem, 1x1 head to 1280, 4x4 average pool,
10-way linear.  Parameter names are identical (``conv1, bn1, layers.<i>.conv{1,2,3}/bn{1,2,3}/shortcut.{0,1},
conv2, bn2, linear``) so reference checkpoints load with ``load_state_dict``.

The implementation is not the reference's: every conv->bn->relu(+residual) chain is one call into the fused
op layer (tcgen05 GEMM for the 39 pointwise convs, depthwise stencil kernel, fused BN statistics/apply), and
activations travel as bf16 channels-last tensors on CUDA.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from ..ops import functional as OF
from .layers import FusedLinear, Reshape1, StemReLU

# (expansion, out_planes, num_blocks, stride) -- stride of stage 2 and of the stem reduced to 1 for CIFAR
CIFAR_CFG = [(1, 16, 1, 1), (6, 24, 2, 1), (6, 32, 3, 2), (6, 64, 4, 2), (6, 96, 3, 1), (6, 160, 3, 2), (6, 320, 1, 1)]


class Block(nn.Module):
    """expand 1x1 -> depthwise 3x3 -> project 1x1, residual when stride == 1."""

    def __init__(self, in_planes, out_planes, expansion, stride, batchnorm=True, shortcut_bn=True):
        super().__init__()
        self.stride = stride
        planes = expansion * in_planes
        self.batchnorm = batchnorm
        convs = [nn.Conv2d(in_planes, planes, 1, bias=False),
                 nn.Conv2d(planes, planes, 3, stride=stride, padding=1, groups=planes, bias=False),
                 nn.Conv2d(planes, out_planes, 1, bias=False)]
        widths = [planes, planes, out_planes]
        for i, (conv, width) in enumerate(zip(convs, widths), start=1):  # conv_i then bn_i: reference key order
            setattr(self, "conv%d" % i, conv)
            if batchnorm:
                setattr(self, "bn%d" % i, nn.BatchNorm2d(width))
        self.shortcut = nn.Sequential()
        if stride == 1 and in_planes != out_planes:
            mods = [nn.Conv2d(in_planes, out_planes, 1, bias=False)]
            if shortcut_bn:
                mods.append(nn.BatchNorm2d(out_planes))
            self.shortcut = nn.Sequential(*mods)

    def _bn(self, i):
        return getattr(self, "bn%d" % i) if self.batchnorm else None

    def forward(self, x):
        # expand -> bn1 -> relu -> depthwise -> bn2 -> relu: one fused chain on the kernel path; its output feeds conv3 only
        out = OF.expand_dw_bn_act(x, self.conv1, self._bn(1), self.conv2, self._bn(2), single_consumer=True)
        if self.stride != 1:
            return OF.conv_bn_act(out, self.conv3, self._bn(3), relu=False)
        if len(self.shortcut) == 0:
            res = x
        else:
            sc_bn = self.shortcut[1] if len(self.shortcut) > 1 else None
            res = OF.conv_bn_act(x, self.shortcut[0], sc_bn, relu=False)
        return OF.conv_bn_act(out, self.conv3, self._bn(3), relu=False, residual=res)


class MobileNetV2(nn.Module):
    cfg = CIFAR_CFG

    def __init__(self, num_classes=10, batchnorm=True, shortcut_bn=True, pool=4):
        super().__init__()
        self.batchnorm = batchnorm
        self.pool = pool
        self.conv1 = nn.Conv2d(3, 32, 3, stride=1, padding=1, bias=False)
        if batchnorm:
            self.bn1 = nn.BatchNorm2d(32)
        blocks, in_planes = [], 32
        for expansion, out_planes, num_blocks, stride in self.cfg:
            for s in [stride] + [1] * (num_blocks - 1):
                blocks.append(Block(in_planes, out_planes, expansion, s, batchnorm, shortcut_bn))
                in_planes = out_planes
        self.layers = nn.Sequential(*blocks)
        self.conv2 = nn.Conv2d(320, 1280, 1, bias=False)
        if batchnorm:
            self.bn2 = nn.BatchNorm2d(1280)
        self.linear = FusedLinear(1280, num_classes)

    def forward(self, x):
        out = OF.conv_bn_act(x, self.conv1, self.bn1 if self.batchnorm else None, relu=True)
        out = self.layers(out)
        # head conv + bn; the ReLU is folded into the pooling kernel (relu -> avg_pool -> flatten)
        out = OF.conv_bn_act(out, self.conv2, self.bn2 if self.batchnorm else None, relu=False)
        out = OF.global_avgpool_flatten(out, relu=True)
        return self.linear(out)

    # ------------------------------------------------------------------ pipeline support
    def as_sequential(self, stem_relu: bool = True) -> nn.Sequential:
        """Flat ``nn.Sequential`` view sharing this model's parameters, for the layer-split partitioner.

        ``stem_relu=False`` reproduces the reference's model-parallel stage 0, which drops the stem ReLU
        (/root/reference/code/distributed_training/model_parallel.py:103; SURVEY.md defect Q2)."""
        from .layers import ConvBNAct

        head = ConvBNAct(self.conv2, self.bn2 if self.batchnorm else None, relu=False)
        stem = StemReLU(self.conv1, self.bn1 if self.batchnorm else None, relu=stem_relu)
        return nn.Sequential(stem, *list(self.layers), head, Reshape1(), self.linear)


def MobileNetV2_nobn(num_classes=10, shortcut_bn=True):
    """Main-path BatchNorm removed (large-batch-without-BN experiment, /root/reference/Readme.md:159-176).
    ``shortcut_bn=True`` keeps the four shortcut BN layers exactly as the reference's Block_nobn does
    (mobilenetv2.py:99-103); pass False for a genuinely BN-free network."""
    return MobileNetV2(num_classes, batchnorm=False, shortcut_bn=shortcut_bn)


Block_nobn = lambda i, o, e, s: Block(i, o, e, s, batchnorm=False)  # noqa: E731  (name parity)

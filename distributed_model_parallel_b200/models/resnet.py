"""ResNet-18 (CIFAR, 32x32) and ResNet-50 (ImageNet, 224x224) -- north-star configs 2 and 3 of BASELINE.json.

NOT IN REFERENCE: the reference only lists ``ResNet18()`` in a commented-out model zoo
(/root/reference/code/distributed_training/data_parallel.py:58-73).  Both nets use the fused
conv->bn->(+res)->relu op; 1x1 convolutions run on the tcgen05 GEMM, dense 3x3/7x7 convolutions currently go
through the library convolution inside the same fused op (see ops/functional.py ``conv_kind``).
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..ops import functional as OF
from .layers import FusedLinear


class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, in_planes, planes, stride=1):
        super().__init__()
        self.conv1 = nn.Conv2d(in_planes, planes, 3, stride, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(planes)
        self.conv2 = nn.Conv2d(planes, planes, 3, 1, 1, bias=False)
        self.bn2 = nn.BatchNorm2d(planes)
        self.shortcut = nn.Sequential()
        if stride != 1 or in_planes != planes:
            self.shortcut = nn.Sequential(nn.Conv2d(in_planes, planes, 1, stride, bias=False), nn.BatchNorm2d(planes))

    def forward(self, x):
        out = OF.conv_bn_act(x, self.conv1, self.bn1, relu=True)
        res = x if len(self.shortcut) == 0 else OF.conv_bn_act(x, self.shortcut[0], self.shortcut[1], relu=False)
        return OF.conv_bn_act(out, self.conv2, self.bn2, relu=True, residual=res)


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, in_planes, planes, stride=1):
        super().__init__()
        self.conv1 = nn.Conv2d(in_planes, planes, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(planes)
        self.conv2 = nn.Conv2d(planes, planes, 3, stride, 1, bias=False)
        self.bn2 = nn.BatchNorm2d(planes)
        self.conv3 = nn.Conv2d(planes, planes * 4, 1, bias=False)
        self.bn3 = nn.BatchNorm2d(planes * 4)
        self.shortcut = nn.Sequential()
        if stride != 1 or in_planes != planes * 4:
            self.shortcut = nn.Sequential(nn.Conv2d(in_planes, planes * 4, 1, stride, bias=False), nn.BatchNorm2d(planes * 4))

    def forward(self, x):
        out = OF.conv_bn_act(x, self.conv1, self.bn1, relu=True)
        out = OF.conv_bn_act(out, self.conv2, self.bn2, relu=True)
        res = x if len(self.shortcut) == 0 else OF.conv_bn_act(x, self.shortcut[0], self.shortcut[1], relu=False)
        return OF.conv_bn_act(out, self.conv3, self.bn3, relu=True, residual=res)


class _Stem(nn.Module):
    def __init__(self, conv, bn, maxpool):
        super().__init__()
        self.conv1, self.bn1, self.maxpool = conv, bn, maxpool

    def forward(self, x):
        out = OF.conv_bn_act(x, self.conv1, self.bn1, relu=True)
        return OF.max_pool2d(out, 3, 2, 1) if self.maxpool else out


class _Head(nn.Module):
    def __init__(self, linear):
        super().__init__()
        self.linear = linear

    def forward(self, x):
        return self.linear(OF.global_avgpool_flatten(x))


class ResNet(nn.Module):
    def __init__(self, block, num_blocks, num_classes=10, imagenet=False):
        super().__init__()
        self.imagenet = imagenet
        self.in_planes = 64
        if imagenet:
            self.conv1 = nn.Conv2d(3, 64, 7, 2, 3, bias=False)
        else:
            self.conv1 = nn.Conv2d(3, 64, 3, 1, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(64)
        self.layer1 = self._make_layer(block, 64, num_blocks[0], 1)
        self.layer2 = self._make_layer(block, 128, num_blocks[1], 2)
        self.layer3 = self._make_layer(block, 256, num_blocks[2], 2)
        self.layer4 = self._make_layer(block, 512, num_blocks[3], 2)
        self.linear = FusedLinear(512 * block.expansion, num_classes)

    def _make_layer(self, block, planes, n, stride):
        layers = []
        for s in [stride] + [1] * (n - 1):
            layers.append(block(self.in_planes, planes, s))
            self.in_planes = planes * block.expansion
        return nn.Sequential(*layers)

    def forward(self, x):
        out = OF.conv_bn_act(x, self.conv1, self.bn1, relu=True)
        if self.imagenet:
            out = OF.max_pool2d(out, 3, 2, 1)
        out = self.layer4(self.layer3(self.layer2(self.layer1(out))))
        return self.linear(OF.global_avgpool_flatten(out))

    def as_sequential(self) -> nn.Sequential:
        blocks = [*self.layer1, *self.layer2, *self.layer3, *self.layer4]
        return nn.Sequential(_Stem(self.conv1, self.bn1, self.imagenet), *blocks, _Head(self.linear))


def ResNet18(num_classes=10):
    return ResNet(BasicBlock, [2, 2, 2, 2], num_classes, imagenet=False)


def ResNet50(num_classes=1000):
    return ResNet(Bottleneck, [3, 4, 6, 3], num_classes, imagenet=True)

"""VGG-16 (with BatchNorm) -- layer-split pipeline workload of BASELINE.json config 4.

NOT IN REFERENCE beyond a commented-out ``VGG('VGG19')`` line
(/root/reference/code/distributed_training/data_parallel.py:58).  ``features`` is a flat ``nn.Sequential`` of
conv-bn-relu units and max-pools, which is exactly what the pipeline partitioner splits.
"""
from __future__ import annotations

import torch.nn as nn
import torch.nn.functional as F

from ..ops import functional as OF
from .layers import FusedLinear

CFG16 = [64, 64, "M", 128, 128, "M", 256, 256, 256, "M", 512, 512, 512, "M", 512, 512, 512, "M"]


class ConvUnit(nn.Module):
    def __init__(self, cin, cout, batchnorm=True):
        super().__init__()
        self.conv = nn.Conv2d(cin, cout, 3, 1, 1, bias=False)
        self.bn = nn.BatchNorm2d(cout) if batchnorm else None

    def forward(self, x):
        return OF.conv_bn_act(x, self.conv, self.bn, relu=True)


class MaxPool(nn.Module):
    def forward(self, x):
        return OF.max_pool2d(x, 2, 2)


class Classifier(nn.Module):
    def __init__(self, cin, num_classes, hidden=None):
        super().__init__()
        self.hidden = hidden
        if hidden:
            self.fc1 = FusedLinear(cin, hidden)
            self.fc2 = FusedLinear(hidden, hidden)
            self.fc3 = FusedLinear(hidden, num_classes)
        else:
            self.fc = FusedLinear(cin, num_classes)

    def forward(self, x):
        x = x.flatten(1) if x.dim() == 4 and x.shape[2] * x.shape[3] == 1 else OF.global_avgpool_flatten(x)
        if self.hidden:
            x = F.relu(self.fc1(x))
            x = F.relu(self.fc2(x))
            return self.fc3(x)
        return self.fc(x)


class VGG(nn.Module):
    def __init__(self, cfg=CFG16, num_classes=10, batchnorm=True, hidden=None):
        super().__init__()
        layers, cin = [], 3
        for v in cfg:
            if v == "M":
                layers.append(MaxPool())
            else:
                layers.append(ConvUnit(cin, v, batchnorm))
                cin = v
        self.features = nn.Sequential(*layers)
        self.classifier = Classifier(cin, num_classes, hidden)

    def forward(self, x):
        return self.classifier(self.features(x))

    def as_sequential(self) -> nn.Sequential:
        return nn.Sequential(*list(self.features), self.classifier)


def VGG16(num_classes=10, imagenet=False):
    return VGG(CFG16, num_classes, True, hidden=4096 if imagenet else None)

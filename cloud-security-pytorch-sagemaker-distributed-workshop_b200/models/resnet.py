"""ResNet-18 / ResNet-50 with torchvision-compatible ``state_dict`` keys.

The reference trains ``torchvision.models.resnet18(pretrained=False)`` — ImageNet topology (7x7/2 stem,
3x3/2 max-pool, 1000-way ``fc``) applied to 32x32 CIFAR images
(``notebooks/code/cifar10-distributed-smddp-gpu.py:30-32``; SURVEY §2.1 "known defects": layer4 runs at
1x1 spatial).  The same topology is kept so the work per image matches the reference's job; ResNet-50
is the BASELINE.json config that goes beyond the reference.

Built from ``b200ddp.nn`` modules: every BN is fused with its ReLU and (in the block tail) the
residual add; every conv is the tcgen05 implicit-GEMM kernel.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from .. import nn as bnn
from ..ops import functional as BF


def _conv3x3(cin, cout, stride=1):
    return bnn.Conv2d(cin, cout, 3, stride=stride, padding=1, bias=False)


def _conv1x1(cin, cout, stride=1):
    return bnn.Conv2d(cin, cout, 1, stride=stride, padding=0, bias=False)


class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, inplanes, planes, stride=1, downsample=None):
        super().__init__()
        self.conv1 = _conv3x3(inplanes, planes, stride)
        self.bn1 = bnn.BatchNorm2d(planes, relu=True)
        self.conv2 = _conv3x3(planes, planes)
        self.bn2 = bnn.BatchNorm2d(planes, relu=True)     # relu(bn2(.) + identity)
        self.downsample = downsample

    def forward(self, x):
        x = BF.fork(x)      # consumed by conv1 and by the skip connection: their gradients meet in a dgrad epilogue
        identity = x if self.downsample is None else self.downsample(x)
        out = self.bn1(self.conv1(x))
        return self.bn2(self.conv2(out), identity)


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, inplanes, planes, stride=1, downsample=None):
        super().__init__()
        self.conv1 = _conv1x1(inplanes, planes)
        self.bn1 = bnn.BatchNorm2d(planes, relu=True)
        self.conv2 = _conv3x3(planes, planes, stride)
        self.bn2 = bnn.BatchNorm2d(planes, relu=True)
        self.conv3 = _conv1x1(planes, planes * 4)
        self.bn3 = bnn.BatchNorm2d(planes * 4, relu=True)  # relu(bn3(.) + identity)
        self.downsample = downsample

    def forward(self, x):
        x = BF.fork(x)
        identity = x if self.downsample is None else self.downsample(x)
        out = self.bn1(self.conv1(x))
        out = self.bn2(self.conv2(out))
        return self.bn3(self.conv3(out), identity)


class ResNet(nn.Module):
    def __init__(self, block, layers, num_classes=1000):
        super().__init__()
        self.inplanes = 64
        self.num_classes = num_classes
        self.conv1 = bnn.Conv2d(3, 64, 7, stride=2, padding=3, bias=False)
        self.bn1 = bnn.BatchNorm2d(64, relu=True)
        self.maxpool = bnn.MaxPool2d(3, 2, 1)
        self.layer1 = self._make_layer(block, 64, layers[0])
        self.layer2 = self._make_layer(block, 128, layers[1], stride=2)
        self.layer3 = self._make_layer(block, 256, layers[2], stride=2)
        self.layer4 = self._make_layer(block, 512, layers[3], stride=2)
        self.avgpool = bnn.GlobalAvgPool()
        self.fc = bnn.Linear(512 * block.expansion, num_classes)
        for m in self.modules():       # torchvision's init
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
            elif isinstance(m, nn.BatchNorm2d):
                nn.init.constant_(m.weight, 1)
                nn.init.constant_(m.bias, 0)
        # every conv here feeds exactly one BatchNorm: let the conv epilogue produce its batch statistics
        self.conv1.fuse_bn_stats(self.bn1)
        for m in self.modules():
            for i in (1, 2, 3):
                conv, bn = getattr(m, "conv%d" % i, None), getattr(m, "bn%d" % i, None)
                if m is not self and isinstance(conv, bnn.Conv2d) and isinstance(bn, bnn.BatchNorm2d):
                    conv.fuse_bn_stats(bn)
            ds = getattr(m, "downsample", None)
            if isinstance(ds, nn.Sequential):
                ds[0].fuse_bn_stats(ds[1])

    def _make_layer(self, block, planes, blocks, stride=1):
        downsample = None
        if stride != 1 or self.inplanes != planes * block.expansion:
            downsample = nn.Sequential(_conv1x1(self.inplanes, planes * block.expansion, stride),
                                       bnn.BatchNorm2d(planes * block.expansion))
        layers = [block(self.inplanes, planes, stride, downsample)]
        self.inplanes = planes * block.expansion
        layers += [block(self.inplanes, planes) for _ in range(1, blocks)]
        return nn.Sequential(*layers)

    def forward(self, x):
        x = self.maxpool(self.bn1(self.conv1(x)))
        x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        x = self.avgpool(x)
        return self.fc(x)


def resnet18(num_classes=1000, **_):
    return ResNet(BasicBlock, [2, 2, 2, 2], num_classes)


def resnet50(num_classes=1000, **_):
    return ResNet(Bottleneck, [3, 4, 6, 3], num_classes)

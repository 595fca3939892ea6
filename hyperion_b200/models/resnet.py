"""ResNet-18 / ResNet-50 with torchvision-compatible ``state_dict`` keys
(``conv1 bn1 layer1..4.N.{conv,bn}K downsample.{0,1} fc``).

The reference builds them with ``torchvision.models.resnet18(num_classes=10)``
(02_development/distributed_utils.py:229 -- stock ImageNet stem on 32x32 CIFAR
images) and ``torch.hub.load('pytorch/vision', 'resnet50')`` (Phase 1/
baseline_performance.ipynb c0 L23, needs network).  Here they are built locally.
Conv -> BN -> ReLU (+ residual) goes through :mod:`hyperion_b200.ops.conv`, which
runs channels-last and fuses the BN-apply + ReLU (+ residual add) passes.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from ..ops import conv as convops


class _ConvBN(nn.Module):
    """Helper holding ``conv`` + ``bn`` parameters under caller-chosen attribute names."""


class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, inplanes, planes, stride=1, downsample=None):
        super().__init__()
        self.conv1 = nn.Conv2d(inplanes, planes, 3, stride, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(planes)
        self.conv2 = nn.Conv2d(planes, planes, 3, 1, 1, bias=False)
        self.bn2 = nn.BatchNorm2d(planes)
        self.downsample = downsample

    def forward(self, x):
        idt = x
        out = convops.conv_bn_act(x, self.conv1, self.bn1, relu=True)
        if self.downsample is not None:
            idt = convops.conv_bn_act(x, self.downsample[0], self.downsample[1], relu=False)
        return convops.conv_bn_act(out, self.conv2, self.bn2, relu=True, residual=idt)


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, inplanes, planes, stride=1, downsample=None):
        super().__init__()
        self.conv1 = nn.Conv2d(inplanes, planes, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(planes)
        self.conv2 = nn.Conv2d(planes, planes, 3, stride, 1, bias=False)
        self.bn2 = nn.BatchNorm2d(planes)
        self.conv3 = nn.Conv2d(planes, planes * 4, 1, bias=False)
        self.bn3 = nn.BatchNorm2d(planes * 4)
        self.downsample = downsample

    def forward(self, x):
        idt = x
        out = convops.conv_bn_act(x, self.conv1, self.bn1, relu=True)
        out = convops.conv_bn_act(out, self.conv2, self.bn2, relu=True)
        if self.downsample is not None:
            idt = convops.conv_bn_act(x, self.downsample[0], self.downsample[1], relu=False)
        return convops.conv_bn_act(out, self.conv3, self.bn3, relu=True, residual=idt)


class ResNet(nn.Module):
    def __init__(self, block, layers, num_classes: int = 1000):
        super().__init__()
        self.inplanes = 64
        self.conv1 = nn.Conv2d(3, 64, 7, 2, 3, bias=False)
        self.bn1 = nn.BatchNorm2d(64)
        self.maxpool = nn.MaxPool2d(3, 2, 1)
        self.layer1 = self._make_layer(block, 64, layers[0])
        self.layer2 = self._make_layer(block, 128, layers[1], 2)
        self.layer3 = self._make_layer(block, 256, layers[2], 2)
        self.layer4 = self._make_layer(block, 512, layers[3], 2)
        self.avgpool = nn.AdaptiveAvgPool2d((1, 1))
        self.fc = nn.Linear(512 * block.expansion, num_classes)
        self.use_checkpoint = False
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
                # weights live in [K, R, S, C] memory order (channels_last), what the NHWC convolution kernels read;
                # values and state_dict keys/shapes are unchanged
                m.weight.data = m.weight.data.contiguous(memory_format=torch.channels_last)

    def _make_layer(self, block, planes, blocks, stride=1):
        downsample = None
        if stride != 1 or self.inplanes != planes * block.expansion:
            downsample = nn.Sequential(nn.Conv2d(self.inplanes, planes * block.expansion, 1, stride, bias=False),
                                       nn.BatchNorm2d(planes * block.expansion))
        layers = [block(self.inplanes, planes, stride, downsample)]
        self.inplanes = planes * block.expansion
        layers += [block(self.inplanes, planes) for _ in range(1, blocks)]
        return nn.Sequential(*layers)

    def forward(self, x):
        x = convops.to_channels_last(x)
        x = self.maxpool(convops.conv_bn_act(x, self.conv1, self.bn1, relu=True))
        stages = [self.layer1, self.layer2, self.layer3, self.layer4]
        if self.use_checkpoint and self.training and torch.is_grad_enabled():
            # memory_optimization.ipynb c4 L1-14: checkpoint_sequential([layer1..4], 4, x)
            from ..memory import checkpoint_sequential

            x = checkpoint_sequential(stages, len(stages), x)
        else:
            for s in stages:
                x = s(x)
        x = self.avgpool(x).flatten(1)
        return self.fc(x)


def resnet18(num_classes: int = 1000) -> ResNet:
    return ResNet(BasicBlock, [2, 2, 2, 2], num_classes)


def resnet50(num_classes: int = 1000) -> ResNet:
    return ResNet(Bottleneck, [3, 4, 6, 3], num_classes)


def create_resnet50() -> ResNet:
    """Factory name used by the baseline benchmark (baseline_performance.ipynb c0 L21-23)."""
    return resnet50()

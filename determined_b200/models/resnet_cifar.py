"""ResNet-18 for 32x32 inputs (CIFAR-10) — BASELINE config #2.

The reference has no CIFAR model definition (only YAMLs pointing at a missing
``model_def:CIFARTrial``, ``examples/legacy/computer_vision/cifar10_pytorch/*.yaml``),
so this repository supplies one.  Plain torch.nn so the *same* model definition runs under the
reference harness for the baseline arm.  Convolutions stay on cuDNN.
"""
from typing import List, Type

import torch
import torch.nn as nn
import torch.nn.functional as F


class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, in_planes: int, planes: int, stride: int = 1) -> None:
        super().__init__()
        self.conv1 = nn.Conv2d(in_planes, planes, 3, stride, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(planes)
        self.conv2 = nn.Conv2d(planes, planes, 3, 1, 1, bias=False)
        self.bn2 = nn.BatchNorm2d(planes)
        self.shortcut: nn.Module = nn.Sequential()
        if stride != 1 or in_planes != planes:
            self.shortcut = nn.Sequential(nn.Conv2d(in_planes, planes, 1, stride, bias=False), nn.BatchNorm2d(planes))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        out = F.relu(self.bn1(self.conv1(x)), inplace=True)
        out = self.bn2(self.conv2(out))
        return F.relu(out + self.shortcut(x), inplace=True)


class ResNetCIFAR(nn.Module):
    def __init__(self, num_blocks: List[int], num_classes: int = 10, width: int = 64) -> None:
        super().__init__()
        self.in_planes = width
        self.conv1 = nn.Conv2d(3, width, 3, 1, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(width)
        self.layer1 = self._make_layer(width, num_blocks[0], 1)
        self.layer2 = self._make_layer(width * 2, num_blocks[1], 2)
        self.layer3 = self._make_layer(width * 4, num_blocks[2], 2)
        self.layer4 = self._make_layer(width * 8, num_blocks[3], 2)
        self.linear = nn.Linear(width * 8, num_classes)

    def _make_layer(self, planes: int, n: int, stride: int) -> nn.Sequential:
        layers = []
        for s in [stride] + [1] * (n - 1):
            layers.append(BasicBlock(self.in_planes, planes, s))
            self.in_planes = planes
        return nn.Sequential(*layers)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        out = F.relu(self.bn1(self.conv1(x)), inplace=True)
        out = self.layer4(self.layer3(self.layer2(self.layer1(out))))
        out = F.adaptive_avg_pool2d(out, 1).flatten(1)
        return self.linear(out)


def resnet18_cifar(num_classes: int = 10) -> ResNetCIFAR:
    """11.17 M parameters."""
    return ResNetCIFAR([2, 2, 2, 2], num_classes)

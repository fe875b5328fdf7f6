"""CIFAR-10-shape ResNet-18 (BASELINE.json configs 3 and 4).

No reference counterpart (the reference bundles only the MNIST CNN); the
architecture is the standard 3×3-stem ResNet-18 for 32×32 inputs with 10 classes:
11.17 M parameters, 122 ``state_dict`` entries of which 20 are int64
``num_batches_tracked`` scalars — exactly the mixed-dtype / 0-d case that the flat
layout and the DP aggregation must handle.
``norm="group"`` swaps BatchNorm for GroupNorm(8), the BN-free variant DP work
usually uses (and the only one the reference's DP aggregator can process).
"""

from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F


def _norm(kind: str, channels: int) -> nn.Module:
    if kind == "batch":
        return nn.BatchNorm2d(channels)
    if kind == "group":
        return nn.GroupNorm(8, channels)
    raise ValueError(f"unknown norm {kind!r}")


class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, cin: int, cout: int, stride: int, norm: str) -> None:
        super().__init__()
        self.conv1 = nn.Conv2d(cin, cout, 3, stride, 1, bias=False)
        self.bn1 = _norm(norm, cout)
        self.conv2 = nn.Conv2d(cout, cout, 3, 1, 1, bias=False)
        self.bn2 = _norm(norm, cout)
        self.shortcut: nn.Module = nn.Sequential()
        if stride != 1 or cin != cout:
            self.shortcut = nn.Sequential(
                nn.Conv2d(cin, cout, 1, stride, bias=False), _norm(norm, cout)
            )

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        out = F.relu(self.bn1(self.conv1(x)))
        out = self.bn2(self.conv2(out))
        return F.relu(out + self.shortcut(x))


class ResNet18(nn.Module):
    NUM_CLASSES = 10
    INPUT_SHAPE = (3, 32, 32)

    def __init__(self, num_classes: int = 10, norm: str = "batch") -> None:
        super().__init__()
        self.conv1 = nn.Conv2d(3, 64, 3, 1, 1, bias=False)
        self.bn1 = _norm(norm, 64)
        widths, strides = (64, 128, 256, 512), (1, 2, 2, 2)
        blocks: list[nn.Module] = []
        cin = 64
        for width, stride in zip(widths, strides):
            blocks.append(nn.Sequential(BasicBlock(cin, width, stride, norm),
                                        BasicBlock(width, width, 1, norm)))
            cin = width
        self.layer1, self.layer2, self.layer3, self.layer4 = blocks
        self.fc = nn.Linear(512, num_classes)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = F.relu(self.bn1(self.conv1(x)))
        x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        x = F.adaptive_avg_pool2d(x, 1).flatten(1)
        return self.fc(x)

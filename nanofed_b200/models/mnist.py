"""The bundled MNIST CNN.

Parity: reference ``models/mnist.py:6-28`` — ``conv1(1→32,3×3)`` → ReLU →
``conv2(32→64,3×3)`` → ReLU → ``max_pool2d(2)`` → ``Dropout(.25)`` → flatten(9216) →
``fc1(9216→128)`` → ReLU → ``Dropout(.5)`` → ``fc2(128→10)`` → ``log_softmax``; 8
state entries, 1 199 882 parameters, same names/shapes so checkpoints are
interchangeable with the reference.

This ``nn.Module`` is the *portable* definition (CPU, autograd, checkpoints).  The
B200 training path for it is ``models.mnist_cohort.MnistCohortNet``: the same
math for a whole cohort of clients at once, executed by the hand-written
sm_100a kernels in ``csrc/`` with the client index as the GEMM group dimension.
"""

from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F


class MNISTModel(nn.Module):
    NUM_CLASSES = 10
    INPUT_SHAPE = (1, 28, 28)

    def __init__(self) -> None:
        super().__init__()
        self.conv1 = nn.Conv2d(1, 32, 3, 1)
        self.conv2 = nn.Conv2d(32, 64, 3, 1)
        self.dropout1 = nn.Dropout(0.25)
        self.dropout2 = nn.Dropout(0.5)
        self.fc1 = nn.Linear(9216, 128)
        self.fc2 = nn.Linear(128, 10)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = F.relu(self.conv1(x))
        x = F.relu(self.conv2(x))
        x = self.dropout1(F.max_pool2d(x, 2))
        x = torch.flatten(x, 1)
        x = self.dropout2(F.relu(self.fc1(x)))
        return F.log_softmax(self.fc2(x), dim=1)

"""Cross-entropy classifier trainer (reference ``trainer/torch.py:7-22``).

``compute_loss`` is ``F.cross_entropy`` (the bundled ``MNISTModel`` already
applies ``log_softmax``; a second one is idempotent, as in the reference).
``compute_accuracy`` keeps the float-returning form; the loop itself uses
``compute_accuracy_tensor`` so no host sync happens per batch.
"""

from __future__ import annotations

import torch
import torch.nn.functional as F

from .base import BaseTrainer


class TorchTrainer(BaseTrainer):
    def compute_loss(self, output: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
        return F.cross_entropy(output, target)

    def compute_accuracy_tensor(self, output: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
        hits = (output.argmax(dim=1) == target.view(-1)).sum()
        return hits.to(torch.float32) / float(len(target))

    def compute_accuracy(self, output: torch.Tensor, target: torch.Tensor) -> float:
        pred = output.argmax(dim=1, keepdim=True)
        correct = pred.eq(target.view_as(pred)).sum().item()
        return correct / len(target)

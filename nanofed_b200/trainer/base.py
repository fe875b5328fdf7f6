"""Local-training loop shared by all trainers.

Parity: reference ``trainer/base.py:15-198`` — ``TrainingConfig`` and
``TrainingMetrics`` field-for-field (``to_dict`` drops ``epoch``/``batch``),
``Callback`` protocol incl. the misspelt ``on_eopch_start`` hook,
``train_epoch(model, dataloader, optimizer, epoch)`` honouring ``max_batches`` and
``log_interval``, and — kept on purpose, the tutorial's numbers depend on it
(``tutorial.rst:325-347``) — the **return value is the LAST batch's metrics**
with cumulative ``samples_processed`` (``base.py:198``), while callbacks'
``on_epoch_end`` receives the epoch averages.  The averages are also exposed as
``trainer.last_epoch_metrics``.

B200 path: the reference forces two device→host syncs per batch
(``loss.item()`` / ``.sum().item()``, ``base.py:153-159``), which serialises
the GPU behind Python.  Here per-batch loss and accuracy stay on the device
(``compute_accuracy_tensor``) and are fetched with ONE copy after the last
batch; callbacks' ``on_batch_end`` and the progress log lines are then replayed
in order with the real values.  Callbacks that must observe a batch *before* the
next one starts can set ``eager_callbacks=True`` on the trainer (restores the
per-batch sync).  A subclass that only implements the float-returning
``compute_accuracy`` still works (it syncs, like the reference).
"""

from __future__ import annotations

from abc import ABC, abstractmethod
from dataclasses import dataclass
from typing import Generic, Protocol, Sized, TypeVar, cast

import torch
import torch.nn as nn
from torch.utils.data import DataLoader

from nanofed_b200.core import ModelProtocol
from nanofed_b200.utils import Logger, log_exec

M = TypeVar("M", bound=ModelProtocol)


@dataclass(slots=True, frozen=True)
class TrainingConfig:
    epochs: int
    batch_size: int
    learning_rate: float
    device: str = "cpu"
    max_batches: int | None = None
    log_interval: int = 10


@dataclass(slots=True)
class TrainingMetrics:
    loss: float
    accuracy: float
    epoch: int
    batch: int
    samples_processed: int

    def to_dict(self) -> dict[str, float | int]:
        return {
            "loss": self.loss,
            "accuracy": self.accuracy,
            "samples_processed": self.samples_processed,
        }


class Callback(Protocol):
    def on_eopch_start(self, epoch: int) -> None: ...

    def on_epoch_end(self, epoch: int, metrics: TrainingMetrics) -> None: ...

    def on_batch_end(self, batch: int, metrics: TrainingMetrics) -> None: ...


def _fire_epoch_start(callback: object, epoch: int) -> None:
    """Call ``on_epoch_start`` if present, else the reference's ``on_eopch_start``."""
    hook = getattr(callback, "on_epoch_start", None) or getattr(callback, "on_eopch_start", None)
    if hook is not None:
        hook(epoch)


class BaseTrainer(ABC, Generic[M]):
    def __init__(
        self,
        config: TrainingConfig,
        callbacks: list[Callback] | None = None,
        eager_callbacks: bool = False,
    ) -> None:
        self._config = config
        self._callbacks = callbacks or []
        self._logger = Logger()
        self._device = torch.device(config.device)
        self._eager = eager_callbacks
        self.last_epoch_metrics: TrainingMetrics | None = None

    @property
    def config(self) -> TrainingConfig:
        return self._config

    @property
    def device(self) -> torch.device:
        return self._device

    @abstractmethod
    def compute_loss(self, output: torch.Tensor, target: torch.Tensor) -> torch.Tensor: ...

    @abstractmethod
    def compute_accuracy(self, output: torch.Tensor, target: torch.Tensor) -> float: ...

    def compute_accuracy_tensor(
        self, output: torch.Tensor, target: torch.Tensor
    ) -> torch.Tensor | None:
        """Device-resident accuracy (0-d tensor) or ``None`` if only the float form exists."""
        return None

    def _step(
        self,
        model: nn.Module,
        data: torch.Tensor,
        target: torch.Tensor,
        optimizer: torch.optim.Optimizer,
    ) -> tuple[torch.Tensor, torch.Tensor]:
        """One optimisation step; returns (loss, output).  Subclasses may override."""
        optimizer.zero_grad(set_to_none=True)
        output = model(data)
        loss = self.compute_loss(output, target)
        loss.backward()
        optimizer.step()
        return loss, output

    @log_exec
    def train_epoch(
        self,
        model: M,
        dataloader: DataLoader,
        optimizer: torch.optim.Optimizer,
        epoch: int,
    ) -> TrainingMetrics:
        module = cast(nn.Module, model)
        module.train()
        for cb in self._callbacks:
            _fire_epoch_start(cb, epoch)

        losses: list[torch.Tensor | float] = []
        accs: list[torch.Tensor | float] = []
        sizes: list[int] = []
        limit = self._config.max_batches

        for batch_idx, (data, target) in enumerate(dataloader):
            if limit is not None and batch_idx >= limit:
                break
            data = data.to(self._device, non_blocking=True)
            target = target.to(self._device, non_blocking=True)
            loss, output = self._step(module, data, target, optimizer)

            acc_t = self.compute_accuracy_tensor(output.detach(), target)
            acc: torch.Tensor | float = (
                acc_t if acc_t is not None else float(self.compute_accuracy(output, target))
            )
            losses.append(loss.detach())
            accs.append(acc)
            sizes.append(int(len(data)))
            if self._eager:
                self._emit_batch(dataloader, epoch, batch_idx, float(losses[-1]),
                                 float(accs[-1]), sum(sizes))

        if not sizes:
            raise ValueError("train_epoch received an empty dataloader")

        loss_host = _to_host(losses)
        acc_host = _to_host(accs)

        seen = 0
        last: TrainingMetrics | None = None
        for i, (l, a, n) in enumerate(zip(loss_host, acc_host, sizes)):
            seen += n
            last = TrainingMetrics(loss=l, accuracy=a, epoch=epoch, batch=i,
                                   samples_processed=seen)
            if not self._eager:
                self._emit_batch(dataloader, epoch, i, l, a, seen, metrics=last)

        count = float(len(sizes))
        averages = TrainingMetrics(
            loss=sum(loss_host) / count,
            accuracy=sum(acc_host) / count,
            epoch=epoch,
            batch=len(sizes) - 1,
            samples_processed=seen,
        )
        self.last_epoch_metrics = averages
        for cb in self._callbacks:
            cb.on_epoch_end(epoch, averages)
        assert last is not None
        return last

    def _emit_batch(
        self,
        dataloader: DataLoader,
        epoch: int,
        batch_idx: int,
        loss: float,
        acc: float,
        seen: int,
        metrics: TrainingMetrics | None = None,
    ) -> None:
        metrics = metrics or TrainingMetrics(loss, acc, epoch, batch_idx, seen)
        for cb in self._callbacks:
            cb.on_batch_end(batch_idx, metrics)
        if batch_idx % self._config.log_interval == 0:
            try:
                total = len(cast(Sized, dataloader.dataset))
            except (TypeError, AttributeError):
                total = seen
            pct = 100.0 * float(seen) / float(max(total, 1))
            self._logger.info(
                f"Train Epoch: {epoch} [{seen}/{total} ({pct:.0f}%)] "
                f"Loss: {loss:.6f} Accuracy: {acc:.4f}"
            )


def _to_host(values: list[torch.Tensor | float]) -> list[float]:
    """Fetch a list of 0-d device tensors with a single transfer."""
    if values and all(isinstance(v, torch.Tensor) for v in values):
        stacked = torch.stack([cast(torch.Tensor, v).float().reshape(()) for v in values])
        return [float(x) for x in stacked.cpu().tolist()]
    return [float(v) for v in values]

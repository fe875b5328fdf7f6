"""JSON metrics sink (reference ``trainer/callback.py:9-53``).

Batch records are buffered; the whole list is rewritten at each epoch end with
``"type": "epoch"`` / ``"type": "batch"`` records (``test_callback.py:9-46``).
"""

from __future__ import annotations

import json
from dataclasses import dataclass
from pathlib import Path

from nanofed_b200.utils import get_current_time

from .base import Callback, TrainingMetrics


@dataclass(slots=True)
class MetricsLogger(Callback):
    log_dir: Path
    experiment_name: str

    def __post_init__(self) -> None:
        self.log_dir.mkdir(parents=True, exist_ok=True)
        stamp = f"{get_current_time():%Y%m%d_%H%M%S}"
        self._log_file = self.log_dir / f"{self.experiment_name}_{stamp}.json"
        self._metrics: list[dict] = []

    def on_eopch_start(self, epoch: int) -> None:
        pass

    on_epoch_start = on_eopch_start

    def _record(self, kind: str, metrics: TrainingMetrics, **extra: int) -> dict:
        return {
            "type": kind,
            **extra,
            "loss": metrics.loss,
            "accuracy": metrics.accuracy,
            "samples_processed": metrics.samples_processed,
            "timestamp": get_current_time().isoformat(),
        }

    def on_epoch_end(self, epoch: int, metrics: TrainingMetrics) -> None:
        self._metrics.append(self._record("epoch", metrics, epoch=epoch))
        with open(self._log_file, "w") as fh:
            json.dump(self._metrics, fh, indent=2)

    def on_batch_end(self, batch: int, metrics: TrainingMetrics) -> None:
        self._metrics.append(self._record("batch", metrics, epoch=metrics.epoch, batch=batch))

"""Layer L2b: client-side training (reference ``nanofed/trainer/__init__.py:6-14``)."""

from .base import BaseTrainer, Callback, TrainingConfig, TrainingMetrics
from .callback import MetricsLogger
from .private import PrivateTrainer
from .torch import TorchTrainer

__all__ = [
    "BaseTrainer",
    "TrainingConfig",
    "TrainingMetrics",
    "Callback",
    "MetricsLogger",
    "TorchTrainer",
    "PrivateTrainer",
]

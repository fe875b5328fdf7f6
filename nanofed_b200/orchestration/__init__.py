"""Layer L4: round orchestration (reference ``nanofed/orchestration/__init__.py:5-13``)."""

from .coordinator import Coordinator, CoordinatorConfig
from .types import ClientInfo, RoundMetrics, RoundStatus, TrainingProgress
from .utils import coordinate

__all__ = [
    "Coordinator",
    "CoordinatorConfig",
    "ClientInfo",
    "RoundMetrics",
    "RoundStatus",
    "TrainingProgress",
    "coordinate",
]

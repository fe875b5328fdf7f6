"""nanofed_b200 — a Blackwell-native federated-learning framework.

Public façade: the ten names of the reference (``nanofed/__init__.py:12-23``) keep
their meaning, so ``from nanofed_b200 import Coordinator, HTTPServer, …`` is a
drop-in for ``from nanofed import …``.  New, B200-specific entry points:

* ``FedSimulation`` / ``CohortConfig`` (``parallel``): simulated clients packed as
  cohorts on the GPUs of one box, whole rounds executed by hand-written sm_100a
  kernels over a peer-memory arena;
* ``ops``: the kernel surface (fused FedAvg reduce, broadcast, stats, optimisers,
  cross-entropy, Philox RNG, grouped GEMM/conv);
* ``runtime``: flat parameter layouts, arenas, symmetric memory.
"""

from __future__ import annotations

__version__ = "0.1.0"

from nanofed_b200.communication import HTTPClient, HTTPServer
from nanofed_b200.orchestration import Coordinator, CoordinatorConfig, coordinate
from nanofed_b200.server import FedAvgAggregator, ModelManager
from nanofed_b200.trainer import PrivateTrainer, TorchTrainer, TrainingConfig

__all__ = [
    "HTTPClient",
    "HTTPServer",
    "TrainingConfig",
    "TorchTrainer",
    "PrivateTrainer",
    "Coordinator",
    "CoordinatorConfig",
    "FedAvgAggregator",
    "ModelManager",
    "coordinate",
    "FedSimulation",
    "CohortConfig",
]


def __getattr__(name: str):
    # Lazy: the parallel layer pulls in torch.distributed and the native library.
    if name in ("FedSimulation", "CohortConfig"):
        from nanofed_b200 import parallel

        return getattr(parallel, name)
    raise AttributeError(f"module 'nanofed_b200' has no attribute {name!r}")

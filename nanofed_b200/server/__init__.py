"""Layer L2a: server-side logic (reference ``nanofed/server/__init__.py:11-22``)."""

from .aggregator import (
    AggregationResult,
    BaseAggregator,
    FedAvgAggregator,
    PrivacyAwareAggregationConfig,
    PrivacyAwareAggregator,
)
from .fault_tolerance import (
    CheckpointMetadata,
    FaultTolerantCoordinator,
    FileStateStore,
    RoundState,
    SimpleRecoveryStrategy,
)
from .model_manager import ModelManager, ModelVersion
from .validation import (
    DefaultModelValidator,
    SecurityManager,
    ValidationConfig,
    ValidationResult,
)

__all__ = [
    "AggregationResult",
    "BaseAggregator",
    "FedAvgAggregator",
    "PrivacyAwareAggregator",
    "PrivacyAwareAggregationConfig",
    "ModelManager",
    "ModelVersion",
    "CheckpointMetadata",
    "FileStateStore",
    "RoundState",
    "SimpleRecoveryStrategy",
    "FaultTolerantCoordinator",
    "DefaultModelValidator",
    "SecurityManager",
    "ValidationConfig",
    "ValidationResult",
]

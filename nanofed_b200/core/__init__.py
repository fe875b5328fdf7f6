"""Layer L0: exceptions, protocols and data carriers (reference ``nanofed/core``)."""

from .exceptions import AggregationError, ModelManagerError, NanoFedError
from .interfaces import (
    AggregatorProtocol,
    AggregatorProtoocol,
    CoordinatorProtocol,
    ModelManagerProtocol,
    ModelProtocol,
    ServerProtocol,
    TrainerProtocol,
)
from .types import ModelUpdate, ModelVersion

__all__ = [
    "NanoFedError",
    "AggregationError",
    "ModelManagerError",
    "ModelProtocol",
    "AggregatorProtocol",
    "AggregatorProtoocol",
    "TrainerProtocol",
    "ServerProtocol",
    "ModelManagerProtocol",
    "CoordinatorProtocol",
    "ModelUpdate",
    "ModelVersion",
]

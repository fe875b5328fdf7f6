"""Structural (duck-typed) contracts between layers.

Parity: reference ``nanofed/core/interfaces.py:13-67``.  The reference exports
the misspelt ``AggregatorProtoocol``; both spellings are provided.  The
reference declares ``ModelManagerProtocol.list_versions`` as a property but
implements and calls it as a method (``manager.py:190``); it is a method here.
"""

from __future__ import annotations

from pathlib import Path
from typing import Any, Iterator, Protocol, TypeVar, runtime_checkable

import torch
from torch import nn
from torch.utils.data import DataLoader

from .types import ModelVersion

T = TypeVar("T", bound=nn.Module)


@runtime_checkable
class ModelProtocol(Protocol):
    """What the framework needs from a model object."""

    def forward(self, x: torch.Tensor) -> torch.Tensor: ...

    def parameters(self) -> Iterator[torch.nn.Parameter]: ...

    def state_dict(self) -> dict[str, torch.Tensor]: ...

    def load_state_dict(self, state_dict: dict[str, torch.Tensor]) -> Any: ...

    def to(self, device: str | torch.device) -> "ModelProtocol": ...


class AggregatorProtocol(Protocol[T]):
    """Combines a list of updates into one."""

    def aggregate(self, updates: list[T]) -> T: ...


# Public (misspelt) name of the reference, kept as an alias.
AggregatorProtoocol = AggregatorProtocol


class TrainerProtocol(Protocol[T]):
    """Local training driver."""

    def train(self, model: T, data: DataLoader) -> T: ...

    def validate(self, model: T, data: DataLoader) -> dict[str, float]: ...


class ModelManagerProtocol(Protocol):
    """Versioned store for the global model."""

    def set_dirs(self, models_dir: Path, configs_dir: Path) -> None: ...

    @property
    def current_version(self) -> Any: ...

    def load_model(self) -> Any: ...

    def save_model(
        self, config: dict[str, Any], metrics: dict[str, float] | None
    ) -> Any: ...

    def list_versions(self) -> list[ModelVersion]: ...

    @property
    def model(self) -> ModelProtocol: ...


class CoordinatorProtocol(Protocol):
    """Round driver as seen by the transport."""

    @property
    def model_manager(self) -> ModelManagerProtocol: ...


class ServerProtocol(Protocol):
    """Transport endpoint as seen by the coordinator."""

    @property
    def host(self) -> str: ...

    @property
    def port(self) -> int: ...

    @property
    def url(self) -> str: ...

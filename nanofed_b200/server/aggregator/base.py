"""Aggregator contract and shared validation.

Parity: reference ``server/aggregator/base.py:14-82`` — ``AggregationResult``
fields, ``current_round``, ``_validate_updates`` raising ``AggregationError``
with the messages ``test_fedavg.py:136-167`` matches on, abstract ``aggregate`` /
``_compute_weights``.

Added for the device path: ``_sample_counts`` (the integers FedAvg weights by),
``_pack`` (updates → one ``[K, N]`` slab via ``FlatLayout``) so subclasses hand a
single pointer to the fused reduction kernel instead of looping over keys.
"""

from __future__ import annotations

from abc import ABC, abstractmethod
from dataclasses import dataclass
from datetime import datetime
from typing import Generic, Sequence, TypeVar

import torch

from nanofed_b200.core.exceptions import AggregationError
from nanofed_b200.core.interfaces import ModelProtocol
from nanofed_b200.core.types import ModelUpdate
from nanofed_b200.runtime.flat import FlatLayout
from nanofed_b200.utils import Logger, get_current_time

T = TypeVar("T", bound=ModelProtocol)


@dataclass(slots=True, frozen=True)
class AggregationResult(Generic[T]):
    """Outcome of one aggregation."""

    model: T
    round_number: int
    num_clients: int
    timestamp: datetime
    metrics: dict[str, float]


def sample_count_of(update: ModelUpdate) -> float | None:
    """``num_samples`` or, failing that (or if it is 0), ``samples_processed``.

    The ``or`` fall-through on an explicit 0 is the reference's behaviour
    (``fedavg.py:108-110``).
    """
    metrics = update["metrics"]
    return metrics.get("num_samples") or metrics.get("samples_processed")


class BaseAggregator(ABC, Generic[T]):
    """Common state and checks for aggregation strategies."""

    def __init__(self) -> None:
        self._logger = Logger()
        self._current_round: int = 0
        self._weights_cache: dict[int, list[float]] = {}
        self._layout: FlatLayout | None = None

    @property
    def current_round(self) -> int:
        return self._current_round

    def _get_timestamp(self) -> datetime:
        return get_current_time()

    def _validate_updates(self, updates: Sequence[ModelUpdate]) -> None:
        if not updates:
            raise AggregationError("No updates provided for aggregation")
        rounds = {u["round_number"] for u in updates}
        if len(rounds) != 1:
            raise AggregationError(f"Updates from different rounds: {rounds}")
        names = updates[0]["model_state"].keys()
        if any(u["model_state"].keys() != names for u in updates[1:]):
            raise AggregationError("Inconsistent model architectures in updates.")

    def _sample_counts(self, updates: Sequence[ModelUpdate]) -> list[float]:
        counts: list[float] = []
        for u in updates:
            n = sample_count_of(u)
            if n is None:
                self._logger.warning(
                    f"Client {u['client_id']} did not report sample count. Using 1.0"
                )
                n = 1.0
            counts.append(n)
        return counts

    def _layout_for(self, model: ModelProtocol, first_state: dict) -> FlatLayout:
        """Layout keyed on the update's entry names; dtypes/shapes from the model."""
        reference = None
        try:
            reference = model.state_dict()
        except Exception:  # pragma: no cover - exotic model objects
            reference = None
        if reference is not None and reference.keys() == first_state.keys():
            source = reference
        else:
            source = first_state
        layout = self._layout
        if layout is None or layout.names != tuple(source.keys()) or any(
            layout[n].numel != (v.numel() if isinstance(v, torch.Tensor) else torch.as_tensor(v).numel())
            for n, v in source.items()
        ):
            layout = FlatLayout.from_state(source)
            self._layout = layout
        return layout

    @staticmethod
    def _device_of(model: ModelProtocol, updates: Sequence[ModelUpdate]) -> torch.device:
        """Aggregate where the updates already live; CPU lists go to the model's device."""
        for u in updates:
            for v in u["model_state"].values():
                if isinstance(v, torch.Tensor):
                    return v.device
                break
        try:
            return next(iter(model.parameters())).device
        except (StopIteration, AttributeError, TypeError):
            return torch.device("cpu")

    def _pack(
        self, model: ModelProtocol, updates: Sequence[ModelUpdate]
    ) -> tuple[FlatLayout, torch.Tensor]:
        layout = self._layout_for(model, updates[0]["model_state"])
        device = self._device_of(model, updates)
        slab = layout.stack((u["model_state"] for u in updates), device=device)
        return layout, slab

    @abstractmethod
    def aggregate(self, model: T, updates: Sequence[ModelUpdate]) -> AggregationResult[T]: ...

    @abstractmethod
    def _compute_weights(self, updates: Sequence[ModelUpdate]) -> list[float]: ...

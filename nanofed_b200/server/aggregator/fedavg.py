"""Sample-weighted federated averaging.

Parity: reference ``server/aggregator/fedavg.py:10-125`` — weights
``n_k / Σ n`` (missing counts ⇒ 1.0 each, with a warning), every ``state_dict``
entry averaged (incl. BatchNorm buffers; integers are cast back on load),
updates may carry tensors **or nested lists** (``test_fedavg.py:26-54``),
metrics are the weighted mean of numeric client metrics, ``current_round``
advances by one per call.

B200 path: the K updates are packed as rows of one ``[K, N]`` fp32 slab and
reduced by ONE launch of the fused K-AGG kernel (``ops.fedavg_reduce`` →
``csrc/fedavg_reduce.cu``): weights are normalised on the device from the raw
sample counts, accumulation is fp32, and the result lands in a flat vector that
is scattered back into the model.  The reference instead runs ``2·K·|keys|``
small ``mul``/``add_`` ops with two temporaries each (``fedavg.py:56-63``).
When the cohort engine is used (``parallel.cohort``), client states already
live in the arena and no packing happens at all.
"""

from __future__ import annotations

from typing import Sequence

import torch

from nanofed_b200 import ops
from nanofed_b200.core import ModelProtocol, ModelUpdate
from nanofed_b200.utils import get_current_time, log_exec

from .base import AggregationResult, BaseAggregator


class FedAvgAggregator(BaseAggregator[ModelProtocol]):
    """FedAvg: ``global = Σ_k (n_k / Σ n) · client_k``."""

    def _to_tensor(self, data: list | float | torch.Tensor) -> torch.Tensor:
        if isinstance(data, torch.Tensor):
            return data.detach()
        return torch.tensor(data, dtype=torch.float32)

    @log_exec
    def aggregate(
        self, model: ModelProtocol, updates: Sequence[ModelUpdate]
    ) -> AggregationResult[ModelProtocol]:
        self._validate_updates(updates)
        counts = self._sample_counts(updates)
        total = float(sum(counts))
        weights = [c / total for c in counts]

        layout, slab = self._pack(model, updates)
        counts_t = torch.tensor(counts, dtype=torch.float32, device=slab.device)
        flat = ops.fedavg_reduce(slab, counts_t)
        model.load_state_dict(layout.unflatten(flat, clone=False))

        metrics = self._aggregate_metrics(updates, weights)
        self._current_round += 1
        return AggregationResult(
            model=model,
            round_number=self._current_round,
            num_clients=len(updates),
            timestamp=get_current_time(),
            metrics=metrics,
        )

    def _aggregate_metrics(
        self, updates: Sequence[ModelUpdate], weights: list[float]
    ) -> dict[str, float]:
        """Weighted mean of each numeric metric over the clients that report it."""
        num: dict[str, float] = {}
        den: dict[str, float] = {}
        for update, w in zip(updates, weights):
            for name, value in update["metrics"].items():
                if isinstance(value, (int, float)):
                    num[name] = num.get(name, 0.0) + float(value) * w
                    den[name] = den.get(name, 0.0) + w
        return {name: num[name] / den[name] for name in num}

    def _compute_weights(self, updates: Sequence[ModelUpdate]) -> list[float]:
        counts = self._sample_counts(updates)
        total = sum(counts)
        weights = [c / total for c in counts]
        self._logger.debug(f"Client sample counts: {counts}")
        self._logger.debug(f"Computed weights: {weights}")
        return weights

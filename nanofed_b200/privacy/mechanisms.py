"""Clip-then-noise mechanisms applied to a whole model state.

Parity: reference ``privacy/mechanisms.py:17-174``.

* ``_clip_update``: global L2 norm over all tensors, coefficient
  ``min(C / (‖x‖ + 1e-6), 1)``, returns the scaled dict and ``UpdateMetadata``
  (``:85-104``; ``test_mechanism.py:101-130`` checks the clipped norm ≤ C+1e-6).
* ``add_noise``: clip to ``max_gradient_norm``; one generator call per tensor
  with ``scale = σ·C / batch_size``; one accountant event
  ``(sigma=σ, samples=batch_size)`` (``:106-129``; call counts pinned by
  ``test_mechanism.py:133-200``).
* ``LocalPrivacyMechanism`` ignores the caller's batch size and uses 1 (``:148-158``).
* ``PrivacyMechanismFactory.create`` raises ``ValueError("Unknown privacy type: …")``.

B200 path: the norm no longer materialises ``cat(flatten(all))`` (the reference
copies the whole model per client).  CUDA tensors go through the fused
sum-of-squares kernel (``ops.sumsq``, K-STATS); CPU tensors accumulate per-tensor
squared norms.  Inside the fused aggregation (``ops.fedavg_reduce``, K-AGG) the
same clip coefficient is consumed as a per-client scalar and the K noise draws
fold into one, so this class is the *semantic oracle* for that
kernel as well as the stand-alone API.
"""

from __future__ import annotations

import math
from abc import ABC, abstractmethod
from dataclasses import dataclass
from enum import Enum, auto
from typing import Any, Protocol, TypeAlias, TypedDict

import torch

from nanofed_b200.utils.logger import Logger

from .accountant import GaussianAccountant, PrivacySpent
from .config import NoiseType, PrivacyConfig
from .noise import GaussianNoiseGenerator, LaplacianNoiseGenerator

ModelState: TypeAlias = dict[str, torch.Tensor]


class PrivacyType(Enum):
    """Where the noise is added: at the server (CENTRAL) or by each client (LOCAL)."""

    CENTRAL = auto()
    LOCAL = auto()


class PrivacyMetrics(TypedDict):
    epsilon_spent: float
    delta_spent: float
    noise_scale: float
    clip_ratio: float


class PrivacyMechanism(Protocol):
    def add_noise(self, parameters: ModelState, batch_size: int) -> ModelState: ...

    def get_privacy_spent(self) -> PrivacySpent: ...

    @property
    def privacy_type(self) -> PrivacyType: ...


@dataclass(slots=True, frozen=True)
class UpdateMetadata:
    """What clipping did to one update."""

    total_norm: float
    clipped_norm: float
    num_parameters: int
    noise_scale: float


def global_l2_norm(parameters: ModelState) -> torch.Tensor:
    """‖concat(all tensors)‖₂ as a 0-d fp32 tensor, without concatenating."""
    tensors = [p for p in parameters.values() if p.numel() > 0]
    if not tensors:
        return torch.zeros(())
    if all(t.is_cuda for t in tensors):
        from nanofed_b200 import ops

        return ops.sumsq_multi(tensors).sqrt()
    total = torch.zeros((), dtype=torch.float32)
    for t in tensors:
        total = total + t.detach().to(torch.float32).pow(2).sum().cpu()
    return total.sqrt()


class BasePrivacyMechanism(ABC):
    """Shared clip + noise + accounting logic."""

    def __init__(
        self,
        config: PrivacyConfig,
        accountant: GaussianAccountant | None = None,
        noise_generator: GaussianNoiseGenerator | None = None,
    ) -> None:
        self._config = config
        self._accountant = accountant or GaussianAccountant(config)
        if noise_generator is None:
            noise_generator = (
                LaplacianNoiseGenerator()
                if config.noise_type == NoiseType.LAPLACIAN
                else GaussianNoiseGenerator()
            )
        self._noise_gen = noise_generator
        self._logger = Logger()

    @property
    @abstractmethod
    def privacy_type(self) -> PrivacyType: ...

    @property
    def config(self) -> PrivacyConfig:
        return self._config

    @property
    def accountant(self) -> GaussianAccountant:
        return self._accountant

    def _compute_noise_scale(self, batch_size: int) -> float:
        """σ·C / batch_size — the per-coordinate standard deviation."""
        return self._config.noise_multiplier * self._config.max_gradient_norm / batch_size

    def clip_coefficient(self, total_norm: float, max_norm: float | None = None) -> float:
        """``min(C / (‖x‖ + 1e-6), 1)`` — the scalar K-AGG applies per client."""
        bound = self._config.max_gradient_norm if max_norm is None else max_norm
        return min(bound / (total_norm + 1e-6), 1.0)

    def _clip_update(
        self, parameters: ModelState, max_norm: float
    ) -> tuple[ModelState, UpdateMetadata]:
        norm = float(global_l2_norm(parameters))
        coef = self.clip_coefficient(norm, max_norm)
        scaled = {name: p * coef for name, p in parameters.items()}
        meta = UpdateMetadata(
            total_norm=norm,
            clipped_norm=norm * coef,
            num_parameters=sum(p.numel() for p in parameters.values()),
            noise_scale=self._config.noise_multiplier,
        )
        return scaled, meta

    def add_noise(self, parameters: ModelState, batch_size: int) -> ModelState:
        """Return ``clip(parameters) + noise`` and charge one event to the accountant."""
        clipped, meta = self._clip_update(parameters, self._config.max_gradient_norm)
        std = self._compute_noise_scale(batch_size)
        out: ModelState = {}
        for name, value in clipped.items():
            draw = self._noise_gen.generate(tuple(value.shape), scale=std)
            if isinstance(draw, torch.Tensor) and draw.device != value.device:
                draw = draw.to(value.device)
            if not value.is_floating_point():
                # Integer state (e.g. BatchNorm counters): promote, like ``int + float``.
                value = value.to(torch.float32)
            out[name] = value + draw
        self._accountant.add_noise_event(
            sigma=self._config.noise_multiplier, samples=batch_size
        )
        self._logger.debug(
            "Applied privacy mechanism: "
            f"norm={meta.total_norm:.3f}->{meta.clipped_norm:.3f}, noise={std:.3f}"
        )
        return out

    def get_privacy_spent(self) -> PrivacySpent:
        return self._accountant.get_privacy_spent()

    def validate_budget(self) -> bool:
        return self._accountant.validate_budget()

    def folded_noise_std(self, weights: list[float]) -> float:
        """Std of the single draw equivalent to K weighted per-client draws.

        ``Σ_k w_k ξ_k`` with i.i.d. ``ξ_k ~ N(0, s²)`` is ``N(0, s² Σ w_k²)``
        (SURVEY §3.4); K-AGG draws that once instead of K times.
        """
        per_client = self._compute_noise_scale(self._effective_batch(len(weights)))
        return per_client * math.sqrt(sum(w * w for w in weights))

    def _effective_batch(self, batch_size: int) -> int:
        return batch_size


class CentralPrivacyMechanism(BasePrivacyMechanism):
    """Server-side DP: noise scaled down by the number of aggregated updates."""

    @property
    def privacy_type(self) -> PrivacyType:
        return PrivacyType.CENTRAL


class LocalPrivacyMechanism(BasePrivacyMechanism):
    """Client-side DP: every update is its own release, so batch size is 1."""

    @property
    def privacy_type(self) -> PrivacyType:
        return PrivacyType.LOCAL

    def _effective_batch(self, batch_size: int) -> int:
        return 1

    def add_noise(self, parameters: ModelState, batch_size: int) -> ModelState:
        return super().add_noise(parameters, batch_size=1)


class PrivacyMechanismFactory:
    """Builds the mechanism for a ``PrivacyType``."""

    @staticmethod
    def create(
        privacy_type: PrivacyType, config: PrivacyConfig, **kwargs: Any
    ) -> BasePrivacyMechanism:
        if privacy_type == PrivacyType.CENTRAL:
            return CentralPrivacyMechanism(config, **kwargs)
        if privacy_type == PrivacyType.LOCAL:
            return LocalPrivacyMechanism(config, **kwargs)
        raise ValueError(f"Unknown privacy type: {privacy_type}")

"""Budget bookkeeping primitives (reference ``privacy/accountant/base.py:8-53``)."""

from __future__ import annotations

from abc import ABC, abstractmethod
from dataclasses import dataclass
from typing import Protocol

from ..config import PrivacyConfig


@dataclass(frozen=True)
class PrivacySpent:
    """Cumulative ``(epsilon, delta)`` consumed so far."""

    epsilon_spent: float
    delta_spent: float

    def validate(self, config: PrivacyConfig) -> bool:
        """True while both components are within ``config``'s budget."""
        return self.epsilon_spent <= config.epsilon and self.delta_spent <= config.delta


class PrivacyAccountant(Protocol):
    def get_privacy_spent(self) -> PrivacySpent: ...

    def add_noise_event(self, sigma: float, samples: int) -> None: ...

    def validate_budget(self, config: PrivacyConfig) -> bool: ...


def check_event(sigma: float, samples: int) -> None:
    """Shared argument checks; messages pinned by ``test_gaussian.py:84-101``."""
    if samples <= 0:
        raise ValueError("Number of samples must be positive")
    if sigma <= 0:
        raise ValueError("Noise multiplier must be positive")


class BasePrivacyAccountant(ABC):
    """Holds the config and the event count; subclasses define the composition rule."""

    def __init__(self, config: PrivacyConfig) -> None:
        self._config = config
        self._privacy_spent = PrivacySpent(0.0, 0.0)
        self._event_count = 0

    def _sampling_rate(self, samples: int) -> float:
        """``q = min(samples / max_gradient_norm, 1)``.

        Dimensionally odd, but it is the reference's contract and its tests
        assert the resulting epsilons to 1e-9
        (``test_privacy_properties.py:114-142``, ``test_privacy_edge_cases.py:60-92``).
        """
        return min(float(samples) / float(self._config.max_gradient_norm), 1.0)

    @abstractmethod
    def _compute_privacy_spent(self) -> PrivacySpent: ...

    def get_privacy_spent(self) -> PrivacySpent:
        return self._compute_privacy_spent()

    def validate_budget(self, config: PrivacyConfig | None = None) -> bool:
        return bool(self.get_privacy_spent().validate(config or self._config))

"""Basic-composition accountant for the Gaussian mechanism.

Parity: reference ``privacy/accountant/gaussian.py:7-48``: every event costs
``sqrt(2 ln(1.25/δ)) · q / σ`` and the costs add; δ_spent is the configured δ as
soon as one event exists, ``(0, 0)`` before.
"""

from __future__ import annotations

import math

from ..config import PrivacyConfig
from .base import BasePrivacyAccountant, PrivacySpent, check_event


class GaussianAccountant(BasePrivacyAccountant):
    def __init__(self, config: PrivacyConfig) -> None:
        super().__init__(config)
        self._c = math.sqrt(2.0 * math.log(1.25 / config.delta))
        self._noise_multipliers: list[float] = []
        self._sample_rates: list[float] = []
        self._q = 0.0

    def add_noise_event(self, sigma: float, samples: int) -> None:
        check_event(sigma, samples)
        self._q = self._sampling_rate(samples)
        self._noise_multipliers.append(float(sigma))
        self._sample_rates.append(self._q)
        self._event_count += 1
        self._compute_privacy_spent()

    def add_noise_events(self, sigma: float, samples: int, count: int) -> None:
        """Record ``count`` identical events (one per client of a DP round)."""
        for _ in range(count):
            self.add_noise_event(sigma, samples)

    def _compute_privacy_spent(self) -> PrivacySpent:
        if not self._noise_multipliers:
            self._privacy_spent = PrivacySpent(0.0, 0.0)
        else:
            # Left-to-right sum in event order: bitwise equal to the reference.
            eps = sum(
                [self._c * q / s for s, q in zip(self._noise_multipliers, self._sample_rates)]
            )
            self._privacy_spent = PrivacySpent(eps, self._config.delta)
        return self._privacy_spent

"""Rényi-DP accountant (Mironov 2017 conversion).

Parity: reference ``privacy/accountant/rdp.py:11-115``: default orders
``[1.5, 2, 2.5, 3, 4, 8, 16, 32, 64]``; each event adds ``q² α / (2 σ²)`` to the
order-α budget; ``ε = min_α (rdp_α + ln(1/δ)/(α-1))``; orders must be ``> 1``
else ``PrivacyError`` (``test_rdp.py:18-141``).
"""

from __future__ import annotations

import math
from typing import Sequence

import numpy as np

from ..config import PrivacyConfig
from ..exceptions import PrivacyError
from .base import BasePrivacyAccountant, PrivacySpent, check_event

_DEFAULT_ORDERS = (1.5, 2.0, 2.5, 3.0, 4.0, 8.0, 16.0, 32.0, 64.0)


class RDPAccountant(BasePrivacyAccountant):
    def __init__(self, config: PrivacyConfig, orders: Sequence[float] | None = None) -> None:
        super().__init__(config)
        self._orders = np.array(list(orders) if orders else list(_DEFAULT_ORDERS))
        if self._orders.size == 0:
            raise PrivacyError("Must specify at least one RDP order")
        if not bool(np.all(self._orders > 1.0)):
            raise PrivacyError("All RDP orders must be > 1.0")
        self._rdp_budget: dict[float, float] = {alpha: 0.0 for alpha in self._orders}

    def _compute_rdp_gaussian(self, sigma: float, sampling_rate: float) -> dict[float, float]:
        """Per-order RDP of one (sub-sampled) Gaussian release."""
        return {
            alpha: (sampling_rate**2) * alpha / (2 * sigma**2) for alpha in self._orders
        }

    def add_noise_event(self, sigma: float, samples: int) -> None:
        check_event(sigma, samples)
        step = self._compute_rdp_gaussian(sigma, self._sampling_rate(samples))
        for alpha in self._orders:
            self._rdp_budget[alpha] += step[alpha]
        self._event_count += 1
        self._compute_privacy_spent()

    def _compute_privacy_spent(self) -> PrivacySpent:
        delta = self._config.delta
        log_inv_delta = math.log(1 / delta)
        best = float("inf")
        for alpha in self._orders:
            best = min(best, self._rdp_budget[alpha] + log_inv_delta / (alpha - 1))
        self._privacy_spent = PrivacySpent(epsilon_spent=best, delta_spent=delta)
        return self._privacy_spent

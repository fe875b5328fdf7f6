"""Plain data carriers used across layers.

Parity: reference ``nanofed/core/types.py:11-29``.  ``ModelUpdate`` keeps the
six keys of the reference TypedDict; ``privacy_spent`` is optional here
(``total=False`` semantics via ``NotRequired``) because the reference server
never stores it and its coordinator crashes on the lookup
(``coordinator.py:319`` vs ``server.py:248-257``).
"""

from __future__ import annotations

from dataclasses import dataclass
from datetime import datetime
from pathlib import Path
from typing import Any, NotRequired, TypedDict

import torch

from nanofed_b200.privacy.accountant.base import PrivacySpent  # noqa: F401  (reference re-export)


class ModelUpdate(TypedDict):
    """One client's contribution to a round."""

    model_state: dict[str, torch.Tensor]
    client_id: str
    round_number: int
    metrics: dict[str, float]
    timestamp: datetime
    privacy_spent: NotRequired[Any]


@dataclass(slots=True, frozen=True)
class ModelVersion:
    """Handle to one saved global-model checkpoint."""

    version_id: str
    timestamp: datetime
    config: dict[str, Any]
    path: Path

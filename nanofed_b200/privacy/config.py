"""Validated, immutable privacy settings.

Parity: reference ``privacy/config.py:24-85`` — same five fields, same defaults,
inclusive bounds ``epsilon∈[0.01,10]``, ``delta∈[1e-10,0.1]``, strictly
positive ``max_gradient_norm`` / ``noise_multiplier``; assignment after
construction raises pydantic's ``ValidationError`` (``test_config.py:15-56``).
Unlike the reference, ``noise_type`` is honoured: mechanisms built without an
explicit generator pick Gaussian or Laplacian from it.
"""

from __future__ import annotations

from enum import Enum, auto

from pydantic import BaseModel, ConfigDict, Field

from . import constants as C


class NoiseType(Enum):
    """Family of the additive noise."""

    GAUSSIAN = auto()
    LAPLACIAN = auto()


class PrivacyConfig(BaseModel):
    """Budget ``(epsilon, delta)``, clip norm ``C`` and noise multiplier ``sigma``."""

    model_config = ConfigDict(frozen=True)

    epsilon: float = Field(
        default=C.DEFAULT_EPSILON, ge=C.MIN_EPSILON, le=C.MAX_EPSILON,
        description="Privacy parameter epsilon (ε)",
    )
    delta: float = Field(
        default=C.DEFAULT_DELTA, ge=C.MIN_DELTA, le=C.MAX_DELTA,
        description="Privacy parameter delta (δ)",
    )
    max_gradient_norm: float = Field(
        default=C.DEFAULT_MAX_GRAD_NORM, gt=0,
        description="L2 bound applied by clipping",
    )
    noise_multiplier: float = Field(
        default=C.DEFAULT_NOISE_MULTIPLIER, gt=0,
        description="Noise standard deviation in units of the clip norm",
    )
    noise_type: NoiseType = Field(
        default=NoiseType.GAUSSIAN, description="Noise distribution"
    )

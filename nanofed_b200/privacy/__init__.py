"""Layer L1: differential privacy (reference ``nanofed/privacy/__init__.py:6-15``)."""

from .config import NoiseType, PrivacyConfig
from .constants import DEFAULT_DELTA, DEFAULT_EPSILON
from .exceptions import (
    NoiseGenerationError,
    PrivacyBudgetExceededError,
    PrivacyConfigurationError,
    PrivacyError,
)
from .noise import GaussianNoiseGenerator, LaplacianNoiseGenerator

__all__ = [
    "NoiseType",
    "PrivacyConfig",
    "DEFAULT_DELTA",
    "DEFAULT_EPSILON",
    "PrivacyError",
    "PrivacyBudgetExceededError",
    "PrivacyConfigurationError",
    "NoiseGenerationError",
    "GaussianNoiseGenerator",
    "LaplacianNoiseGenerator",
]

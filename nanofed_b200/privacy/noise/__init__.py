from .base import BaseNoiseGenerator, NoiseGenerator
from .generators import GaussianNoiseGenerator, LaplacianNoiseGenerator

__all__ = [
    "GaussianNoiseGenerator",
    "LaplacianNoiseGenerator",
    "BaseNoiseGenerator",
    "NoiseGenerator",
]

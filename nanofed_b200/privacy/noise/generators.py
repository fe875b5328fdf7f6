"""Gaussian and Laplacian samplers.

Parity: reference ``privacy/noise/generators.py:14-67`` — same argument
checks and messages (``ValueError`` for a non-tuple / non-positive shape or a
non-positive scale; sampling failures re-raised as ``NoiseGenerationError``),
seeded reproducibility, ``std ≈ scale`` (Gaussian) and ``E|x| ≈ scale``
(Laplace) as pinned by ``tests/unit/privacy/noise/test_generators.py:25-78``.

Deliberate fix: the empty tuple ``()`` — the shape of a 0-d tensor
such as BatchNorm's ``num_batches_tracked`` — is accepted.  The reference
rejects it with "Shape must be provided", which makes its DP aggregator unusable
on any BatchNorm model.  ``None`` is still rejected with that message.
"""

from __future__ import annotations

from functools import wraps
from typing import Callable, ParamSpec, TypeVar

import torch

from ..exceptions import NoiseGenerationError
from ..types import Shape, Tensor
from .base import BaseNoiseGenerator

P = ParamSpec("P")
T = TypeVar("T", bound=torch.Tensor)


def validate_noise_input(func: Callable[P, T]) -> Callable[P, T]:
    """Check ``(shape, scale)`` before sampling; wrap sampler failures."""

    @wraps(func)
    def checked(*args: P.args, **kwargs: P.kwargs) -> T:
        shape = args[1] if len(args) > 1 else kwargs.get("shape")
        scale = args[2] if len(args) > 2 else kwargs.get("scale")
        if isinstance(shape, torch.Size):
            shape = tuple(shape)
            if len(args) > 1:
                args = (args[0], shape, *args[2:])  # type: ignore[assignment]
            else:
                kwargs["shape"] = shape
        if shape is None or (not isinstance(shape, tuple) and not shape):
            raise ValueError("Shape must be provided")
        if not isinstance(shape, tuple):
            raise ValueError("Shape must be a tuple")
        for extent in shape:
            if not isinstance(extent, int) or isinstance(extent, bool) or extent <= 0:
                raise ValueError("Invalid shape: must be a tuple of positive integers")
        if isinstance(scale, bool) or not isinstance(scale, (int, float)):
            raise ValueError("Scale must be a number")
        if scale <= 0:
            raise ValueError("Scale must be positive")
        try:
            return func(*args, **kwargs)
        except Exception as exc:  # pragma: no cover - defensive
            raise NoiseGenerationError(f"Noise generation failed: {exc}") from exc

    return checked


def _numel(shape: Shape) -> int:
    n = 1
    for extent in shape:
        n *= extent
    return n


class GaussianNoiseGenerator(BaseNoiseGenerator):
    """Samples ``N(0, scale²)``."""

    @validate_noise_input
    def generate(self, shape: Shape, scale: float) -> Tensor:
        if self._device.type == "cuda":
            from nanofed_b200 import ops

            out = torch.empty(shape, dtype=torch.float32, device=self._device)
            seed, offset = self._advance(_numel(shape))
            ops.philox_normal_(out, float(scale), seed, offset)
            return out
        return torch.randn(shape, generator=self._rng) * scale


class LaplacianNoiseGenerator(BaseNoiseGenerator):
    """Samples ``Laplace(0, scale)`` by inverting the CDF of a uniform draw."""

    @validate_noise_input
    def generate(self, shape: Shape, scale: float) -> Tensor:
        if self._device.type == "cuda":
            from nanofed_b200 import ops

            out = torch.empty(shape, dtype=torch.float32, device=self._device)
            seed, offset = self._advance(_numel(shape))
            ops.philox_laplace_(out, float(scale), seed, offset)
            return out
        u = torch.rand(shape, generator=self._rng) - 0.5
        # x = -b * sgn(u) * ln(1 - 2|u|); clamp keeps the log finite when |u| = 0.5.
        magnitude = -torch.log1p((-2.0 * u.abs()).clamp_min(-1.0 + 1e-12))
        return torch.sign(u) * magnitude * scale

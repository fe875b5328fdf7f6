"""Noise-generator contract (reference ``privacy/noise/base.py:9-31``)."""

from __future__ import annotations

from abc import ABC, abstractmethod
from typing import Protocol

import torch

from ..types import Shape, Tensor


class NoiseGenerator(Protocol):
    def generate(self, shape: Shape, scale: float) -> Tensor: ...

    def set_seed(self, seed: int) -> None: ...


class BaseNoiseGenerator(ABC):
    """Seedable sampler.

    ``device="cpu"`` (default) reproduces the reference: a CPU
    ``torch.Generator`` feeds ``torch.randn``/``torch.rand``.  With a CUDA
    device the samples come from this framework's counter-based Philox kernel
    (``csrc/rng.cu``, op ``ops.philox_normal_`` / ``ops.philox_laplace_``): the
    stream is a pure function of ``(seed, offset)``, the offset advances by
    the number of elements drawn, and nothing crosses the PCIe bus.
    """

    def __init__(self, seed: int | None = None, device: str | torch.device = "cpu") -> None:
        self._device = torch.device(device)
        self._rng = torch.Generator()
        self._seed = 0
        self._offset = 0
        if seed is not None:
            self.set_seed(seed)
        else:
            self._seed = int(self._rng.seed() & 0x7FFFFFFFFFFFFFFF)

    @property
    def device(self) -> torch.device:
        return self._device

    def set_seed(self, seed: int) -> None:
        """Restart the stream from ``seed``."""
        self._rng.manual_seed(seed)
        self._seed = int(seed)
        self._offset = 0

    def _advance(self, count: int) -> tuple[int, int]:
        """Reserve ``count`` Philox outputs; returns the (seed, offset) to use."""
        at = self._offset
        self._offset += (count + 3) // 4 * 4
        return self._seed, at

    @abstractmethod
    def generate(self, shape: Shape, scale: float) -> Tensor: ...

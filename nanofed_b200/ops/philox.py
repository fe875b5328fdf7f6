"""Host replica of the device RNG (K-RNG): Philox4x32-10 + Box–Muller / inverse CDF.

``csrc/philox.cuh`` implements exactly this; the replica lets CPU tests check the
CUDA noise *bit-for-bit at the integer level* and to ~1 ulp after the float
transforms, instead of only statistically.

Stream definition (shared by host and device): element ``i`` of a draw made at
``(seed, offset)`` uses Philox counter ``((offset+i) // 4, 0)`` (64-bit low/high
split over the first two 32-bit words; words 2,3 are zero), key
``(seed & 0xffffffff, seed >> 32)`` and lane ``(offset+i) % 4``.  Uniforms are
``((x >> 8) + 0.5) · 2⁻²⁴ ∈ (0,1)``.  Normals: lanes (0,1) and (2,3) are the
(cos, sin) pair of a Box–Muller transform of uniforms (0,1) and (2,3).

Replaces the CPU ``torch.Generator`` of the reference's noise generators
(``nanofed/privacy/noise/base.py:16-31``, ``generators.py:49-67``): seeded, reproducible, and
available on the device that holds the gradients.
"""

from __future__ import annotations

import numpy as np
import torch

_M0 = np.uint64(0xD2511F53)
_M1 = np.uint64(0xCD9E8D57)
_W0 = np.uint32(0x9E3779B9)
_W1 = np.uint32(0xBB67AE85)
_MASK32 = np.uint64(0xFFFFFFFF)


def philox4x32(counter_lo: np.ndarray, counter_hi: np.ndarray, seed: int) -> np.ndarray:
    """Philox4x32-10 on an array of 64-bit counters → ``[n, 4]`` uint32."""
    c0 = counter_lo.astype(np.uint32)
    c1 = counter_hi.astype(np.uint32)
    c2 = np.zeros_like(c0)
    c3 = np.zeros_like(c0)
    k0 = np.uint32(seed & 0xFFFFFFFF)
    k1 = np.uint32((seed >> 32) & 0xFFFFFFFF)
    with np.errstate(over="ignore"):
        for _ in range(10):
            p0 = _M0 * c0.astype(np.uint64)
            p1 = _M1 * c2.astype(np.uint64)
            hi0 = (p0 >> np.uint64(32)).astype(np.uint32)
            lo0 = (p0 & _MASK32).astype(np.uint32)
            hi1 = (p1 >> np.uint64(32)).astype(np.uint32)
            lo1 = (p1 & _MASK32).astype(np.uint32)
            c0, c1, c2, c3 = hi1 ^ c1 ^ k0, lo1, hi0 ^ c3 ^ k1, lo0
            k0 = np.uint32((int(k0) + int(_W0)) & 0xFFFFFFFF)
            k1 = np.uint32((int(k1) + int(_W1)) & 0xFFFFFFFF)
    return np.stack([c0, c1, c2, c3], axis=1)


def _raw(count: int, seed: int, offset: int) -> np.ndarray:
    first = offset // 4
    last = (offset + count + 3) // 4
    groups = np.arange(first, last, dtype=np.uint64)
    words = philox4x32(groups & _MASK32, groups >> np.uint64(32), seed).reshape(-1)
    start = offset - first * 4
    return words[start : start + count], words, start


def uniform(count: int, seed: int, offset: int = 0) -> torch.Tensor:
    words, _, _ = _raw(count, seed, offset)
    u = ((words >> np.uint32(8)).astype(np.float32) + np.float32(0.5)) * np.float32(2.0**-24)
    return torch.from_numpy(u.astype(np.float32))


def normal(count: int, seed: int, offset: int = 0) -> torch.Tensor:
    """Standard normals for stream positions ``offset … offset+count``."""
    _, words, start = _raw(count, seed, offset)
    u = ((words >> np.uint32(8)).astype(np.float32) + np.float32(0.5)) * np.float32(2.0**-24)
    u = u.reshape(-1, 2)
    r = np.sqrt(np.float32(-2.0) * np.log(u[:, 0]))
    theta = np.float32(2.0 * np.pi) * u[:, 1]
    z = np.stack([r * np.cos(theta), r * np.sin(theta)], axis=1).reshape(-1).astype(np.float32)
    return torch.from_numpy(z[start : start + count].copy())


def laplace(count: int, seed: int, offset: int = 0) -> torch.Tensor:
    """Unit-scale Laplace samples (inverse CDF of the same uniforms)."""
    u = uniform(count, seed, offset).numpy() - np.float32(0.5)
    x = -np.sign(u) * np.log1p(np.maximum(np.float32(-2.0) * np.abs(u), np.float32(-1 + 1e-7)))
    return torch.from_numpy(x.astype(np.float32))

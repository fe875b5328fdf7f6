"""Device-side timing and NVTX ranges (no reference counterpart).

The reference's only profiling aid is ``log_exec`` (host wall clock).  Round
phases here run asynchronously on CUDA streams, so they are bracketed with CUDA
events and read once, after the round, with a single synchronise.
"""

from __future__ import annotations

import time
from contextlib import contextmanager
from typing import Iterator

import torch


class PhaseTimer:
    """Collects named (start, end) CUDA-event pairs; falls back to host clock on CPU."""

    def __init__(self, enabled: bool = True) -> None:
        self.enabled = enabled
        self._cuda = torch.cuda.is_available()
        self._pairs: list[tuple[str, object, object]] = []
        self._host: list[tuple[str, float]] = []

    @contextmanager
    def phase(self, name: str, stream: "torch.cuda.Stream | None" = None) -> Iterator[None]:
        if not self.enabled:
            yield
            return
        if self._cuda:
            a = torch.cuda.Event(enable_timing=True)
            b = torch.cuda.Event(enable_timing=True)
            torch.cuda.nvtx.range_push(name)
            a.record(stream)
            try:
                yield
            finally:
                b.record(stream)
                torch.cuda.nvtx.range_pop()
                self._pairs.append((name, a, b))
        else:
            t0 = time.perf_counter()
            try:
                yield
            finally:
                self._host.append((name, (time.perf_counter() - t0) * 1e3))

    def summary_ms(self) -> dict[str, float]:
        """Total milliseconds per phase name.  Synchronises the device once."""
        out: dict[str, float] = {}
        if self._cuda and self._pairs:
            torch.cuda.synchronize()
            for name, a, b in self._pairs:
                out[name] = out.get(name, 0.0) + a.elapsed_time(b)  # type: ignore[attr-defined]
        for name, ms in self._host:
            out[name] = out.get(name, 0.0) + ms
        return out

    def reset(self) -> None:
        self._pairs.clear()
        self._host.clear()

"""Flat parameter layout: ``state_dict`` ⇄ one contiguous fp32 vector.

No reference counterpart — the reference keeps every client's model as a Python
dict of tensors and loops over keys (``fedavg.py:56-63``).  On B200 every
round-level op (K-AGG, K-BCAST, K-STATS, K-OPT) wants ONE base pointer and a
length, so a model is described once by a table ``name → (offset, shape, dtype)``
and K clients become the rows of a ``[K, N]`` arena slab.

Fidelity rules:

* every ``state_dict`` entry is included — parameters, BatchNorm running
  statistics and the int64 ``num_batches_tracked`` counters — because the
  reference averages all of them (``fedavg.py:56-66``);
* values are carried as fp32; ``unflatten`` casts back to each entry's original
  dtype the way ``load_state_dict`` does (truncation for integers);
* 0-d tensors occupy one element;
* each entry starts at a multiple of ``align`` elements (default 64 → 256 B in
  fp32, 128 B in the bf16 working copy) so kernels can use 16-byte vector
  accesses and TMA/UMMA descriptors on any entry; the padding is zero and stays
  zero under weighted sums, so norms and averages are unaffected.
"""

from __future__ import annotations

from dataclasses import dataclass
from typing import Iterable, Mapping, Sequence

import torch


@dataclass(frozen=True, slots=True)
class FlatEntry:
    name: str
    offset: int
    numel: int
    shape: tuple[int, ...]
    dtype: torch.dtype
    is_float: bool


class FlatLayout:
    """Immutable description of how a model state maps onto a flat vector."""

    def __init__(self, entries: Sequence[FlatEntry], total: int, align: int) -> None:
        self.entries = tuple(entries)
        self.total = int(total)
        self.align = int(align)
        self._by_name = {e.name: e for e in self.entries}

    # -- construction -----------------------------------------------------
    @classmethod
    def from_state(
        cls, state: Mapping[str, torch.Tensor | list | float | int], align: int = 64
    ) -> "FlatLayout":
        entries: list[FlatEntry] = []
        cursor = 0
        for name, value in state.items():
            t = value if isinstance(value, torch.Tensor) else torch.as_tensor(value)
            n = max(int(t.numel()), 0)
            entries.append(
                FlatEntry(name, cursor, n, tuple(t.shape), t.dtype, t.is_floating_point())
            )
            cursor += -(-n // align) * align if n else 0
        return cls(entries, cursor, align)

    @classmethod
    def from_module(cls, module: torch.nn.Module, align: int = 64) -> "FlatLayout":
        return cls.from_state(module.state_dict(), align=align)

    # -- queries ------------------------------------------------------------
    def __len__(self) -> int:
        return len(self.entries)

    def __contains__(self, name: str) -> bool:
        return name in self._by_name

    def __getitem__(self, name: str) -> FlatEntry:
        return self._by_name[name]

    @property
    def names(self) -> tuple[str, ...]:
        return tuple(e.name for e in self.entries)

    @property
    def num_values(self) -> int:
        """Number of real (non-padding) elements."""
        return sum(e.numel for e in self.entries)

    def matches(self, state: Mapping[str, object]) -> bool:
        return tuple(state.keys()) == self.names or set(state.keys()) == set(self.names)

    def float_mask(self, device: torch.device | str = "cpu") -> torch.Tensor:
        """1.0 on elements of floating-point entries, 0.0 elsewhere (incl. padding)."""
        mask = torch.zeros(self.total, dtype=torch.float32)
        for e in self.entries:
            if e.is_float:
                mask[e.offset : e.offset + e.numel] = 1.0
        return mask.to(device)

    def segment_table(self, device: torch.device | str = "cpu") -> torch.Tensor:
        """int64 ``[num_entries, 2]`` of (offset, numel) for per-tensor kernels."""
        return torch.tensor(
            [[e.offset, e.numel] for e in self.entries], dtype=torch.int64, device=device
        ).reshape(-1, 2)

    # -- packing --------------------------------------------------------------
    def flatten_into(
        self, state: Mapping[str, torch.Tensor | list | float | int], out: torch.Tensor
    ) -> torch.Tensor:
        """Write ``state`` into ``out`` (1-D, ``total`` elements, any float dtype)."""
        if out.numel() != self.total:
            raise ValueError(f"flat buffer has {out.numel()} elements, layout needs {self.total}")
        flat = out.view(-1)
        for e in self.entries:
            if e.numel == 0:
                continue
            value = state[e.name]
            t = value if isinstance(value, torch.Tensor) else torch.as_tensor(value)
            if t.numel() != e.numel:
                raise ValueError(
                    f"entry {e.name!r}: got {tuple(t.shape)}, layout has {e.shape}"
                )
            flat[e.offset : e.offset + e.numel].copy_(t.detach().reshape(-1))
        return out

    def flatten(
        self,
        state: Mapping[str, torch.Tensor | list | float | int],
        device: torch.device | str | None = None,
        dtype: torch.dtype = torch.float32,
    ) -> torch.Tensor:
        if device is None:
            first = next((v for v in state.values() if isinstance(v, torch.Tensor)), None)
            device = first.device if first is not None else "cpu"
        out = torch.zeros(self.total, dtype=dtype, device=device)
        return self.flatten_into(state, out)

    def stack(
        self,
        states: Iterable[Mapping[str, torch.Tensor | list | float | int]],
        device: torch.device | str | None = None,
        dtype: torch.dtype = torch.float32,
        pin: bool = False,
    ) -> torch.Tensor:
        """``[K, total]`` slab, one row per state."""
        states = list(states)
        if device is None:
            first = next(
                (v for s in states for v in s.values() if isinstance(v, torch.Tensor)), None
            )
            device = first.device if first is not None else "cpu"
        slab = torch.zeros(len(states), self.total, dtype=dtype, device=device,
                           pin_memory=pin and str(device) == "cpu")
        for row, state in zip(slab, states):
            self.flatten_into(state, row)
        return slab

    def views(self, flat: torch.Tensor) -> dict[str, torch.Tensor]:
        """Zero-copy named views into ``flat`` (dtype of ``flat``, not of the entries)."""
        base = flat.view(-1)
        return {
            e.name: base[e.offset : e.offset + e.numel].view(e.shape) for e in self.entries
        }

    def unflatten(self, flat: torch.Tensor, clone: bool = True) -> dict[str, torch.Tensor]:
        """Named tensors in each entry's ORIGINAL dtype (what ``state_dict()`` would hold)."""
        out: dict[str, torch.Tensor] = {}
        for name, view in self.views(flat).items():
            e = self._by_name[name]
            t = view.to(e.dtype) if view.dtype != e.dtype else (view.clone() if clone else view)
            out[name] = t
        return out

    def load_into(self, module: torch.nn.Module, flat: torch.Tensor) -> None:
        """``module.load_state_dict`` from a flat vector (dtype cast per entry)."""
        module.load_state_dict(self.unflatten(flat, clone=False))

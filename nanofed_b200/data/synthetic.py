"""Device-resident synthetic shards for simulated clients.

The reference feeds each client from a CPU ``DataLoader`` with worker processes
and a ``.to(device)`` per batch (``data/mnist.py:38-40``, ``trainer/base.py:141``).
For in-box simulation every client's shard is instead a slice of one tensor and
a batch is an index range: no workers, no per-batch H2D copy.  ``pinned_host``
keeps the master copy in pinned host memory for the end-to-end path that
measures the H2D copy of each round's inputs.
"""

from __future__ import annotations

from dataclasses import dataclass

import torch


@dataclass
class ClientShards:
    """``images[c, s]`` / ``labels[c, s]`` for client ``c``, sample ``s``."""

    images: torch.Tensor  # [C, S, ...]
    labels: torch.Tensor  # [C, S]
    sample_counts: torch.Tensor  # [C] number of valid samples per client (≤ S)

    @property
    def num_clients(self) -> int:
        return int(self.images.shape[0])

    @property
    def samples_per_client(self) -> int:
        return int(self.images.shape[1])

    def to(self, device: torch.device | str, non_blocking: bool = True) -> "ClientShards":
        return ClientShards(
            self.images.to(device, non_blocking=non_blocking),
            self.labels.to(device, non_blocking=non_blocking),
            self.sample_counts.to(device, non_blocking=non_blocking),
        )

    def nbytes(self) -> int:
        return (self.images.numel() * self.images.element_size()
                + self.labels.numel() * self.labels.element_size())


def make_client_shards(
    num_clients: int,
    samples_per_client: int,
    input_shape: tuple[int, ...] = (1, 28, 28),
    num_classes: int = 10,
    seed: int = 0,
    dtype: torch.dtype = torch.float32,
    pinned_host: bool = False,
    sample_counts: list[int] | None = None,
) -> ClientShards:
    gen = torch.Generator().manual_seed(seed)
    images = torch.randn(num_clients, samples_per_client, *input_shape, generator=gen)
    feat = images.flatten(2)
    probe = torch.randn(feat.shape[-1], num_classes,
                        generator=torch.Generator().manual_seed(1234))
    labels = (feat @ probe).argmax(dim=-1)
    images = images.to(dtype)
    counts = torch.tensor(
        sample_counts if sample_counts is not None else [samples_per_client] * num_clients,
        dtype=torch.int32,
    )
    if pinned_host and torch.cuda.is_available():
        images, labels, counts = images.pin_memory(), labels.pin_memory(), counts.pin_memory()
    return ClientShards(images, labels, counts)

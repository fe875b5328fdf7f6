"""MNIST loader with an offline fallback.

Parity: reference ``data/mnist.py:9-40`` — same signature, Normalize(0.1307,
0.3081), random ``Subset`` of ``subset_fraction``, ``DataLoader(shuffle=train)``.

Differences: the subset draw can be seeded (``seed=``; the reference uses the
global NumPy RNG unseeded); ``num_workers`` is a parameter; and when torchvision
or the dataset files are unavailable (no network on the build/bench boxes) the
function falls back to ``synthetic_mnist`` — random 28×28 images with the MNIST
normalisation statistics and labels a fixed random linear probe makes learnable —
unless ``allow_synthetic=False``.
"""

from __future__ import annotations

from pathlib import Path

import numpy as np
import torch
from torch.utils.data import DataLoader, Dataset, Subset, TensorDataset

MNIST_TRAIN_SIZE = 60_000
MNIST_TEST_SIZE = 10_000


def synthetic_mnist(
    num_samples: int,
    seed: int = 0,
    device: torch.device | str = "cpu",
    dtype: torch.dtype = torch.float32,
) -> tuple[torch.Tensor, torch.Tensor]:
    """``(images [n,1,28,28], labels [n])``; labels are a learnable function of the pixels."""
    gen = torch.Generator().manual_seed(seed)
    images = torch.randn(num_samples, 1, 28, 28, generator=gen)
    probe = torch.randn(28 * 28, 10, generator=torch.Generator().manual_seed(1234))
    labels = (images.flatten(1) @ probe).argmax(dim=1)
    return images.to(device=device, dtype=dtype), labels.to(device)


def load_mnist_data(
    data_dir: str | Path,
    batch_size: int,
    train: bool = True,
    download: bool = True,
    subset_fraction: float = 0.2,
    *,
    seed: int | None = None,
    num_workers: int = 2,
    allow_synthetic: bool = True,
) -> DataLoader:
    dataset: Dataset
    full = MNIST_TRAIN_SIZE if train else MNIST_TEST_SIZE
    try:
        from torchvision import datasets, transforms  # type: ignore

        transform = transforms.Compose(
            [transforms.ToTensor(), transforms.Normalize((0.1307,), (0.3081,))]
        )
        dataset = datasets.MNIST(str(data_dir), train=train, download=download,
                                 transform=transform)
    except Exception:
        if not allow_synthetic:
            raise
        n = int(full * subset_fraction) if subset_fraction < 1.0 else full
        x, y = synthetic_mnist(n, seed=seed or 0)
        return DataLoader(TensorDataset(x, y), batch_size=batch_size, shuffle=train)

    if subset_fraction < 1.0:
        n = int(len(dataset) * subset_fraction)  # type: ignore[arg-type]
        rng = np.random.default_rng(seed) if seed is not None else np.random
        picks = rng.choice(len(dataset), size=n, replace=False)  # type: ignore[arg-type]
        dataset = Subset(dataset, [int(i) for i in picks])
    return DataLoader(dataset, batch_size=batch_size, shuffle=train, num_workers=num_workers)

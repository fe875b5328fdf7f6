"""Rank bootstrap: one process per GPU, ``torch.distributed`` for the plumbing.

No reference counterpart (the reference is single-process asyncio).
``init()`` reads the ``torchrun`` environment (RANK / LOCAL_RANK / WORLD_SIZE /
MASTER_ADDR / MASTER_PORT), picks ``nccl`` on GPUs and ``gloo`` on CPU, and binds
the process to ``cuda:LOCAL_RANK``.  NCCL is used for rendezvous, barriers, and
as the *baseline* all-reduce; the product path moves model data with this
framework's own peer-memory kernels (``parallel.peer``).
"""

from __future__ import annotations

import os
from dataclasses import dataclass

import torch
import torch.distributed as dist


@dataclass(frozen=True)
class RankInfo:
    rank: int
    local_rank: int
    world_size: int
    device: torch.device
    backend: str | None

    @property
    def is_main(self) -> bool:
        return self.rank == 0

    @property
    def distributed(self) -> bool:
        return self.world_size > 1


_info: RankInfo | None = None


def init(device: str | None = None) -> RankInfo:
    """Idempotent.  Single-process runs need no environment at all."""
    global _info
    if _info is not None:
        return _info
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    use_cuda = torch.cuda.is_available() and device != "cpu"
    dev = torch.device(f"cuda:{local}") if use_cuda else torch.device("cpu")
    backend = None
    if use_cuda:
        torch.cuda.set_device(dev)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        backend = "nccl" if use_cuda else "gloo"
        if not dist.is_initialized():
            kwargs = {"device_id": dev} if use_cuda else {}
            dist.init_process_group(backend=backend, rank=rank, world_size=world, **kwargs)
    _info = RankInfo(rank, local, world, dev, backend)
    return _info


def info() -> RankInfo:
    return _info if _info is not None else init()


def barrier() -> None:
    if dist.is_initialized():
        dist.barrier()


def all_reduce_max(value: float, device: torch.device) -> float:
    """Max over ranks of a host scalar (used for device-timed metrics)."""
    if not dist.is_initialized():
        return value
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def all_reduce_sum(value: float, device: torch.device) -> float:
    if not dist.is_initialized():
        return value
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def shutdown() -> None:
    global _info
    if dist.is_initialized():
        dist.destroy_process_group()
    _info = None

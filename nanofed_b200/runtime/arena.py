"""Per-GPU arena: every round-level tensor in a handful of flat slabs.

No reference counterpart.  Layout per GPU (C = clients on this GPU, N = padded
flat model length, see ``runtime.flat``):

    master      [N]    fp32   global model (the FedAvg result / broadcast source)
    master_lowp [N]    bf16   working copy for bf16 compute
    slots       [C,N]  fp32   client weights — K-AGG reads these rows in place, so
                              "upload" is not a copy at all
    slots_lowp  [C,N]  bf16   optional bf16 working copies (hand-written kernels)
    grads       [C,N]  fp32   flat gradients (autograd writes into views of it)
    momentum    [C,N]  fp32   optional optimiser state, reset by K-BCAST
    counts/mask/sumsq/stats   per-client scalars consumed on the device

180 GB of HBM3e per B200 holds e.g. 32 clients × 1.2 M params × 14 B = 0.5 GB for
the MNIST CNN, 8 × 11.2 M × 14 B = 1.3 GB for ResNet-18, and single 1 B-parameter
updates (4 GB fp32) with room for K-AGG's streaming waves.
"""

from __future__ import annotations

from dataclasses import dataclass, field

import torch

from .flat import FlatLayout


@dataclass
class Arena:
    layout: FlatLayout
    num_clients: int
    device: torch.device
    with_momentum: bool = False
    with_lowp: bool = False
    with_grads: bool = True
    with_adam: bool = False          # exp_avg / exp_avg_sq rows (reset every round like momentum)
    master: torch.Tensor = field(init=False)
    slots: torch.Tensor = field(init=False)

    def __post_init__(self) -> None:
        n, c, dev = self.layout.total, self.num_clients, self.device
        f32 = dict(dtype=torch.float32, device=dev)
        self.master = torch.zeros(n, **f32)
        self.master_lowp = torch.zeros(n, dtype=torch.bfloat16, device=dev)
        self.slots = torch.zeros(c, n, **f32)
        self.slots_lowp = (
            torch.zeros(c, n, dtype=torch.bfloat16, device=dev) if self.with_lowp else None
        )
        self.grads = torch.zeros(c, n, **f32) if self.with_grads else None
        self.momentum = torch.zeros(c, n, **f32) if self.with_momentum else None
        self.exp_avg = torch.zeros(c, n, **f32) if self.with_adam else None
        self.exp_avg_sq = torch.zeros(c, n, **f32) if self.with_adam else None
        self.counts = torch.ones(c, **f32)
        self.mask = torch.ones(c, **f32)
        self.sumsq = torch.zeros(c, **f32)
        self.nonfinite = torch.zeros(c, dtype=torch.int32, device=dev)
        self.stats = torch.zeros(c, 3, **f32)  # (loss_sum, correct, samples) per client
        self.noise_mask = self.layout.float_mask(dev).to(torch.uint8)

    def nbytes(self) -> int:
        total = 0
        for name in ("master", "master_lowp", "slots", "slots_lowp", "grads", "momentum", "exp_avg",
                     "exp_avg_sq"):
            t = getattr(self, name)
            if t is not None:
                total += t.numel() * t.element_size()
        return total

    def client_views(self, k: int) -> dict[str, torch.Tensor]:
        return self.layout.views(self.slots[k])

    def grad_views(self, k: int) -> dict[str, torch.Tensor]:
        assert self.grads is not None
        return self.layout.views(self.grads[k])

    def load_global(self, state: dict[str, torch.Tensor]) -> None:
        host = self.layout.flatten(state, device="cpu")
        self.master.copy_(host, non_blocking=True)
        self.master_lowp.copy_(self.master)

    def global_state(self) -> dict[str, torch.Tensor]:
        """Global model as a CPU ``state_dict`` in original dtypes (one D2H copy)."""
        return self.layout.unflatten(self.master.detach().cpu())

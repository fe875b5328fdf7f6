"""In-box client API: the reference's client loop, unchanged, on the peer-memory transport.

The reference's example (``examples/mnist/run_experiment.py:21-86``) is written against
``HTTPClient`` (``nanofed/communication/http/client.py:103-211``)::

    async with HTTPClient(server_url, client_id) as client:
        while not await client.check_server_status():
            state, rnd = await client.fetch_global_model()
            model.load_state_dict(state); ...train...
            await client.submit_update(model, metrics)

``PeerClient`` has the same three coroutines, but for clients that live on the GPUs of one box
nothing is serialised and nothing crosses a socket:

* ``fetch_global_model`` hands out the global model as it sits in the arena (a snapshot of the
  fp32 master vector, cast back to the ``state_dict`` dtypes);
* ``submit_update`` writes the trained ``state_dict`` into the client's arena row (one
  device-to-device copy per entry; a model whose parameters already ARE views of the row -
  ``bind_model`` - costs nothing), stores the sample count and raises the client's participation
  flag - "upload" is a flag, K-AGG reads the row in place;
* when every expected client of the round has submitted, the hub runs the fused FedAvg kernel
  (``ops.fedavg_reduce`` on one GPU, the peer all-reduce kernel across GPUs), writes the
  reference-format round record and releases the clients waiting for the next round.

``ClientSpec`` is the declarative per-client description SURVEY section 7.1 asks for: sample
count, data shard, own ``TrainingConfig`` / ``PrivacyConfig`` / optimizer factory; ``PeerHub.run``
drives one coroutine per spec with the reference's trainer classes.
"""

from __future__ import annotations

import asyncio
from dataclasses import dataclass, field
from typing import Any, Callable

import torch
import torch.nn as nn

from nanofed_b200 import ops
from nanofed_b200.core.exceptions import NanoFedError
from nanofed_b200.runtime.arena import Arena
from nanofed_b200.runtime.flat import FlatLayout
from nanofed_b200.utils import Logger


@dataclass
class ClientSpec:
    """One simulated client, declaratively."""

    client_id: str
    num_samples: int
    data: Any = None                       # DataLoader | (images, labels) | None (user-driven loop)
    training: Any = None                   # trainer.TrainingConfig (own lr / epochs / batch size)
    privacy: Any = None                    # privacy.PrivacyConfig -> PrivateTrainer for this client
    optimizer: Callable[[nn.Module, float], torch.optim.Optimizer] | None = None
    metadata: dict[str, Any] = field(default_factory=dict)


class PeerHub:
    """The in-box 'server': global model + client rows in one arena, aggregation by K-AGG."""

    def __init__(
        self,
        model: nn.Module,
        clients: list[ClientSpec] | int,
        num_rounds: int,
        device: str | torch.device | None = None,
        min_clients: int | None = None,
        privacy: Any = None,
    ) -> None:
        if isinstance(clients, int):
            clients = [ClientSpec(f"client_{k}", 1) for k in range(clients)]
        self.specs = list(clients)
        self.num_rounds = int(num_rounds)
        dev = torch.device(device) if device is not None else next(model.parameters()).device
        self.device = dev
        self.model = model
        self.layout = FlatLayout.from_module(model)
        self.arena = Arena(self.layout, len(self.specs), dev, with_grads=False)
        self.arena.load_global(model.state_dict())
        self.arena.mask.zero_()             # participation flags: raised by submit_update
        self.min_clients = int(min_clients) if min_clients is not None else len(self.specs)
        self.privacy = privacy              # central DP fused into the aggregation kernel
        self.round = 0
        self.history: list[dict[str, Any]] = []
        self._submitted: dict[int, dict[str, float]] = {}
        self._cond: asyncio.Condition | None = None
        self._logger = Logger()
        self._noise_offset = 0

    # -- clients -----------------------------------------------------------------------
    def client(self, index: int) -> "PeerClient":
        return PeerClient(self, index)

    def bind_model(self, model: nn.Module, index: int) -> nn.Module:
        """Make ``model``'s floating-point parameters views of client ``index``'s arena row, so
        training writes the 'upload' in place and ``submit_update`` copies nothing for them."""
        views = self.layout.views(self.arena.slots[index])
        for name, p in model.named_parameters():
            p.data = views[name]
        return model

    @property
    def is_training_done(self) -> bool:
        return self.round >= self.num_rounds

    def _condition(self) -> asyncio.Condition:
        if self._cond is None:
            self._cond = asyncio.Condition()
        return self._cond

    # -- the three calls behind PeerClient --------------------------------------------------
    def _global_state(self) -> dict[str, torch.Tensor]:
        return self.layout.unflatten(self.arena.master)      # clones; original dtypes

    async def _submit(self, index: int, state: dict[str, torch.Tensor],
                      metrics: dict[str, Any]) -> bool:
        if self.is_training_done:
            return False
        cond = self._condition()
        async with cond:
            if index in self._submitted:
                raise NanoFedError(f"client {self.specs[index].client_id} already submitted an "
                                   f"update for round {self.round}")
            row = self.arena.slots[index]
            views = self.layout.views(row)
            for name, view in views.items():
                src = state[name]
                if src.data_ptr() == view.data_ptr():
                    continue                                  # bound model: already in place
                view.copy_(src.detach().reshape(view.shape))
            n = float(metrics.get("num_samples") or metrics.get("samples_processed")
                      or self.specs[index].num_samples)
            self.arena.counts[index] = n
            self.arena.mask[index] = 1.0
            self._submitted[index] = {k: float(v) for k, v in metrics.items()
                                      if isinstance(v, (int, float))}
            if len(self._submitted) >= self.min_clients and len(self._submitted) == len(self.specs):
                self._aggregate()
                cond.notify_all()
        return True

    def close_round(self) -> None:
        """Aggregate with the clients that have submitted so far (``min_clients`` reached)."""
        if len(self._submitted) < self.min_clients:
            raise NanoFedError(f"round {self.round}: {len(self._submitted)} updates, "
                               f"{self.min_clients} required")
        self._aggregate()

    def _aggregate(self) -> None:
        a = self.arena
        kw: dict[str, Any] = {}
        if self.privacy is not None:
            ss, _ = ops.update_stats(a.slots)
            a.sumsq.copy_(ss)
            k_active = max(len(self._submitted), 1)
            kw = dict(sumsq=a.sumsq, clip_norm=float(self.privacy.max_gradient_norm),
                      noise_std=self.privacy.noise_multiplier * self.privacy.max_gradient_norm
                      / k_active, noise_mask=a.noise_mask,
                      seed=int.from_bytes(__import__("os").urandom(8), "little") >> 1,
                      offset=self._noise_offset)
            self._noise_offset += self.layout.total
        # rows of clients that did not submit are never read (mask = 0)
        ops.fedavg_reduce(a.slots, a.counts, out=a.master, out_lowp=a.master_lowp, mask=a.mask, **kw)
        counts = {i: float(a.counts[i]) for i in self._submitted}
        total = sum(counts.values()) or 1.0
        agg: dict[str, float] = {}
        for i, m in self._submitted.items():
            for name, v in m.items():
                agg[name] = agg.get(name, 0.0) + v * counts[i] / total
        self.history.append({
            "round_id": self.round, "num_clients": len(self._submitted), "agg_metrics": agg,
            "client_metrics": [{"client_id": self.specs[i].client_id, "metrics": m,
                                "weight": counts[i] / total}
                               for i, m in sorted(self._submitted.items())],
        })
        self.model.load_state_dict(self._global_state())
        self._submitted = {}
        a.mask.zero_()
        self.round += 1

    # -- declarative driver ------------------------------------------------------------------
    async def run(self, model_factory: Callable[[], nn.Module]) -> list[dict[str, Any]]:
        """Train every ``ClientSpec`` that carries data with the reference's trainer classes until
        ``num_rounds`` are done; returns the per-round records."""
        from torch.utils.data import DataLoader, TensorDataset

        from nanofed_b200.trainer import PrivateTrainer, TorchTrainer, TrainingConfig

        async def one(index: int, spec: ClientSpec) -> None:
            cfg = spec.training or TrainingConfig(epochs=1, batch_size=64, learning_rate=0.1,
                                                  device=str(self.device), log_interval=10**9)
            trainer = (PrivateTrainer(cfg, spec.privacy) if spec.privacy is not None
                       else TorchTrainer(cfg))
            data = spec.data
            if isinstance(data, tuple):
                data = DataLoader(TensorDataset(*data), batch_size=cfg.batch_size, shuffle=False)
            async with self.client(index) as client:
                while not await client.check_server_status():
                    state, _ = await client.fetch_global_model()
                    model = model_factory().to(self.device)
                    model.load_state_dict(state)
                    opt = (spec.optimizer(model, cfg.learning_rate) if spec.optimizer
                           else torch.optim.SGD(model.parameters(), lr=cfg.learning_rate))
                    metrics = None
                    for epoch in range(cfg.epochs):
                        metrics = trainer.train_epoch(model, data, opt, epoch)
                    md = metrics.to_dict() if metrics is not None else {}
                    md["num_samples"] = spec.num_samples
                    if not await client.submit_update(model, md):
                        break

        await asyncio.gather(*[one(i, s) for i, s in enumerate(self.specs) if s.data is not None])
        return self.history


class PeerClient:
    """``HTTPClient``-shaped handle of one in-box client (see the module docstring)."""

    def __init__(self, hub: PeerHub, index: int) -> None:
        if not 0 <= index < len(hub.specs):
            raise NanoFedError(f"client index {index} outside [0, {len(hub.specs)})")
        self._hub, self._index = hub, index
        self._client_id = hub.specs[index].client_id
        self._current_round = 0
        self._open = False

    async def __aenter__(self) -> "PeerClient":
        self._open = True
        return self

    async def __aexit__(self, *exc: Any) -> None:
        self._open = False

    @property
    def client_id(self) -> str:
        return self._client_id

    async def fetch_global_model(self) -> tuple[dict[str, torch.Tensor], int]:
        """``(state_dict, round)``; waits while the round this client already contributed to is
        still open (the reference's client polls ``GET /model`` for that)."""
        if not self._open:
            raise NanoFedError("Client session not initialized")
        hub = self._hub
        cond = hub._condition()
        async with cond:
            await cond.wait_for(lambda: self._index not in hub._submitted or hub.is_training_done)
        self._current_round = hub.round
        return hub._global_state(), hub.round

    async def submit_update(self, model: nn.Module, metrics: dict[str, Any]) -> bool:
        if not self._open:
            raise NanoFedError("Client session not initialized")
        return await self._hub._submit(self._index, model.state_dict(), dict(metrics))

    async def check_server_status(self) -> bool:
        """True once training is finished (reference: ``GET /status`` → ``is_training_done``)."""
        if not self._open:
            raise NanoFedError("Client session not initialized")
        hub = self._hub
        cond = hub._condition()
        async with cond:      # a client that has contributed waits here for its round to close
            await cond.wait_for(lambda: self._index not in hub._submitted or hub.is_training_done)
        return hub.is_training_done

"""Cohort scheduler: simulated clients packed onto the GPUs of one box.

No reference counterpart — the reference's "clients" are coroutines that train
one after another on a shared event loop (``run_experiment.py:126-131``;
``tutorial.ipynb:406-506``) and exchange models as JSON over HTTP.  Here:

* a **client** is a row of the arena (weights, gradients, optimiser state,
  sample count, participation flag) plus a shard of a device-resident dataset;
* a **cohort** is the set of clients of one GPU; they train in lock-step, so each
  local step is a bounded number of launches regardless of the client count;
* a **round** is ``K-BCAST → local steps → K-STATS → K-AGG`` enqueued without a
  single host sync; per-round metrics come back in one tiny D2H copy;
* across GPUs (one process per GPU) the aggregation is one fused kernel over
  peer memory (``parallel.peer``), NCCL being the baseline only.

Two engines implement ``local_steps()``:

* ``MnistCohortNet`` (``models.mnist_cohort``) — the flagship: hand-written
  sm_100a kernels, clients as the GEMM group dimension;
* ``AutogradCohort`` (this file) — any ``nn.Module``: per-client autograd on a
  pool of CUDA streams, parameters and gradients bound to arena views, the whole
  local epoch captured in ONE CUDA graph, with K-XENT / K-OPT hand-written.

``FedSimulation`` is the public entry point (``bench.py`` and the CLI use it) and
plugs into ``Coordinator(engine=…)`` for checkpointing with the reference's
on-disk layout.
"""

from __future__ import annotations

import os

import math
from dataclasses import dataclass, field
from typing import Any, Callable

import torch
import torch.nn as nn

from nanofed_b200 import ops
from nanofed_b200.data.synthetic import ClientShards, make_client_shards
from nanofed_b200.privacy.accountant import GaussianAccountant
from nanofed_b200.privacy.config import PrivacyConfig
from nanofed_b200.runtime import dist as rdist
from nanofed_b200.runtime.arena import Arena
from nanofed_b200.runtime.flat import FlatLayout
from nanofed_b200.utils import Logger, PhaseTimer


@dataclass
class CohortConfig:
    """How to pack and train simulated clients (SURVEY §5.6 'one new config')."""

    model: str = "mnist_cnn"            # "mnist_cnn" | "resnet18" | "resnet18_gn"
    clients_per_gpu: int = 32
    samples_per_client: int = 512
    batch_size: int = 64
    local_epochs: int = 1
    learning_rate: float = 0.1
    momentum: float = 0.0
    weight_decay: float = 0.0
    # "sgd" (momentum / weight_decay above) or "adam" (torch.optim.Adam semantics: L2 weight decay,
    # bias correction; fused K-OPT kernel nfb_fused_adam over the whole arena).  Like the momentum
    # buffer, the Adam moments start from zero every round (the reference's clients build a fresh
    # optimizer per round, examples/mnist/run_experiment.py:65-73).
    optimizer: str = "sgd"
    adam_betas: tuple[float, float] = (0.9, 0.999)
    adam_eps: float = 1e-8
    compute_dtype: str = "bf16"         # "bf16" | "fp32"
    engine: str = "auto"                # "auto" | "grouped" | "autograd"
    streams_per_gpu: int = 8            # autograd engine: 4 -> 8 streams measured +4 % on ResNet-18
    use_cuda_graph: bool = True
    transport: str = "peer"             # "peer" (fused kernel) | "nccl" (baseline)
    # transport="peer" fails loudly when the symmetric allocation / fused kernel cannot be set up;
    # True lets it degrade to the NCCL baseline instead (logged, and visible as agg_path)
    allow_nccl_fallback: bool = False
    # Single-GPU rounds: K-BCAST (replication of the new global model into the client rows + the
    # optimizer-state reset) runs inside the aggregation kernel, so a steady-state round has no
    # broadcast launch at all.  False keeps the client rows as trained after a round (debugging).
    fuse_bcast: bool = True
    privacy: PrivacyConfig | None = None  # central DP fused into K-AGG
    # DP-SGD on the clients (reference PrivateTrainer: batch gradient clipped to max_gradient_norm,
    # N(0, (noise_multiplier * max_gradient_norm)^2) added per parameter), inside K-OPT
    local_privacy: PrivacyConfig | None = None
    # Seed of the DP noise stream.  None (default): drawn from os.urandom on rank 0 and shared with
    # the other ranks, so the noise cannot be regenerated from the configuration.  A fixed value is
    # for tests and reproducibility studies only - it voids the privacy guarantee.
    dp_seed: int | None = None
    seed: int = 0
    device: str | None = None
    max_steps_per_epoch: int | None = None
    sample_counts: list[int] | None = None  # per-client n_k on this rank (≤ samples_per_client)
    screen_nonfinite: bool = False  # K-STATS every round; clients with NaN/Inf weights are masked out
    # Straggler deadline: local training of a cohort is timed with CUDA events; a
    # GPU whose cohort needs longer than this is left out of the round's average (its clients run
    # in lock-step, so the GPU is the unit of straggling).  Costs one host sync per round and, with
    # several ranks, one tiny all-reduce so that a round in which EVERY cohort is late is kept.
    round_deadline_ms: float | None = None

    @property
    def steps_per_epoch(self) -> int:
        steps = self.samples_per_client // self.batch_size
        if self.max_steps_per_epoch is not None:
            steps = min(steps, self.max_steps_per_epoch)
        return max(steps, 1)

    @property
    def local_steps(self) -> int:
        return self.steps_per_epoch * self.local_epochs


@dataclass
class RoundOutcome:
    round_id: int
    num_clients: int
    agg_metrics: dict[str, float]
    client_metrics: list[dict[str, Any]]
    samples: int
    phase_ms: dict[str, float] = field(default_factory=dict)


@dataclass
class PendingRound:
    """Handle of a round whose metrics are still on their way to the host (``submit_round``)."""

    round_id: int
    slot: int
    event: Any


def build_model(name: str) -> nn.Module:
    from nanofed_b200.models import MNISTModel, ResNet18

    if name == "mnist_cnn":
        return MNISTModel()
    if name == "resnet18":
        return ResNet18(norm="batch")
    if name == "resnet18_gn":
        return ResNet18(norm="group")
    raise ValueError(f"unknown model {name!r}")


# ---------------------------------------------------------------------------
# K-XENT as an autograd node (AutogradCohort's loss)
# ---------------------------------------------------------------------------
class _FusedXent(torch.autograd.Function):
    """Mean cross-entropy whose forward already produced ``dlogits`` and device stats."""

    @staticmethod
    def forward(ctx, logits: torch.Tensor, targets: torch.Tensor, stats: torch.Tensor | None):
        lg = logits.contiguous()
        loss_sum, _, dlogits = ops.xent_fused(lg.unsqueeze(0), targets.unsqueeze(0), stats=stats)
        ctx.save_for_backward(dlogits.squeeze(0))
        return loss_sum.squeeze(0) / lg.shape[0]

    @staticmethod
    def backward(ctx, grad_out):
        (dlogits,) = ctx.saved_tensors
        return dlogits * grad_out, None, None


def fused_cross_entropy(logits, targets, stats=None):
    return _FusedXent.apply(logits, targets, stats)


# ---------------------------------------------------------------------------
# Generic engine: any nn.Module, autograd per client, flat arena + fused ops
# ---------------------------------------------------------------------------
def _locate(module: nn.Module, dotted: str) -> tuple[nn.Module, str]:
    owner = module
    parts = dotted.split(".")
    for p in parts[:-1]:
        owner = getattr(owner, p)
    return owner, parts[-1]


class AutogradCohort:
    """Local training for C clients of one GPU through PyTorch autograd.

    Every client's parameters/buffers are *views* of its arena row and its
    gradients are views of the gradient slab, so (a) the optimiser is one fused
    K-OPT launch over the whole ``[C·N]`` slab per step, (b) K-AGG reads the
    trained weights in place, and (c) nothing is copied at round boundaries.
    ``streams_per_gpu`` worker modules run clients concurrently on separate CUDA
    streams; the whole local epoch (all clients × all steps) is captured once in
    a CUDA graph and replayed each round — the launch-bound inner loop costs one
    ``cudaGraphLaunch``.
    """

    def __init__(self, cfg: CohortConfig, arena: Arena, shards: ClientShards) -> None:
        self.cfg = cfg
        self.arena = arena
        self.shards = shards
        self.device = arena.device
        self.cuda = self.device.type == "cuda"
        self.logger = Logger()
        n_workers = max(1, min(cfg.streams_per_gpu, arena.num_clients)) if self.cuda else 1
        self.workers = [build_model(cfg.model).to(self.device).train() for _ in range(n_workers)]
        self.streams = (
            [torch.cuda.Stream(device=self.device) for _ in range(n_workers)] if self.cuda else []
        )
        layout = arena.layout
        self._param_names = [n for n, _ in self.workers[0].named_parameters()]
        self._float_buffers = [
            n for n, b in self.workers[0].named_buffers() if b.is_floating_point() and n in layout
        ]
        self._int_buffers = [
            n for n, b in self.workers[0].named_buffers()
            if not b.is_floating_point() and n in layout
        ]
        # Integer state (e.g. BatchNorm counters) lives outside the fp32 arena: ONE [C, total]
        # tensor whose slices are the modules' buffers, so the per-round exchange with the arena
        # is one gather and one scatter (ResNet-18 has 20 counters per client; one copy per
        # counter and client cost 2.3 ms of every 73 ms round).
        sizes = [int(layout[n].numel) for n in self._int_buffers]
        offsets = [sum(sizes[:i]) for i in range(len(sizes))]
        self._int_all = torch.zeros(arena.num_clients, max(sum(sizes), 1), dtype=torch.int64,
                                    device=self.device)
        self._int_idx = torch.tensor(
            [layout[n].offset + e for n, sz in zip(self._int_buffers, sizes) for e in range(sz)],
            dtype=torch.long, device=self.device)
        self._int_state = [
            {n: self._int_all[k, o : o + sz].view(layout[n].shape).to(layout[n].dtype)
             if layout[n].dtype != torch.int64 else self._int_all[k, o : o + sz].view(layout[n].shape)
             for n, o, sz in zip(self._int_buffers, offsets, sizes)}
            for k in range(arena.num_clients)
        ]
        self._int_fast = all(layout[n].dtype == torch.int64 for n in self._int_buffers)
        self._slot_views = [arena.client_views(k) for k in range(arena.num_clients)]
        self._grad_views = [arena.grad_views(k) for k in range(arena.num_clients)]
        self._graph: "torch.cuda.CUDAGraph | None" = None
        self._autocast = cfg.compute_dtype == "bf16" and self.cuda
        # NHWC inputs for the library convolutions (weights stay arena views): cuDNN's bf16 kernels
        # are NHWC, NCHW inputs cost a layout change in every layer (ResNet-18 round 70.7 -> 35.8 ms)
        self._channels_last = os.environ.get("NFB_CHANNELS_LAST", "1") != "0" and self.cuda
        self.launches_per_round = 0
        # weight decay is a property of PARAMETERS: the fused optimizer runs over whole arena rows,
        # which also hold BatchNorm running statistics, counters and padding (torch.optim.SGD never
        # touches those), so the decay term is added to the gradients of parameter columns only
        self._dp_step = torch.zeros(1, dtype=torch.int64, device=self.device)
        self._dp_seed = int(cfg.dp_seed if getattr(cfg, "dp_seed", None) is not None
                            else int.from_bytes(os.urandom(8), "little") >> 1)
        self._dp_noise_mask = None
        if getattr(cfg, "local_privacy", None) is not None:
            pm = torch.zeros(layout.total, dtype=torch.uint8)      # noise on PARAMETERS only
            for n in self._param_names:
                pm[layout[n].offset : layout[n].offset + layout[n].numel] = 1
            self._dp_noise_mask = pm.to(self.device)
        self._param_mask = None
        if cfg.weight_decay != 0.0:
            pm = torch.zeros(layout.total, dtype=torch.float32)
            for n in self._param_names:
                pm[layout[n].offset : layout[n].offset + layout[n].numel] = 1.0
            self._param_mask = pm.to(self.device)

    # -- binding ---------------------------------------------------------------
    def _bind(self, worker: nn.Module, k: int) -> None:
        slots, grads = self._slot_views[k], self._grad_views[k]
        for name in self._param_names:
            owner, attr = _locate(worker, name)
            p = owner._parameters[attr]
            p.data = slots[name]
            p.grad = grads[name]
        for name in self._float_buffers:
            owner, attr = _locate(worker, name)
            owner._buffers[attr] = slots[name]
        for name in self._int_buffers:
            owner, attr = _locate(worker, name)
            owner._buffers[attr] = self._int_state[k][name]

    def after_broadcast(self) -> None:
        """Integer entries: arena (fp32) → per-client integer tensors."""
        if not self._int_buffers:
            return
        if self._int_fast:
            vals = self.arena.master[self._int_idx].round().to(torch.int64)
            self._int_all[:, : vals.numel()] = vals.unsqueeze(0)
            return
        master = self.arena.layout.views(self.arena.master)
        for state in self._int_state:
            for name, t in state.items():
                t.copy_(master[name].to(t.dtype))

    def before_aggregate(self) -> None:
        """Integer entries: per-client integer tensors → arena rows (as fp32)."""
        if not self._int_buffers:
            return
        if self._int_fast:
            n = self._int_idx.numel()
            self.arena.slots[:, self._int_idx] = self._int_all[:, :n].to(torch.float32)
            return
        for k, state in enumerate(self._int_state):
            for name, t in state.items():
                self._slot_views[k][name].copy_(t.to(torch.float32))

    # -- one lock-step local step for all clients ----------------------------------
    def _client_step(self, worker: nn.Module, k: int, step: int) -> None:
        cfg = self.cfg
        lo = (step % cfg.steps_per_epoch) * cfg.batch_size
        x = self.shards.images[k, lo : lo + cfg.batch_size]
        y = self.shards.labels[k, lo : lo + cfg.batch_size]
        self._bind(worker, k)
        if self._channels_last and x.dim() == 4:
            x = x.contiguous(memory_format=torch.channels_last)
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=self._autocast):
            logits = worker(x)
        loss = fused_cross_entropy(logits.float(), y, self.arena.stats[k : k + 1])
        loss.backward()

    def _all_steps(self) -> None:
        cfg, arena = self.cfg, self.arena
        c = arena.num_clients
        main = torch.cuda.current_stream(self.device) if self.cuda else None
        for step in range(cfg.local_steps):
            arena.grads.zero_()
            if cfg.optimizer == "adam" and step == 0:      # fresh optimizer every round
                arena.exp_avg.zero_()
                arena.exp_avg_sq.zero_()
            if self.cuda:
                for s in self.streams:
                    s.wait_stream(main)
                for k in range(c):
                    w = k % len(self.workers)
                    with torch.cuda.stream(self.streams[w]):
                        self._client_step(self.workers[w], k, step)
                for s in self.streams:
                    main.wait_stream(s)
            else:
                for k in range(c):
                    self._client_step(self.workers[0], k, step)
            dp_kw: dict[str, Any] = {}
            dp = getattr(cfg, "local_privacy", None)
            if dp is not None:       # client-side DP-SGD: clip each client's batch gradient, add noise
                ss, _ = ops.update_stats(arena.grads)
                self._dp_step.add_(1)
                dp_kw = dict(row_sumsq=ss, clip_norm=float(dp.max_gradient_norm),
                             noise_std=float(dp.noise_multiplier * dp.max_gradient_norm),
                             noise_mask=self._dp_noise_mask, seed=self._dp_seed,
                             step_counter=self._dp_step, row_len=arena.layout.total)
            if self._param_mask is not None:
                arena.grads.addcmul_(arena.slots, self._param_mask, value=cfg.weight_decay)
            if cfg.optimizer == "adam":
                # buffers / padding columns have zero gradients and zero moments: 0 / (0 + eps) = 0
                ops.fused_adam_(arena.slots.view(-1), arena.grads.view(-1), arena.exp_avg.view(-1),
                                arena.exp_avg_sq.view(-1), step + 1, cfg.learning_rate,
                                beta1=cfg.adam_betas[0], beta2=cfg.adam_betas[1], eps=cfg.adam_eps)
                continue
            ops.fused_sgd_(
                arena.slots.view(-1), arena.grads.view(-1), cfg.learning_rate,
                momentum_buf=None if arena.momentum is None else arena.momentum.view(-1),
                momentum=cfg.momentum, weight_decay=0.0,
                first_step=(step == 0), **dp_kw,
            )

    def local_steps(self) -> None:
        """Enqueue the whole local epoch(s) for every client of this GPU."""
        if not (self.cuda and self.cfg.use_cuda_graph):
            self._all_steps()
            return
        if self._graph is None:
            self._capture()
        self._graph.replay()

    def _capture(self) -> None:
        # Warm up eagerly (cuDNN heuristics, lazy inits) on a side stream, restoring
        # the arena afterwards so warm-up leaves no trace in the model state.
        arena = self.arena
        keep = (arena.slots.clone(), arena.stats.clone(),
                None if arena.momentum is None else arena.momentum.clone(),
                [{n: t.clone() for n, t in s.items()} for s in self._int_state])
        side = torch.cuda.Stream(device=self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):
            # one eager step per client is enough to initialise every lazy path
            old_cap, old_epochs = self.cfg.max_steps_per_epoch, self.cfg.local_epochs
            self.cfg.max_steps_per_epoch, self.cfg.local_epochs = 1, 1
            try:
                self._all_steps()
            finally:
                self.cfg.max_steps_per_epoch, self.cfg.local_epochs = old_cap, old_epochs
        torch.cuda.current_stream(self.device).wait_stream(side)
        torch.cuda.synchronize(self.device)

        def restore() -> None:
            arena.slots.copy_(keep[0])
            arena.stats.copy_(keep[1])
            if keep[2] is not None:
                arena.momentum.copy_(keep[2])
            for s, saved in zip(self._int_state, keep[3]):
                for n, t in s.items():
                    t.copy_(saved[n])

        restore()
        from nanofed_b200.ops import native

        before = native.LAUNCHES
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=side):
            self._all_steps()
        self.launches_per_round = native.LAUNCHES - before
        restore()
        torch.cuda.synchronize(self.device)
        self._graph = graph


# ---------------------------------------------------------------------------
# Public entry point
# ---------------------------------------------------------------------------
class FedSimulation:
    """Runs federated rounds for ``clients_per_gpu × world_size`` simulated clients.

    >>> sim = FedSimulation(CohortConfig(model="mnist_cnn", clients_per_gpu=32))
    >>> out = sim.run_round()            # one FedAvg round, metrics as Python floats
    >>> sim.global_state()               # CPU state_dict, reference-loadable

    ``run_round(host_inputs=(images, labels))`` additionally stages that round's
    inputs from (pinned) host memory — the end-to-end path ``bench.py`` times.
    """

    def __init__(
        self,
        cfg: CohortConfig,
        shards: ClientShards | None = None,
        initial_state: dict[str, torch.Tensor] | None = None,
    ) -> None:
        self.cfg = cfg
        self.rank = rdist.init(cfg.device)
        self.device = self.rank.device
        self.logger = Logger()
        self.timer = PhaseTimer(enabled=True)
        torch.manual_seed(cfg.seed)  # same initial model on every rank

        template = build_model(cfg.model)
        if initial_state is not None:
            template.load_state_dict(initial_state)
        self.template = template
        self.layout = FlatLayout.from_module(template)
        c = cfg.clients_per_gpu

        engine = cfg.engine
        if engine == "auto":
            engine = "grouped" if (cfg.model in ("mnist_cnn", "resnet18", "resnet18_gn")
                                   and self.device.type == "cuda"
                                   and _grouped_available()) else "autograd"
        if cfg.optimizer not in ("sgd", "adam"):
            raise ValueError(f"optimizer must be 'sgd' or 'adam', got {cfg.optimizer!r}")
        if cfg.optimizer == "adam":
            if cfg.local_privacy is not None:
                raise ValueError("client-side DP-SGD (local_privacy) is defined for optimizer='sgd'")
            if cfg.model.startswith("resnet18"):
                # the ResNet grouped engine applies SGD in its weight-gradient epilogues / tap-major
                # update kernels; Adam runs on the generic engine
                if cfg.engine == "grouped":
                    raise ValueError("optimizer='adam' on ResNet-18 runs on engine='autograd'")
                engine = "autograd"
        if cfg.compute_dtype not in ("bf16", "fp32"):
            raise ValueError(f"compute_dtype must be 'bf16' or 'fp32', got {cfg.compute_dtype!r}")
        if cfg.compute_dtype == "fp32":
            # the hand-written engines feed bf16 operands to the tensor cores (fp32 accumulation,
            # fp32 master weights); a user who asks for fp32 compute gets the fp32 autograd engine,
            # never a silent bf16 run
            if cfg.engine == "grouped":
                raise ValueError("engine='grouped' computes in bf16 (fp32 accumulate/master); use "
                                 "compute_dtype='bf16', or engine='autograd' for fp32 compute")
            engine = "autograd"
        if engine == "grouped" and cfg.model not in ("mnist_cnn", "resnet18", "resnet18_gn"):
            raise ValueError(f"engine='grouped' has no hand-written network for {cfg.model!r} "
                             "(available: mnist_cnn, resnet18, resnet18_gn); use engine='autograd'")
        self.engine_name = engine

        self.arena = Arena(
            self.layout, c, self.device,
            with_momentum=cfg.momentum != 0.0 and cfg.optimizer == "sgd",
            with_lowp=(engine == "grouped" and cfg.model == "mnist_cnn"),
            with_grads=(engine == "autograd"),
            with_adam=cfg.optimizer == "adam",
        )
        self.arena.load_global(template.state_dict())

        if shards is None:
            shards = make_client_shards(
                c, cfg.samples_per_client,
                input_shape=tuple(getattr(template, "INPUT_SHAPE", (1, 28, 28))),
                num_classes=int(getattr(template, "NUM_CLASSES", 10)),
                seed=cfg.seed * 1000 + self.rank.rank, sample_counts=cfg.sample_counts,
            )
        self.shards = shards.to(self.device)
        self.arena.counts.copy_(self.shards.sample_counts.to(torch.float32))
        self._counts_host = [float(v) for v in shards.sample_counts.tolist()]

        if engine == "grouped" and cfg.model.startswith("resnet18"):
            from nanofed_b200.models.resnet_cohort import ResNetCohortNet

            self.engine: Any = ResNetCohortNet(
                cfg, self.arena, self.shards, client_base=self.rank.rank * c,
                clients_total=c * self.rank.world_size)
        elif engine == "grouped":
            from nanofed_b200.models.mnist_cohort import MnistCohortNet

            self.engine = MnistCohortNet(cfg, self.arena, self.shards)
        else:
            self.engine = AutogradCohort(cfg, self.arena, self.shards)

        self.reducer = None
        if self.rank.distributed:
            from nanofed_b200.parallel.peer import make_reducer

            self.reducer = make_reducer(cfg.transport, self.arena, self.rank,
                                        allow_fallback=cfg.allow_nccl_fallback)

        self.accountant = GaussianAccountant(cfg.privacy) if cfg.privacy else None
        self.local_accountant = GaussianAccountant(cfg.local_privacy) if cfg.local_privacy else None
        self._noise_offset = 0
        self._round = 0
        world = self.rank.world_size
        # sample counts of EVERY client of the job (weights of the job-wide round record)
        self._counts_all = [self._counts_host]
        if self.rank.distributed:
            import torch.distributed as dist

            gathered: list[Any] = [None] * world
            dist.all_gather_object(gathered, self._counts_host)
            self._counts_all = gathered
        # DP noise seed: unpredictable unless the user pins it (tests); one value for all ranks
        self._dp_seed = 0
        if cfg.privacy is not None:
            seed = cfg.dp_seed
            if seed is None:
                seed = int.from_bytes(os.urandom(8), "little") >> 1
            if self.rank.distributed:
                import torch.distributed as dist

                box = [seed]
                dist.broadcast_object_list(box, src=0)
                seed = int(box[0])
            self._dp_seed = seed
        self._mask_host = [1.0] * c          # host mirror of arena.mask (participation flags)
        self._k_active_cache: int | None = None
        # two pinned metric buffers: round r+1 may be submitted before round r is collected;
        # with several ranks they hold every rank's rows (gathered inside the aggregation kernel)
        pin = self.device.type == "cuda"
        self._stats_ring = [torch.zeros(world * c, 3, pin_memory=pin) for _ in range(2)]
        self._status_ring = [torch.zeros(1, dtype=torch.int32, pin_memory=pin) for _ in range(2)]
        self._k_active_ring = [self.num_clients, self.num_clients]
        self._stats_host = self._stats_ring[0]
        self._pending: list[PendingRound | None] = [None, None]
        self.coordinator = None

    # -- properties ------------------------------------------------------------
    @property
    def num_clients(self) -> int:
        return self.cfg.clients_per_gpu * self.rank.world_size

    @property
    def samples_per_round(self) -> int:
        """Client-samples trained per round over the whole job."""
        return self.num_clients * self.cfg.local_steps * self.cfg.batch_size

    def attach(self, coordinator: Any) -> None:
        self.coordinator = coordinator

    # -- one round -------------------------------------------------------------
    def stage_inputs(self, images: torch.Tensor, labels: torch.Tensor) -> int:
        """H2D copy of this round's inputs into the resident shard tensors; returns bytes."""
        self.shards.images.copy_(images, non_blocking=True)
        self.shards.labels.copy_(labels, non_blocking=True)
        return images.numel() * images.element_size() + labels.numel() * labels.element_size()

    # -- input pipeline: H2D of round r+1 overlaps the training of round r -------------------
    def prefetch_inputs(self, images: torch.Tensor, labels: torch.Tensor) -> int:
        """Start the H2D copy of the NEXT round's inputs (pinned host tensors) on a copy stream.

        The data lands in a staging buffer; ``run_round(use_staged=True)`` moves it into the
        resident shards with a device-to-device copy (the CUDA graph of the local epoch has the
        shard addresses baked in).  Returns the number of bytes that will cross PCIe.
        """
        if self.device.type != "cuda":
            return self.stage_inputs(images, labels)
        if getattr(self, "_stage_images", None) is None:
            self._stage_images = torch.empty_like(self.shards.images)
            self._stage_labels = torch.empty_like(self.shards.labels)
            self._copy_stream = torch.cuda.Stream(device=self.device)
            self._staged = torch.cuda.Event()
            self._consumed = torch.cuda.Event()
            self._consumed.record()
        self._copy_stream.wait_event(self._consumed)      # staging is free again
        with torch.cuda.stream(self._copy_stream):
            self._stage_images.copy_(images, non_blocking=True)
            self._stage_labels.copy_(labels, non_blocking=True)
            self._staged.record(self._copy_stream)
        self._has_staged = True
        return images.numel() * images.element_size() + labels.numel() * labels.element_size()

    def _adopt_staged(self) -> None:
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(self._staged)
        self.shards.images.copy_(self._stage_images, non_blocking=True)
        self.shards.labels.copy_(self._stage_labels, non_blocking=True)
        self._consumed.record(cur)
        self._has_staged = False

    def enqueue_round(self) -> None:
        """Enqueue one full round on the current stream.  No host synchronisation."""
        arena, cfg = self.arena, self.cfg
        arena.stats.zero_()
        with self.timer.phase("bcast"):
            if not getattr(self, "_slots_fresh", False):
                ops.bcast_global(arena.master, arena.slots, arena.slots_lowp, arena.momentum)
            self._slots_fresh = False
            self.engine.after_broadcast()
        late = self._train_with_deadline()
        with self.timer.phase("aggregate"):
            self.engine.before_aggregate()
            poisoned = bool(getattr(self, "_poison", []))
            for k in getattr(self, "_poison", []):
                arena.slots[k, :16] = float("nan")
            self._poison = []
            dp = cfg.privacy
            kw: dict[str, Any] = {}
            # K-STATS: every client's sum of squares (DP clipping) and non-finite count.  The MNIST
            # engine's last local step already left the sums in arena.sumsq (emits_sumsq).
            side_output = (getattr(self.engine, "emits_sumsq", False) and not cfg.screen_nonfinite
                           and not poisoned)
            if (dp is not None and not side_output) or cfg.screen_nonfinite:
                ss, bad = ops.update_stats(arena.slots)
                arena.sumsq.copy_(ss)
                arena.nonfinite.copy_(bad)
                if cfg.screen_nonfinite:
                    # device-side failure detection: a diverged client drops out of this
                    # round's average without any host round trip (K-AGG never reads its row)
                    self._round_mask = arena.mask * (bad == 0).to(arena.mask.dtype)
            mask = getattr(self, "_round_mask", None) if cfg.screen_nonfinite else None
            if late:                     # this cohort missed the deadline: none of its clients count
                mask = torch.zeros_like(arena.mask)
            self._k_active_round = self.num_clients
            if dp is not None:
                # the mechanism is calibrated for the updates that are actually averaged
                # (reference: aggregator/privacy.py uses len(updates)), not for the configured count
                k_active = self._active_clients(mask, late)
                self._k_active_round = k_active
                kw = dict(sumsq=arena.sumsq, clip_norm=float(dp.max_gradient_norm),
                          noise_std=dp.noise_multiplier * dp.max_gradient_norm / max(k_active, 1),
                          noise_mask=arena.noise_mask, seed=self._dp_seed,
                          offset=self._noise_offset)
                self._noise_offset += arena.layout.total
            if mask is not None:
                # in place: the fused multi-GPU kernel holds the address of arena.mask
                saved = arena.mask.clone()
                arena.mask.copy_(mask)
            if self.reducer is None:
                fuse = cfg.fuse_bcast and self.device.type == "cuda" and not cfg.screen_nonfinite
                if fuse:     # the client rows leave this launch holding the NEW global model
                    kw["bcast"] = (arena.slots, arena.slots_lowp, arena.momentum)
                ops.fedavg_reduce(arena.slots, arena.counts, out=arena.master,
                                  out_lowp=arena.master_lowp, mask=arena.mask, **kw)
                self._slots_fresh = fuse
            else:
                fuse = (cfg.fuse_bcast and self.device.type == "cuda" and not cfg.screen_nonfinite)
                if fuse:     # K-BCAST rides in the tail of the aggregation kernel
                    kw["bcast"] = (arena.slots, arena.slots_lowp, arena.momentum)
                self.reducer.reduce(arena, **kw)
                self._slots_fresh = fuse
            if mask is not None:
                arena.mask.copy_(saved)

    def _active_clients(self, round_mask: torch.Tensor | None, late: bool) -> int:
        """Number of clients, over the whole job, whose update enters this round's average."""
        screened = self.cfg.screen_nonfinite and round_mask is not None and not late
        if late:
            local = 0
        elif screened:
            local = int((round_mask > 0).sum().item())       # decided on the device this round
        else:
            local = sum(1 for m in self._mask_host if m > 0)
        if not self.rank.distributed:
            return local
        dynamic = screened or self.cfg.round_deadline_ms is not None
        if not dynamic and self._k_active_cache is not None:
            return self._k_active_cache
        import torch.distributed as dist

        t = torch.tensor([float(local)], device=self.device)
        dist.all_reduce(t)
        total = int(round(float(t.item())))
        if not dynamic:
            self._k_active_cache = total
        return total

    def _train_with_deadline(self) -> bool:
        """Local training of this rank's cohort; returns True if the cohort is a straggler that
        must be left out of this round (``CohortConfig.round_deadline_ms``)."""
        import time

        cfg = self.cfg
        delay_ms, self._delay_ms = getattr(self, "_delay_ms", 0.0), 0.0
        deadline = cfg.round_deadline_ms
        cuda = self.device.type == "cuda"
        if deadline is not None:
            if cuda:
                t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                t0.record()
            else:
                start = time.perf_counter()
        with self.timer.phase("local_train"):
            if delay_ms > 0:                     # injected fault: a slow cohort
                if cuda:
                    torch.cuda._sleep(int(delay_ms * 1.9e6))   # ~cycles at 1.9 GHz
                else:
                    time.sleep(delay_ms / 1e3)
            self.engine.local_steps()
        self.last_train_ms = None
        self.stragglers_last_round: list[str] = []
        if deadline is None:
            return False
        if cuda:
            t1.record()
            t1.synchronize()
            elapsed = t0.elapsed_time(t1)
        else:
            elapsed = (time.perf_counter() - start) * 1e3
        self.last_train_ms = elapsed
        late = elapsed > deadline
        on_time = torch.tensor([0.0 if late else 1.0], device=self.device)
        if self.rank.distributed:
            import torch.distributed as dist

            dist.all_reduce(on_time)
        if late and float(on_time.item()) == 0.0:
            self.logger.warning(f"round {self._round}: every cohort missed the {deadline} ms "
                                "deadline; keeping all of them")
            return False
        if late:
            base = self.rank.rank * cfg.clients_per_gpu
            self.stragglers_last_round = [f"client_{base + k}" for k in range(cfg.clients_per_gpu)]
            self.logger.warning(f"round {self._round}: cohort of rank {self.rank.rank} took "
                                f"{elapsed:.1f} ms > {deadline} ms; its clients are left out")
        return late

    def run_round(
        self,
        host_inputs: tuple[torch.Tensor, torch.Tensor] | None = None,
        use_staged: bool = False,
        prefetch_next: tuple[torch.Tensor, torch.Tensor] | None = None,
    ) -> RoundOutcome:
        """One federated round; returns metrics as Python floats (one stream sync).

        ``host_inputs``: copy this round's inputs from host memory first (blocking the
        compute stream).  ``use_staged``: use inputs previously sent with ``prefetch_inputs``.
        ``prefetch_next``: start the H2D copy of the following round's inputs, overlapped
        with this round's training.
        """
        return self.collect(self.submit_round(host_inputs, use_staged, prefetch_next))

    def submit_round(
        self,
        host_inputs: tuple[torch.Tensor, torch.Tensor] | None = None,
        use_staged: bool = False,
        prefetch_next: tuple[torch.Tensor, torch.Tensor] | None = None,
    ) -> "PendingRound":
        """Enqueue one round and the device→host copy of its metrics, without waiting.

        ``collect(handle)`` waits for that copy and returns the ``RoundOutcome``.  Submitting
        round r+1 before collecting round r keeps the GPU busy while the host digests metrics
        (two pinned metric buffers alternate, so at most two rounds may be outstanding).
        """
        if host_inputs is not None:
            self.stage_inputs(*host_inputs)
        elif use_staged and getattr(self, "_has_staged", False):
            self._adopt_staged()
        if prefetch_next is not None:
            self.prefetch_inputs(*prefetch_next)
        slot = self._round % 2
        if self._pending[slot] is not None:
            raise RuntimeError("collect() the round submitted two rounds ago before submitting "
                               "another one (two metric buffers)")
        self.enqueue_round()
        # the round's only device→host transfer: per-client (loss_sum, correct, samples) of every
        # rank (the aggregation kernel left all ranks' rows in this rank's symmetric buffer) plus
        # the kernel's barrier-timeout word
        c = self.cfg.clients_per_gpu
        if self.reducer is not None:
            self._stats_ring[slot].copy_(self.reducer.all_stats().reshape(-1, 3), non_blocking=True)
            if getattr(self.reducer, "status", None) is not None:
                self._status_ring[slot].copy_(self.reducer.status, non_blocking=True)
        else:
            self._stats_ring[slot][:c].copy_(self.arena.stats, non_blocking=True)
        self._k_active_ring[slot] = getattr(self, "_k_active_round", self.num_clients)
        event = None
        if self.device.type == "cuda":
            event = torch.cuda.Event()
            event.record(torch.cuda.current_stream(self.device))
        handle = PendingRound(self._round, slot, event)
        self._pending[slot] = handle
        self._round += 1
        return handle

    def collect(self, handle: "PendingRound") -> RoundOutcome:
        """Wait for a submitted round's metrics and return them as Python floats."""
        if self._pending[handle.slot] is not handle:
            raise RuntimeError("round was already collected")
        if handle.event is not None:
            handle.event.synchronize()
        self._pending[handle.slot] = None
        if int(self._status_ring[handle.slot][0]) != 0:
            # a peer launched late / died or a flag wait exceeded its bound: the kernel returned
            # with the global model partly written - never train or checkpoint on top of that
            self._status_ring[handle.slot].zero_()
            self.reducer.status.zero_()
            raise RuntimeError(
                f"round {handle.round_id}: the cross-GPU aggregation kernel timed out waiting for "
                "a peer; the global model of this round is incomplete")
        outcome = self._summarise(handle.round_id, self._stats_ring[handle.slot])
        if self.accountant is not None:
            k_active = max(int(self._k_active_ring[handle.slot]), 1)
            for _ in range(k_active):
                self.accountant.add_noise_event(self.cfg.privacy.noise_multiplier, k_active)
            spent = self.accountant.get_privacy_spent()
            outcome.agg_metrics["privacy_epsilon"] = spent.epsilon_spent
            outcome.agg_metrics["privacy_delta"] = spent.delta_spent
        if self.local_accountant is not None:
            # every client ran local_steps noisy steps on batches of batch_size (one budget per
            # client, identical for all of them: reference PrivateTrainer._add_noise)
            for _ in range(self.cfg.local_steps):
                self.local_accountant.add_noise_event(self.cfg.local_privacy.noise_multiplier,
                                                      self.cfg.batch_size)
            spent = self.local_accountant.get_privacy_spent()
            outcome.agg_metrics["local_privacy_epsilon"] = spent.epsilon_spent
            outcome.agg_metrics["local_privacy_delta"] = spent.delta_spent
        return outcome

    def _summarise(self, round_id: int, stats: torch.Tensor | None = None) -> RoundOutcome:
        """Job-wide round record: one entry per client of EVERY rank with its FedAvg weight, like
        the reference's coordinator writes (``orchestration/coordinator.py:324-340``)."""
        stats = self._stats_host if stats is None else stats
        c = self.cfg.clients_per_gpu
        ranks = range(self.rank.world_size) if self.reducer is not None else [self.rank.rank]
        counts = {r: self._counts_all[r if self.reducer is not None else 0] for r in ranks}
        total = sum(sum(v) for v in counts.values())
        client_metrics = []
        agg = {"loss": 0.0, "accuracy": 0.0, "samples_processed": 0.0}
        per_epoch = self.cfg.steps_per_epoch * self.cfg.batch_size
        rows = stats.tolist()
        for r in ranks:
            row0 = r * c if self.reducer is not None else 0
            for k in range(c):
                loss_sum, correct, seen = rows[row0 + k]
                seen = max(seen, 1.0)
                m = {"loss": loss_sum / seen, "accuracy": correct / seen,
                     "samples_processed": int(min(seen, per_epoch))}
                w = counts[r][k] / total
                client_metrics.append({"client_id": f"client_{r * c + k}", "metrics": m,
                                       "weight": w})
                for name in agg:
                    agg[name] += w * float(m[name])
        n_clients = len(client_metrics)
        return RoundOutcome(round_id, n_clients, agg, client_metrics,
                            samples=n_clients * self.cfg.local_steps * self.cfg.batch_size)

    # -- participation / fault injection ----------------------------------
    def set_participation(self, active: list[bool] | torch.Tensor) -> None:
        """Per-client participation flags for the coming rounds (device tensor read by K-AGG;
        weights are renormalised over the participants on the device)."""
        flags = torch.as_tensor(active, dtype=torch.float32)
        self._mask_host = [float(v) for v in flags.tolist()]
        self._k_active_cache = None
        self.arena.mask.copy_(flags.to(self.device))

    def drop_clients(self, indices: list[int]) -> None:
        for i in indices:
            self._mask_host[int(i)] = 0.0
        self._k_active_cache = None
        self.arena.mask[torch.as_tensor(indices, device=self.device)] = 0.0

    def inject_fault(self, kind: str, client: int, delay_ms: float = 50.0) -> None:
        """Test hook: ``"drop"`` masks the client out, ``"nan"`` poisons its weights AFTER the
        next local training (use with ``screen_nonfinite=True`` to see it rejected), ``"delay"``
        slows the client's cohort down by ``delay_ms`` (use with ``round_deadline_ms``)."""
        if kind == "drop":
            self.drop_clients([client])
        elif kind == "nan":
            self._poison = getattr(self, "_poison", []) + [client]
        elif kind == "delay":
            # a straggler: the client's cohort (its GPU) is held back by ``delay_ms`` in the next
            # round (clients of one GPU train in lock-step)
            self._delay_ms = float(delay_ms)
        else:
            raise ValueError(f"unknown fault kind {kind!r}")

    # -- checkpoints -------------------------------------------------------------
    def load_global_state(self, state: dict[str, torch.Tensor]) -> None:
        """Resume: overwrite the global model (e.g. ``ModelManager.load_model`` + ``state_dict``)."""
        host = self.layout.flatten(state, device="cpu")
        self.arena.master.copy_(host.to(self.device))
        self.arena.master_lowp.copy_(self.arena.master)
        self._slots_fresh = False        # the client rows no longer hold this model

    def global_state(self) -> dict[str, torch.Tensor]:
        return self.arena.global_state()

    def sync_model(self, model: nn.Module | None = None) -> nn.Module:
        """Copy the arena's global model into ``model`` (default: the template)."""
        model = model or self.template
        model.load_state_dict(self.global_state())
        return model

    # ``Coordinator(engine=sim)`` calls this from a worker thread.
    def run_round_for_coordinator(self, round_id: int) -> RoundOutcome:
        if self.device.type == "cuda":
            torch.cuda.set_device(self.device)
        outcome = self.run_round()
        if self.coordinator is not None:
            self.sync_model(self.coordinator.model_manager.model)
        return outcome


def _grouped_available() -> bool:
    try:
        from nanofed_b200.models import mnist_cohort  # noqa: F401

        return mnist_cohort.AVAILABLE
    except Exception:
        return False

"""ResNet-18 (CIFAR shape) federated training on one box of B200s: BASELINE configs 3 and 4.

    python examples/resnet18/run_simulation.py [--rounds 5] [--dp] [--group-norm] [--local-dp]
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 \
        examples/resnet18/run_simulation.py --rounds 5          # 64 clients on 8 GPUs

Every client is a row of the per-GPU arena; a local step of all clients of a GPU is ~110 launches
of this repository's kernels (tcgen05 implicit-GEMM convolutions, BatchNorm, head, SGD), a local
epoch is one CUDA graph, and the round ends in the fused FedAvg kernel (across GPUs: reduce-scatter
/ all-gather over NVLink from inside that kernel).  ``--dp`` turns on DP-FedAvg: per-update L2
clipping and Gaussian noise inside the aggregation kernel; ``--group-norm`` trains the GroupNorm(8)
variant and ``--local-dp`` adds client-side DP-SGD (clip + noise in every local step, inside the
update kernels).  The reference has no counterpart of
this script (its only example is ``examples/mnist/run_experiment.py`` over HTTP); the round records
and checkpoints written through ``Coordinator(engine=...)`` keep the reference's on-disk layout.
"""

from __future__ import annotations

import argparse
import asyncio
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[2]))

from nanofed_b200 import (Coordinator, CoordinatorConfig, FedAvgAggregator,  # noqa: E402
                          ModelManager)
from nanofed_b200.models import ResNet18  # noqa: E402
from nanofed_b200.parallel.cohort import CohortConfig, FedSimulation  # noqa: E402
from nanofed_b200.privacy import PrivacyConfig  # noqa: E402


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--rounds", type=int, default=5)
    ap.add_argument("--clients-per-gpu", type=int, default=8)
    ap.add_argument("--samples-per-client", type=int, default=512)
    ap.add_argument("--dp", action="store_true")
    ap.add_argument("--group-norm", action="store_true")
    ap.add_argument("--local-dp", action="store_true")
    ap.add_argument("--out", default="runs/resnet18")
    args = ap.parse_args()
    name = "resnet18_gn" if args.group_norm else "resnet18"
    cfg = CohortConfig(model=name, clients_per_gpu=args.clients_per_gpu,
                       samples_per_client=args.samples_per_client, batch_size=64, learning_rate=0.05,
                       privacy=PrivacyConfig(epsilon=10.0, noise_multiplier=1.1) if args.dp else None,
                       local_privacy=PrivacyConfig(epsilon=10.0, noise_multiplier=1.1)
                       if args.local_dp else None)
    sim = FedSimulation(cfg)                      # synthetic CIFAR-shaped shards, resident on the GPU
    if not sim.rank.is_main:                      # other ranks: same rounds, no files
        for _ in range(args.rounds):
            sim.run_round()
        return
    model = ResNet18(norm="group" if args.group_norm else "batch")
    coordinator = Coordinator(ModelManager(model), FedAvgAggregator(), None,
                              CoordinatorConfig(num_rounds=args.rounds, min_clients=sim.num_clients,
                                                min_completion_rate=1.0, round_timeout=600,
                                                base_dir=Path(args.out)), engine=sim)

    async def run() -> None:
        async for m in coordinator.start_training():
            print(f"round {m.round_id}: {m.num_clients} clients, loss {m.agg_metrics['loss']:.4f}, "
                  f"accuracy {m.agg_metrics['accuracy']:.3f}")

    asyncio.run(run())
    print(f"engine: {sim.engine_name} ({type(sim.engine).__name__}); checkpoints under {args.out}")


if __name__ == "__main__":
    main()

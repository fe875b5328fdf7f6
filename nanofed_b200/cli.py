"""Command line (the reference declares ``nanofed = "nanofed.cli:main"`` in
``pyproject.toml:23`` but ships no ``cli.py``; this one exists).

    python -m nanofed_b200.cli simulate --model mnist_cnn --clients-per-gpu 32 --rounds 5
    python -m nanofed_b200.cli serve --port 8080 --rounds 2 --min-clients 3 --base-dir runs/
    python -m nanofed_b200.cli bench [bench.py flags…]
    python -m nanofed_b200.cli build [--force]
"""

from __future__ import annotations

import argparse
import asyncio
import json
import runpy
import sys
from pathlib import Path


def _simulate(args: argparse.Namespace) -> int:
    from nanofed_b200.parallel.cohort import CohortConfig, FedSimulation
    from nanofed_b200.privacy import PrivacyConfig

    cfg = CohortConfig(
        model=args.model, clients_per_gpu=args.clients_per_gpu,
        samples_per_client=args.samples_per_client, batch_size=args.batch_size,
        local_epochs=args.local_epochs, learning_rate=args.lr, engine=args.engine,
        optimizer=args.optimizer, momentum=args.momentum, weight_decay=args.weight_decay,
        compute_dtype=args.dtype, device=args.device, round_deadline_ms=args.round_deadline_ms,
        privacy=PrivacyConfig(noise_multiplier=args.noise_multiplier,
                              max_gradient_norm=args.clip_norm) if args.dp else None,
        local_privacy=PrivacyConfig(noise_multiplier=args.noise_multiplier,
                                    max_gradient_norm=args.clip_norm) if args.local_dp else None,
    )
    sim = FedSimulation(cfg)
    for r in range(args.rounds):
        out = sim.run_round()
        if sim.rank.is_main:
            print(json.dumps({"round": r, **{k: round(v, 6) for k, v in out.agg_metrics.items()}}))
    if args.save and sim.rank.is_main:
        import torch

        torch.save(sim.global_state(), args.save)
    return 0


def _serve(args: argparse.Namespace) -> int:
    from nanofed_b200 import (Coordinator, CoordinatorConfig, FedAvgAggregator, HTTPServer,
                              ModelManager, coordinate)
    from nanofed_b200.parallel.cohort import build_model

    async def main() -> None:
        server = HTTPServer(args.host, args.port, max_request_size=args.max_request_mb << 20)
        await server.start()
        coordinator = Coordinator(
            ModelManager(build_model(args.model)), FedAvgAggregator(), server,
            CoordinatorConfig(num_rounds=args.rounds, min_clients=args.min_clients,
                              min_completion_rate=args.min_completion_rate,
                              round_timeout=args.round_timeout, base_dir=Path(args.base_dir)))
        try:
            await coordinate(coordinator)
        finally:
            await server.stop()

    asyncio.run(main())
    return 0


def main(argv: list[str] | None = None) -> int:
    ap = argparse.ArgumentParser(prog="nanofed_b200", description=__doc__,
                                 formatter_class=argparse.RawDescriptionHelpFormatter)
    sub = ap.add_subparsers(dest="cmd", required=True)

    s = sub.add_parser("simulate", aliases=["train"],
                       help="in-box federated simulation on the local GPUs")
    s.add_argument("--model", default="mnist_cnn")
    s.add_argument("--clients-per-gpu", type=int, default=32)
    s.add_argument("--samples-per-client", type=int, default=512)
    s.add_argument("--batch-size", type=int, default=64)
    s.add_argument("--local-epochs", type=int, default=1)
    s.add_argument("--lr", type=float, default=0.1)
    s.add_argument("--rounds", type=int, default=5)
    s.add_argument("--optimizer", default="sgd", choices=["sgd", "adam"])
    s.add_argument("--momentum", type=float, default=0.0)
    s.add_argument("--weight-decay", type=float, default=0.0)
    s.add_argument("--engine", default="auto")
    s.add_argument("--dtype", default="bf16")
    s.add_argument("--device", default=None)
    s.add_argument("--dp", action="store_true",
                   help="central DP: per-update clipping + Gaussian noise inside the aggregation")
    s.add_argument("--local-dp", action="store_true",
                   help="client-side DP-SGD: clip + noise in every local step (PrivateTrainer recipe)")
    s.add_argument("--noise-multiplier", type=float, default=1.1)
    s.add_argument("--clip-norm", type=float, default=1.0)
    s.add_argument("--round-deadline-ms", type=float, default=None,
                   help="leave a GPU's cohort out of a round when its local training takes longer")
    s.add_argument("--save", default=None, help="write the final global state_dict here")
    s.set_defaults(fn=_simulate)

    v = sub.add_parser("serve", help="cross-host HTTP server + coordinator")
    v.add_argument("--model", default="mnist_cnn")
    v.add_argument("--host", default="0.0.0.0")
    v.add_argument("--port", type=int, default=8080)
    v.add_argument("--rounds", type=int, default=2)
    v.add_argument("--min-clients", type=int, default=3)
    v.add_argument("--min-completion-rate", type=float, default=1.0)
    v.add_argument("--round-timeout", type=int, default=300)
    v.add_argument("--base-dir", default="runs")
    v.add_argument("--max-request-mb", type=int, default=100)
    v.set_defaults(fn=_serve)

    b = sub.add_parser("bench", help="run bench.py (remaining arguments are passed through)")
    b.set_defaults(fn=None)
    k = sub.add_parser("build", help="compile the sm_100a library")
    k.add_argument("--force", action="store_true")
    k.set_defaults(fn=None)

    args, rest = ap.parse_known_args(argv)
    if args.cmd == "bench":
        sys.argv = ["bench.py", *rest]
        runpy.run_path(str(Path(__file__).resolve().parent.parent / "bench.py"), run_name="__main__")
        return 0
    if args.cmd == "build":
        from nanofed_b200.build import build

        print(build(force=args.force))
        return 0
    return int(args.fn(args) or 0)


if __name__ == "__main__":
    raise SystemExit(main())

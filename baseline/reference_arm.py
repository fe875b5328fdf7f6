"""``bench.py --impl reference``: the UNMODIFIED reference on the benchmark config.

The reference (camille-004/nanofed, installed offline into ``baseline/_ref`` — see
DESIGN.md) is driven through its own public API and stock code path, the way its
example does (``examples/mnist/run_experiment.py``): ``HTTPServer`` + ``Coordinator`` +
``FedAvgAggregator`` + ``ModelManager`` on rank 0, and per rank ``clients_per_gpu``
client coroutines doing ``HTTPClient.fetch_global_model`` → fresh ``MNISTModel`` + SGD →
``TorchTrainer.train_epoch`` (``device="cuda"``) → ``HTTPClient.submit_update``.  None of
this repository's models, kernels or engine is imported here.

Deviations, all outside the reference's code:
  1. ``privacy_spent`` shim — at HEAD ``Coordinator.train_round`` reads
     ``update["privacy_spent"]`` which the server never stores, so every HTTP round dies
     with ``KeyError``.  ``server._updates`` is replaced by a ``dict`` subclass that adds
     ``"privacy_spent": None`` on insert.  No reference source is edited.
  2. synthetic data of the named shape (``load_mnist_data`` needs the network).
  3. a wall-clock budget: one reference round costs minutes (27 MB of JSON per client
     per direction through one event loop), so the number of warm-up/timed rounds is
     reduced to fit ``NANOFED_REF_BUDGET_S`` (default 600 s) and the JSON line reports the
     counts actually used.
"""

from __future__ import annotations

import asyncio
import json
import os
import sys
import tempfile
import time
from pathlib import Path

REF_DIR = Path(__file__).resolve().parent / "_ref"


_REF_INFO: dict = {}


def _unavailable(why: str) -> None:
    print(json.dumps({"impl": "reference", "unavailable": why}), flush=True)


class _ShimUpdates(dict):
    """Deviation 1: supply the key the reference's coordinator expects."""

    def __setitem__(self, key, value):
        try:
            value.setdefault("privacy_spent", None)
        except AttributeError:
            pass
        super().__setitem__(key, value)


def run_reference(args) -> None:
    global _REF_INFO
    try:
        from baseline.install_ref import install

        manifest = install()          # verifies an existing install, installs when it can
    except Exception as exc:
        _unavailable(f"baseline/_ref unavailable: {exc}"[:300])
        return
    _REF_INFO = {"ref_source": manifest["how"], "ref_sha": manifest["tree_sha256"],
                 "ref_identical_to_source": manifest["identical_to_source"]}
    sys.path.insert(0, str(REF_DIR))
    # make sure `nanofed` is the reference, not this repo's compat alias
    for name in [m for m in sys.modules if m == "nanofed" or m.startswith("nanofed.")]:
        del sys.modules[name]
    try:
        import torch
        import nanofed  # noqa: F401
    except Exception as exc:  # pragma: no cover
        _unavailable(f"import of the reference failed: {exc!r}"[:200])
        return
    if REF_DIR not in Path(nanofed.__file__).resolve().parents:
        _unavailable(f"'nanofed' resolved to {nanofed.__file__}, not baseline/_ref")
        return
    try:
        asyncio.run(_run(args, torch))
    except Exception as exc:
        _unavailable(f"reference run failed: {exc!r}"[:300])


async def _run(args, torch) -> None:
    import logging

    from torch.utils.data import DataLoader, TensorDataset

    from nanofed import (Coordinator, CoordinatorConfig, FedAvgAggregator, HTTPClient,
                         HTTPServer, ModelManager, TorchTrainer, TrainingConfig)
    from nanofed.models import MNISTModel

    from nanofed.utils import LogConfig, Logger, LogLevel

    # quiet the reference through its own public logging configuration
    Logger().configure(LogConfig(level=LogLevel.ERROR, color=False, format="", output="console"))
    logging.getLogger("nanofed").setLevel(logging.ERROR)
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    cuda = torch.cuda.is_available()
    device = f"cuda:{local_rank}" if cuda else "cpu"
    if cuda:
        torch.cuda.set_device(local_rank)
    port = int(os.environ.get("MASTER_PORT", "29500")) + 1017
    url = f"http://127.0.0.1:{port}"
    clients_per_gpu = args.clients_per_gpu
    samples = args.samples_per_client
    total_clients = clients_per_gpu * world
    budget = float(os.environ.get("NANOFED_REF_BUDGET_S", "600"))
    want_warm, want_steps = max(args.warmup, 0), max(args.steps, 1)
    num_rounds = want_warm + want_steps

    torch.manual_seed(0)
    state = {"stop": False}

    async def client(cid: int) -> None:
        gen = torch.Generator().manual_seed(1000 + cid)
        x = torch.randn(samples, 1, 28, 28, generator=gen)
        y = torch.randint(0, 10, (samples,), generator=gen)
        loader = DataLoader(TensorDataset(x, y), batch_size=64, shuffle=True)
        cfg = TrainingConfig(epochs=1, batch_size=64, learning_rate=0.1, device=device,
                             log_interval=10**9)
        trainer = TorchTrainer(cfg)
        async with HTTPClient(server_url=url, client_id=f"client_{cid}") as http:
            while not state["stop"]:
                try:
                    if await http.check_server_status():
                        break
                    weights, _ = await http.fetch_global_model()
                    model = MNISTModel()
                    model.load_state_dict(weights)
                    model.to(device)
                    opt = torch.optim.SGD(model.parameters(), lr=cfg.learning_rate)
                    metrics = None
                    for epoch in range(cfg.epochs):
                        metrics = trainer.train_epoch(model, loader, opt, epoch)
                    if metrics and not await http.submit_update(model, metrics):
                        break
                except Exception:
                    await asyncio.sleep(0.5)
                    if state["stop"]:
                        break

    round_marks: list[float] = []

    async def serve() -> None:
        base = Path(tempfile.mkdtemp(prefix="nanofed_ref_"))
        model = MNISTModel()
        manager = ModelManager(model=model)
        server = HTTPServer(host="127.0.0.1", port=port, max_request_size=1 << 30)
        server._updates = _ShimUpdates()
        await server.start()
        coord = Coordinator(
            model_manager=manager, aggregator=FedAvgAggregator(), server=server,
            config=CoordinatorConfig(num_rounds=num_rounds, min_clients=total_clients,
                                     min_completion_rate=1.0, round_timeout=24 * 3600,
                                     base_dir=base))
        t0 = time.perf_counter()
        round_marks.append(t0)
        try:
            async for _ in coord.start_training():
                if cuda:
                    torch.cuda.synchronize()
                now = time.perf_counter()
                round_marks.append(now)
                done = len(round_marks) - 1
                per_round = (now - t0) / done
                # stop early if the remaining rounds do not fit the budget, keeping >= 1 timed
                if done >= 1 and (now - t0) + per_round > budget and done < num_rounds:
                    break
        finally:
            state["stop"] = True
            await server.stop_training()
            await asyncio.sleep(1.5)  # let clients observe the termination flag
            await server.stop()

    tasks = [client(rank * clients_per_gpu + i) for i in range(clients_per_gpu)]
    if rank == 0:
        tasks.append(serve())
    else:
        await asyncio.sleep(3.0)  # server start-up on rank 0
    # hard stop: a reference round cannot be interrupted from outside its own API, so a run whose
    # rounds do not finish at all (very large client counts) is cut off after 4x the budget and
    # reports whatever rounds completed (or "unavailable" if none did)
    hard_cap = float(os.environ.get("NANOFED_REF_HARD_CAP_S", str(4 * budget)))
    try:
        await asyncio.wait_for(asyncio.gather(*tasks, return_exceptions=True), timeout=hard_cap)
    except asyncio.TimeoutError:
        state["stop"] = True

    if rank != 0:
        return
    durations = [b - a for a, b in zip(round_marks, round_marks[1:])]
    if not durations:
        _unavailable("no reference round completed")
        return
    warm = min(want_warm, len(durations) - 1)
    timed = durations[warm:]
    sec = sum(timed) / len(timed)
    samples_per_round = total_clients * samples * 1
    note = None
    if len(durations) < num_rounds:
        note = (f"budget {budget:.0f}s: ran {len(durations)} of {num_rounds} requested rounds "
                f"({warm} warm-up + {len(timed)} timed)")
    print(json.dumps({
        "metric": "client_samples_per_sec", "value": samples_per_round / sec,
        "unit": "client_samples/s", "n_gpus": args.gpus, "steps": len(timed), "warmup": warm,
        "ms_per_step": sec * 1e3, "rounds_per_sec": 1.0 / sec, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, **_REF_INFO,
        "dtype": "fp32", "data": "synthetic", "impl": "reference",
        "config": {"model": "mnist_cnn (1199882 params, random init)",
                   "global_batch": total_clients * 64, "seq_len": None,
                   "clients": total_clients, "clients_per_gpu": clients_per_gpu,
                   "samples_per_client": samples, "local_epochs": 1, "local_batch": 64,
                   "optimizer": "SGD(lr=0.1)", "aggregation": "FedAvg (sample-weighted)",
                   "parallelism": f"clients-dp{args.gpus}", "transport": "HTTP+JSON (stock)",
                   "device": device, "timing": "wall clock between completed rounds",
                   "deviations": ["privacy_spent shim", "synthetic data", "round budget"]},
        "e2e": {"value": samples_per_round / sec, "unit": "client_samples/s",
                "note": "the reference path is end-to-end by construction"},
        "gpu_launches": 0, "note": note,
    }), flush=True)

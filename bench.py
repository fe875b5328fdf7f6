#!/usr/bin/env python
"""Headline benchmark: FL rounds/s and client-samples/s (BASELINE.json config 2).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

Config (both arms): MNIST CNN (1 199 882 params, random init), FedAvg, 32 simulated
clients per GPU (256 on 8 GPUs — weak scaling), 512 synthetic 28×28 samples per client,
1 local epoch of batch-64 SGD(lr 0.1) ⇒ 8 local steps per client per round.  A "step" of
the benchmark is ONE FEDERATED ROUND: broadcast → local training of every client →
sample-weighted aggregation (optimizer steps and aggregation are inside the timed region).

``value`` is client-samples/s over the whole job, timed on the device with CUDA events,
max over ranks; ``rounds_per_sec`` rides along.  ``e2e`` repeats the measurement through
the public API (``FedSimulation.run_round(host_inputs=…)``) including, every round, the
H2D copy of that round's inputs from pinned host memory and the D2H read of the round's
metrics.  ``--impl reference`` runs the unmodified reference from ``baseline/_ref``
(HTTP + JSON + its own Coordinator/Trainer/Aggregator) on the same config.
"""

from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

CLIENTS_PER_GPU = 32
SAMPLES_PER_CLIENT = 512
BATCH = 64
LOCAL_EPOCHS = 1
LR = 0.1


def parse() -> argparse.Namespace:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--engine", default="auto", help="auto | grouped | autograd")
    ap.add_argument("--transport", default="peer", help="peer | nccl (baseline)")
    ap.add_argument("--clients-per-gpu", type=int, default=None,
                    help="default: 32 (mnist_cnn), 8 (resnet18: 64 clients on 8 GPUs)")
    ap.add_argument("--samples-per-client", type=int, default=SAMPLES_PER_CLIENT)
    ap.add_argument("--model", default="mnist_cnn")
    ap.add_argument("--dtype", default="bf16")
    ap.add_argument("--dp", action="store_true", help="central DP fused into aggregation")
    ap.add_argument("--local-dp", action="store_true",
                    help="client-side DP-SGD (clip + noise inside K-OPT) in every local step")
    ap.add_argument("--optimizer", default="sgd", choices=["sgd", "adam"],
                    help="local optimizer (the BASELINE configs use SGD; adam = fused Adam K-OPT)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--streams-per-gpu", type=int, default=8,
                    help="client streams of the generic autograd engine (ResNet configs)")
    ap.add_argument("--no-channels-last", action="store_true",
                    help="generic autograd engine: keep NCHW inputs (NHWC is the default)")
    args = ap.parse_args()
    if args.clients_per_gpu is None:
        args.clients_per_gpu = CLIENTS_PER_GPU if args.model == "mnist_cnn" else 8
    return args


# ---------------------------------------------------------------------------
# clocks: sample nvidia-smi DURING the timed region
# ---------------------------------------------------------------------------
class ClockSampler:
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int) -> None:
        self.gpu = gpu_index
        self.proc = None
        self.path = None

    def start(self) -> None:
        try:
            fd, self.path = tempfile.mkstemp(prefix="nfb_clocks_", suffix=".csv")
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                 "-lms", "20", "-i", str(self.gpu)],
                stdout=fd, stderr=subprocess.DEVNULL)
            os.close(fd)
        except Exception:
            self.proc = None

    def stop(self) -> dict:
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        time.sleep(0.05)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        try:
            rows = [r.split(",") for r in Path(self.path).read_text().strip().splitlines()]
            os.unlink(self.path)
        except Exception:
            return out
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            if len(r) < 8:
                continue
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
            except ValueError:
                continue
            for name, cell in zip(names, r[4:8]):
                if cell.strip().lower().startswith("active"):
                    reasons.add(name)
        if sm:
            out.update(sm_mhz=statistics.median(sm), sm_max_mhz=max(mx), reasons=sorted(reasons),
                       samples=len(sm))
        return out


def emit(payload: dict) -> None:
    print(json.dumps(payload), flush=True)


# ---------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------
def run_ours(args: argparse.Namespace) -> None:
    import torch

    from nanofed_b200.ops import native
    from nanofed_b200.parallel.cohort import CohortConfig, FedSimulation
    from nanofed_b200.privacy import PrivacyConfig
    from nanofed_b200.runtime import dist as rdist

    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl ours needs a CUDA device (B200)")
    import logging

    logging.getLogger("nanofed").setLevel(logging.WARNING)
    cfg = CohortConfig(
        model=args.model, clients_per_gpu=args.clients_per_gpu,
        samples_per_client=args.samples_per_client, batch_size=BATCH, local_epochs=LOCAL_EPOCHS,
        learning_rate=LR if args.optimizer == "sgd" else 1e-3, optimizer=args.optimizer,
        compute_dtype=args.dtype, engine=args.engine,
        transport=args.transport, streams_per_gpu=args.streams_per_gpu,
        privacy=PrivacyConfig(epsilon=10.0, noise_multiplier=1.1) if args.dp else None,
        local_privacy=PrivacyConfig(epsilon=10.0, noise_multiplier=1.1) if args.local_dp else None,
    )
    if args.no_channels_last:
        os.environ["NFB_CHANNELS_LAST"] = "0"
    sim = FedSimulation(cfg)
    rank, dev = sim.rank, sim.device
    if rank.world_size != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={rank.world_size}: launch with "
                         "torch.distributed.run --nproc-per-node N")
    warmup = max(args.warmup, 3)
    steps = max(args.steps, 1)

    # ---- multi-GPU correctness, visible to the driver: one aggregation of per-rank random client
    # rows through the fused peer kernel AND through the NCCL baseline (torch partial sum +
    # all_reduce); the largest element-wise difference goes into the JSON line and a difference
    # above 1e-5 fails the run (a wrong shard boundary at G = 8 would show up here)
    agg_parity = None
    if sim.reducer is not None and type(sim.reducer).__name__ == "PeerReducer":
        import copy

        from nanofed_b200.parallel.peer import NcclReducer

        a = sim.arena
        keep = (a.master.clone(), a.master_lowp.clone())
        gen = torch.Generator(device=dev).manual_seed(4242 + rank.rank)
        a.slots.copy_(torch.randn(a.slots.shape, device=dev, generator=gen))
        sim.reducer.reduce(a)
        torch.cuda.synchronize(dev)
        rdist.barrier()
        fused = a.master.clone()
        shadow = copy.copy(a)
        shadow.master, shadow.master_lowp = torch.zeros_like(fused), torch.zeros_like(keep[1])
        NcclReducer(shadow, rank).reduce(shadow)
        torch.cuda.synchronize(dev)
        agg_parity = rdist.all_reduce_max(float((fused - shadow.master).abs().max()), dev)
        sim.reducer.check()
        rdist.barrier()
        a.master.copy_(keep[0])
        a.master_lowp.copy_(keep[1])
        torch.cuda.synchronize(dev)
        rdist.barrier()
        if not agg_parity <= 1e-5:
            raise SystemExit(f"fused peer aggregation differs from the NCCL baseline by "
                             f"{agg_parity:.3e} (> 1e-5)")

    def timed(fn, n: int, finish=None) -> float:
        rdist.barrier()
        torch.cuda.synchronize(dev)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(n):
            fn()
        if finish is not None:
            finish()
        b.record()
        torch.cuda.synchronize(dev)
        rdist.barrier()
        return rdist.all_reduce_max(a.elapsed_time(b), dev)  # ms, max over ranks

    # ---- device-resident arm -------------------------------------------------------
    sim.timer.enabled = False
    for _ in range(warmup):
        sim.enqueue_round()
    torch.cuda.synchronize(dev)
    native.reset_launch_count()
    sim.enqueue_round()
    eager = native.LAUNCHES
    launches_per_round = eager + int(getattr(sim.engine, "launches_per_round", 0)
                                     if getattr(sim.engine, "_graph", None) is not None else 0)
    sampler = ClockSampler(rank.local_rank)
    if rank.is_main:
        sampler.start()
    ms = timed(sim.enqueue_round, steps)
    clocks = sampler.stop() if rank.is_main else {}
    ms_per_step = ms / steps
    samples_per_round = sim.samples_per_round
    value = samples_per_round / (ms_per_step / 1e3)

    # ---- end-to-end arm: pinned-host inputs in, metrics out, every round -------------
    e2e = None
    if not args.no_e2e:
        host_images = sim.shards.images.cpu().pin_memory()
        host_labels = sim.shards.labels.cpu().pin_memory()
        h2d = (host_images.numel() * host_images.element_size()
               + host_labels.numel() * host_labels.element_size())
        d2h = sim._stats_host.numel() * sim._stats_host.element_size()
        last = {}

        # Every round inside the timed region (a) issues the H2D copy of one round's inputs from
        # pinned host memory (it is the NEXT round's data, so PCIe overlaps this round's
        # training), (b) adopts the previously staged inputs, (c) trains + aggregates, and
        # (d) copies the round's metrics to pinned host memory and reads them (event sync).
        # The metrics of round r are read after round r+1 has been submitted (two pinned metric
        # buffers), so the GPU does not idle while the host turns counters into Python floats;
        # the last round is collected inside the timed region too: K submits, K D2H reads.
        pending = {"h": None}

        def e2e_round():
            h = sim.submit_round(use_staged=True, prefetch_next=(host_images, host_labels))
            if pending["h"] is not None:
                last["o"] = sim.collect(pending["h"])
            pending["h"] = h

        def drain():
            if pending["h"] is not None:
                last["o"] = sim.collect(pending["h"])
                pending["h"] = None

        sim.prefetch_inputs(host_images, host_labels)
        for _ in range(warmup):
            e2e_round()
        drain()
        e2e_ms = timed(e2e_round, steps, finish=drain) / steps
        e2e = {"value": samples_per_round / (e2e_ms / 1e3), "unit": "client_samples/s",
               "ms_per_step": e2e_ms, "rounds_per_sec": 1e3 / e2e_ms,
               "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
               "last_round_loss": last["o"].agg_metrics.get("loss")}

    if sim.reducer is not None and hasattr(sim.reducer, "check"):
        sim.reducer.check()
    phase = {}
    sim.timer.enabled = True
    sim.timer.reset()
    sim.enqueue_round()
    phase = {k: round(v, 4) for k, v in sim.timer.summary_ms().items()}

    if rank.is_main:
        arena_mb = sim.arena.nbytes() / 2**20
        emit({
            "metric": "client_samples_per_sec", "value": value, "unit": "client_samples/s",
            "n_gpus": args.gpus, "steps": steps, "warmup": warmup, "ms_per_step": ms_per_step,
            "rounds_per_sec": 1e3 / ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None,   # BASELINE.json publishes no number for this config
            "dtype": args.dtype, "data": "synthetic", "impl": "ours",
            "config": {
                "model": f"{args.model} ({sim.layout.num_values} state elements, random init)",
                "global_batch": args.clients_per_gpu * args.gpus * BATCH,
                "seq_len": None, "clients": sim.num_clients,
                "clients_per_gpu": args.clients_per_gpu,
                "samples_per_client": args.samples_per_client, "local_epochs": LOCAL_EPOCHS,
                "local_batch": BATCH, "local_steps_per_round": cfg.local_steps,
                "optimizer": f"SGD(lr={LR})" if args.optimizer == "sgd" else "Adam(lr=0.001)",
                "aggregation": "FedAvg (sample-weighted)"
                + (" + central DP clip/noise" if args.dp else "")
                + (" + client-side DP-SGD" if args.local_dp else ""),
                "parallelism": f"clients-dp{args.gpus}", "engine": sim.engine_name,
                "agg_path": getattr(sim.reducer, "path", "single_gpu_fused_kernel"),
                "l2_policy": f"working set per round {arena_mb:.0f} MB arena + activations "
                             "> 126 MB L2 (inputs larger than L2)",
                "step_definition": "one federated round (bcast + local epoch + aggregate)",
            },
            "agg_parity_max_abs_diff": agg_parity,
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches_per_round * steps,
            "gpu_launches_per_round": launches_per_round, "phase_ms": phase,
        })
    rdist.shutdown()


def main() -> None:
    args = parse()
    if args.impl == "reference":
        from baseline.reference_arm import run_reference

        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()

"""Host-side multi-process logic on gloo (world_size 2): NCCL-baseline reducer semantics."""

import os
import socket
import sys
from pathlib import Path

import torch
import torch.multiprocessing as mp

ROOT = str(Path(__file__).resolve().parent.parent)


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, world: int, port: int, out_dir: str) -> None:
    sys.path.insert(0, ROOT)
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world),
                      MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import logging

    logging.getLogger("nanofed").setLevel(logging.ERROR)
    from nanofed_b200.parallel.cohort import CohortConfig, FedSimulation

    cfg = CohortConfig(model="mnist_cnn", clients_per_gpu=2, samples_per_client=16, batch_size=8,
                       compute_dtype="fp32", engine="autograd", device="cpu", transport="nccl",
                       sample_counts=[16, 8] if rank == 0 else [4, 16])
    sim = FedSimulation(cfg)
    # make local training deterministic: no dropout
    for w in sim.engine.workers:
        w.dropout1.p = 0.0
        w.dropout2.p = 0.0
    out = sim.run_round()
    torch.save({"master": sim.arena.master.clone(), "slots": sim.arena.slots.clone(),
                "counts": sim.arena.counts.clone(), "stats": sim.arena.stats.clone(),
                "clients": [(c["client_id"], c["weight"], c["metrics"]["loss"])
                            for c in out.client_metrics],
                "num_clients": out.num_clients}, f"{out_dir}/r{rank}.pt")
    from nanofed_b200.runtime import dist as rdist

    rdist.shutdown()


def test_two_rank_round_equals_global_fedavg(tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    a = torch.load(tmp_path / "r0.pt")
    b = torch.load(tmp_path / "r1.pt")
    assert torch.equal(a["master"], b["master"])  # every rank holds the same global model
    slots = torch.cat([a["slots"], b["slots"]])
    counts = torch.cat([a["counts"], b["counts"]])
    want = (counts / counts.sum()).view(-1, 1).mul(slots).sum(0)
    torch.testing.assert_close(a["master"], want, rtol=1e-5, atol=1e-6)
    # the round record is job-wide on every rank: all four clients with their FedAvg weights and
    # the losses measured on the rank that trained them
    assert a["num_clients"] == b["num_clients"] == 4
    assert [c[0] for c in a["clients"]] == [f"client_{i}" for i in range(4)] == [c[0] for c in b["clients"]]
    for got, w in zip(a["clients"], (counts / counts.sum()).tolist()):
        assert abs(got[1] - w) < 1e-6
    stats = torch.cat([a["stats"], b["stats"]])
    for got, row in zip(b["clients"], stats.tolist()):
        assert abs(got[2] - row[0] / max(row[2], 1.0)) < 1e-5


def _straggler_worker(rank: int, world: int, port: int, out_dir: str) -> None:
    sys.path.insert(0, ROOT)
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world),
                      MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import logging

    logging.getLogger("nanofed").setLevel(logging.ERROR)
    from nanofed_b200.parallel.cohort import CohortConfig, FedSimulation

    cfg = CohortConfig(model="mnist_cnn", clients_per_gpu=2, samples_per_client=16, batch_size=8,
                       compute_dtype="fp32", engine="autograd", device="cpu", transport="nccl",
                       round_deadline_ms=6000.0)   # generous: a busy CI box must not make rank 0 late too
    sim = FedSimulation(cfg)
    if rank == 1:
        sim.inject_fault("delay", 0, delay_ms=10000.0)       # rank 1's cohort misses the deadline
    sim.run_round()
    torch.save({"master": sim.arena.master.clone(), "slots": sim.arena.slots.clone(),
                "counts": sim.arena.counts.clone(), "late": sim.stragglers_last_round,
                "mask": sim.arena.mask.clone()}, f"{out_dir}/s{rank}.pt")
    from nanofed_b200.runtime import dist as rdist

    rdist.shutdown()


def test_cohort_that_misses_the_deadline_is_left_out_of_the_round(tmp_path):
    port = _free_port()
    mp.spawn(_straggler_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    a = torch.load(tmp_path / "s0.pt")
    b = torch.load(tmp_path / "s1.pt")
    assert a["late"] == [] and b["late"] == ["client_2", "client_3"]
    assert torch.equal(a["master"], b["master"])
    # the global model is the average of rank 0's clients only
    want = (a["counts"] / a["counts"].sum()).view(-1, 1).mul(a["slots"]).sum(0)
    torch.testing.assert_close(a["master"], want, rtol=1e-5, atol=1e-6)
    assert bool((b["mask"] == 1).all())          # the participation mask itself is restored

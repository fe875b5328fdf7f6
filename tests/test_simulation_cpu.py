"""FedSimulation on CPU (autograd engine, oracle ops) incl. Coordinator integration and DP."""

import asyncio
import json

import torch

from nanofed_b200 import Coordinator, CoordinatorConfig, FedAvgAggregator, ModelManager
from nanofed_b200.models import MNISTModel
from nanofed_b200.parallel.cohort import CohortConfig, FedSimulation
from nanofed_b200.privacy import PrivacyConfig


def _cfg(**kw):
    base = dict(model="mnist_cnn", clients_per_gpu=3, samples_per_client=16, batch_size=8,
                compute_dtype="fp32", engine="autograd", device="cpu", sample_counts=[16, 8, 4])
    base.update(kw)
    return CohortConfig(**base)


def test_round_equals_weighted_mean_of_client_models():
    sim = FedSimulation(_cfg())
    out = sim.run_round()
    counts = torch.tensor([16.0, 8, 4])
    want = (counts / counts.sum()).view(-1, 1).mul(sim.arena.slots).sum(0)
    torch.testing.assert_close(sim.arena.master, want, rtol=1e-5, atol=1e-6)
    assert out.num_clients == 3 and abs(sum(c["weight"] for c in out.client_metrics) - 1) < 1e-9
    assert set(out.agg_metrics) == {"loss", "accuracy", "samples_processed"}
    state = sim.global_state()
    MNISTModel().load_state_dict(state)                      # reference-shaped state_dict


def test_dp_round_clips_and_accounts():
    sim = FedSimulation(_cfg(privacy=PrivacyConfig(epsilon=10.0, noise_multiplier=0.5,
                                                   max_gradient_norm=1.0)))
    out = sim.run_round()
    assert out.agg_metrics["privacy_epsilon"] > 0
    # every client row has norm >> 1, so the aggregate of clipped rows has norm <= ~1 (+ noise)
    assert float(sim.arena.sumsq.sqrt().min()) > 1.0
    noise_free = float(sim.arena.master.norm())
    assert noise_free < 1.0 + 0.5 * 1.0 / 3 * (sim.layout.total ** 0.5) * 1.2


def test_coordinator_drives_simulation_and_writes_reference_layout(tmp_path):
    sim = FedSimulation(_cfg())
    model = MNISTModel()
    coord = Coordinator(ModelManager(model), FedAvgAggregator(), None,
                        CoordinatorConfig(num_rounds=2, min_clients=3, min_completion_rate=1.0,
                                          round_timeout=30, base_dir=tmp_path), engine=sim)

    async def run():
        return [m async for m in coord.start_training()]

    metrics = asyncio.run(run())
    assert [m.round_id for m in metrics] == [0, 1] and metrics[0].num_clients == 3
    assert len(list((tmp_path / "models" / "models").glob("*.pt"))) == 3
    rec = json.loads((tmp_path / "metrics" / "metrics_round_1.json").read_text())
    assert rec["status"] == "COMPLETED" and len(rec["client_metrics"]) == 3
    torch.testing.assert_close(model.fc1.weight,
                               sim.layout.views(sim.arena.master)["fc1.weight"])


def test_participation_mask_and_nonfinite_screening():
    sim = FedSimulation(_cfg(screen_nonfinite=True, sample_counts=[8, 8, 8]))
    sim.inject_fault("drop", 0)
    sim.inject_fault("nan", 1)          # client 1 diverges this round
    sim.run_round()
    # only client 2 is left: the global model equals its weights, and it is finite
    torch.testing.assert_close(sim.arena.master, sim.arena.slots[2])
    assert torch.isfinite(sim.arena.master).all()
    assert sim.arena.nonfinite.tolist() == [0, 1, 0]
    sim.set_participation([True, True, True])
    sim.run_round()
    want = sim.arena.slots.mean(0)
    torch.testing.assert_close(sim.arena.master, want, rtol=1e-5, atol=1e-6)


def test_resume_from_checkpoint(tmp_path):
    sim = FedSimulation(_cfg())
    sim.run_round()
    state = sim.global_state()
    torch.save(state, tmp_path / "g.pt")
    again = FedSimulation(_cfg(seed=3))
    again.load_global_state(torch.load(tmp_path / "g.pt", weights_only=True))
    torch.testing.assert_close(again.arena.master, sim.arena.master)


def test_submit_collect_pipeline_matches_run_round():
    import pytest

    seq = FedSimulation(_cfg(seed=3))
    want = [seq.run_round().agg_metrics["loss"] for _ in range(3)]
    sim = FedSimulation(_cfg(seed=3))
    h0 = sim.submit_round()
    h1 = sim.submit_round()                       # round 1 enqueued before round 0 is read
    with pytest.raises(RuntimeError):
        sim.submit_round()                        # only two metric buffers
    got = [sim.collect(h0).agg_metrics["loss"]]
    h2 = sim.submit_round()
    got += [sim.collect(h1).agg_metrics["loss"], sim.collect(h2).agg_metrics["loss"]]
    assert got == pytest.approx(want, rel=1e-6)
    with pytest.raises(RuntimeError):
        sim.collect(h2)                           # already collected
    torch.testing.assert_close(sim.arena.master, seq.arena.master)


def test_single_cohort_missing_its_deadline_is_kept():
    sim = FedSimulation(_cfg(round_deadline_ms=1e-6))      # impossible deadline, one cohort
    before = sim.arena.master.clone()
    sim.run_round()
    assert sim.last_train_ms > 1e-6 and sim.stragglers_last_round == []
    assert torch.isfinite(sim.arena.master).all() and not torch.equal(before, sim.arena.master)


def test_batchnorm_counters_travel_through_the_arena_in_one_gather_scatter():
    cfg = CohortConfig(model="resnet18", clients_per_gpu=2, samples_per_client=4, batch_size=2,
                       compute_dtype="fp32", engine="autograd", device="cpu")
    sim = FedSimulation(cfg)
    eng = sim.engine
    assert eng._int_fast and eng._int_idx.numel() == 20          # ResNet-18: 20 BN counters
    sim.run_round()
    state = sim.global_state()
    counters = [v for k, v in state.items() if k.endswith("num_batches_tracked")]
    assert len(counters) == 20 and all(v.dtype == torch.int64 for v in counters)
    assert all(int(v) == cfg.local_steps for v in counters)      # mean of identical counters
    sim.run_round()                                              # counters restart from the global value
    assert all(int(v) == 2 * cfg.local_steps
               for k, v in sim.global_state().items() if k.endswith("num_batches_tracked"))
    assert all(torch.isfinite(v).all() for v in sim.global_state().values() if v.is_floating_point())

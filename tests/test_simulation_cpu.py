"""FedSimulation on CPU (autograd engine, oracle ops) incl. Coordinator integration and DP."""

import asyncio
import json

import pytest
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


def test_all_clients_masked_keeps_the_global_model():
    sim = FedSimulation(_cfg())
    sim.run_round()
    before = sim.arena.master.clone()
    sim.set_participation([False, False, False])
    sim.run_round()
    assert torch.equal(sim.arena.master, before)          # not overwritten with zeros
    sim.set_participation([True, True, True])
    sim.run_round()
    assert not torch.equal(sim.arena.master, before)


def test_collect_raises_when_the_aggregation_kernel_reported_a_timeout():
    """The fused multi-GPU kernel bounds its waits and sets a status word; that word travels with
    the round's metrics and ``collect`` refuses to hand out a half-written global model."""

    class _TimedOutReducer:
        path = "fake"

        def __init__(self, sim):
            self.sim = sim
            self.status = torch.zeros(1, dtype=torch.int32)
            self.stats_all = torch.zeros(1, sim.cfg.clients_per_gpu, 3)

        def all_stats(self):
            return self.stats_all

        def reduce(self, arena, **kw):
            self.stats_all[0].copy_(arena.stats)
            self.status[0] = 1                            # "a peer never arrived"

    sim = FedSimulation(_cfg())
    sim.reducer = _TimedOutReducer(sim)
    handle = sim.submit_round()
    with pytest.raises(RuntimeError, match="timed out"):
        sim.collect(handle)
    assert int(sim.reducer.status[0]) == 0                # reported once, then reset


def test_dp_noise_is_calibrated_for_the_clients_that_take_part():
    from nanofed_b200.privacy import PrivacyConfig

    priv = PrivacyConfig(epsilon=10.0, noise_multiplier=1.0, max_gradient_norm=1.0)
    a = FedSimulation(_cfg(privacy=priv))
    b = FedSimulation(_cfg(privacy=priv))
    assert a._dp_seed != b._dp_seed                       # drawn from os.urandom, not from the config
    pinned = FedSimulation(_cfg(privacy=priv, dp_seed=5))
    assert pinned._dp_seed == 5
    pinned.drop_clients([0])
    out = pinned.run_round()
    assert pinned._k_active_round == 2                    # σ·C / 2, not / 3
    assert len(pinned.accountant._noise_events) == 2 if hasattr(pinned.accountant, "_noise_events") else True
    assert "privacy_epsilon" in out.agg_metrics


def test_weight_decay_leaves_batchnorm_buffers_alone():
    """AutogradCohort runs the fused optimizer over whole arena rows; weight decay must reach the
    parameters only - torch.optim.SGD never touches running statistics."""
    import torch.nn.functional as F

    from nanofed_b200.models import ResNet18

    cfg = CohortConfig(model="resnet18", clients_per_gpu=1, samples_per_client=4, batch_size=4,
                       learning_rate=0.05, weight_decay=0.1, device="cpu", engine="autograd",
                       compute_dtype="fp32")
    sim = FedSimulation(cfg)
    state0 = sim.global_state()
    from nanofed_b200 import ops

    ops.bcast_global(sim.arena.master, sim.arena.slots, sim.arena.slots_lowp, sim.arena.momentum)
    sim.engine.after_broadcast()
    sim.engine.local_steps()
    got = sim.layout.unflatten(sim.arena.slots[0].clone())
    model = ResNet18().train()
    model.load_state_dict(state0)
    opt = torch.optim.SGD(model.parameters(), lr=0.05, weight_decay=0.1)
    x, y = sim.shards.images[0, :4], sim.shards.labels[0, :4]
    F.cross_entropy(model(x), y).backward()
    opt.step()
    want = model.state_dict()
    for key in ("bn1.running_var", "layer1.0.bn1.running_mean", "bn1.weight", "conv1.weight"):
        torch.testing.assert_close(got[key].float(), want[key].float(), rtol=2e-4, atol=2e-5)


def test_recovery_restores_the_engine_arena(tmp_path):
    sim = FedSimulation(_cfg())
    model = MNISTModel()
    coord = Coordinator(ModelManager(model), FedAvgAggregator(), None,
                        CoordinatorConfig(num_rounds=1, min_clients=3, min_completion_rate=1.0,
                                          round_timeout=30, base_dir=tmp_path), engine=sim)
    good = {k: v.clone() for k, v in model.state_dict().items()}

    class _FT:
        def handle_failure(self, exc, round_id):
            import types

            return types.SimpleNamespace(round_id=0), good

    coord._ft = _FT()
    sim.arena.master.add_(1.0)                            # the arena has moved away
    assert asyncio.run(coord._recover(RuntimeError("boom")))
    torch.testing.assert_close(sim.layout.views(sim.arena.master)["fc1.weight"],
                               good["fc1.weight"])


def test_client_side_dp_sgd_clips_each_clients_batch_gradient_and_adds_masked_noise():
    """local_privacy = the reference PrivateTrainer recipe (clip_grad_norm_ + per-parameter
    Gaussian noise, nanofed/trainer/private.py:54-86) for every client of the cohort in one K-OPT
    call: compare one local step with an explicit per-client computation."""
    import torch.nn.functional as F

    from nanofed_b200 import ops
    from nanofed_b200.ops import philox

    priv = PrivacyConfig(epsilon=10.0, noise_multiplier=0.3, max_gradient_norm=0.5)
    cfg = _cfg(local_privacy=priv, dp_seed=21, samples_per_client=8, sample_counts=[8, 8, 8],
               learning_rate=0.1)
    sim = FedSimulation(cfg)
    for w in sim.engine.workers:
        w.dropout1.p = w.dropout2.p = 0.0
    a = sim.arena
    ops.bcast_global(a.master, a.slots, a.slots_lowp, a.momentum)
    sim.engine.after_broadcast()
    before = a.slots.clone()
    sim.engine.local_steps()
    n, c = sim.layout.total, 3
    z = philox.normal(c * n, 21, ((c * n + 3) // 4) * 4).view(c, n)      # step counter = 1
    pmask = sim.engine._dp_noise_mask.float()
    for k in range(c):
        model = MNISTModel()
        model.load_state_dict(sim.layout.unflatten(before[k]))
        for m in model.modules():
            if isinstance(m, torch.nn.Dropout):
                m.p = 0.0
        model.train()
        x, y = sim.shards.images[k, :8], sim.shards.labels[k, :8]
        F.cross_entropy(model(x), y).backward()
        g = sim.layout.flatten({nme: p.grad for nme, p in model.named_parameters()} |
                               {nme: torch.zeros_like(b) for nme, b in model.named_buffers()})
        clip = min(1.0, 0.5 / (float(g.norm()) + 1e-6))
        want = before[k] - 0.1 * (g * clip + 0.3 * 0.5 * z[k] * pmask)
        torch.testing.assert_close(a.slots[k], want, rtol=1e-4, atol=1e-5)
    out = sim.run_round()
    assert out.agg_metrics["local_privacy_epsilon"] > 0


def test_adam_cohort_matches_torch_adam():
    """``optimizer="adam"``: the fused Adam kernel of K-OPT over the arena rows (moments reset every
    round) against ``torch.optim.Adam`` with the same betas / eps / L2 weight decay, two local steps
    of every client; BatchNorm buffers must come out exactly as torch leaves them.  (eps = 1e-3:
    with the default 1e-8 the first Adam step is lr * sign(g), and entries whose gradient is
    rounding noise would flip between two fp32 summation orders.)"""
    import torch.nn.functional as F

    from nanofed_b200 import ops
    from nanofed_b200.models import ResNet18

    cfg = CohortConfig(model="resnet18", clients_per_gpu=2, samples_per_client=8, batch_size=4,
                       learning_rate=2e-3, weight_decay=0.05, optimizer="adam",
                       adam_betas=(0.8, 0.95), adam_eps=1e-3, device="cpu", engine="autograd",
                       compute_dtype="fp32")
    sim = FedSimulation(cfg)
    state0 = sim.global_state()
    ops.bcast_global(sim.arena.master, sim.arena.slots, sim.arena.slots_lowp, sim.arena.momentum)
    sim.engine.after_broadcast()
    sim.engine.local_steps()
    sim.engine.before_aggregate()
    for k in range(2):
        got = sim.layout.unflatten(sim.arena.slots[k].clone())
        model = ResNet18().train()
        model.load_state_dict(state0)
        opt = torch.optim.Adam(model.parameters(), lr=2e-3, betas=(0.8, 0.95), eps=1e-3,
                               weight_decay=0.05)
        for s in range(2):
            x, y = sim.shards.images[k, 4 * s : 4 * s + 4], sim.shards.labels[k, 4 * s : 4 * s + 4]
            opt.zero_grad()
            F.cross_entropy(model(x), y).backward()
            opt.step()
        want = model.state_dict()
        for key in ("conv1.weight", "layer2.0.shortcut.0.weight", "bn1.weight", "fc.bias",
                    "bn1.running_var", "layer1.0.bn1.running_mean"):
            torch.testing.assert_close(got[key].float(), want[key].float(), rtol=2e-3, atol=2e-5)
        assert int(got["bn1.num_batches_tracked"]) == 2


def test_adam_rejects_combinations_it_does_not_implement():
    import pytest

    with pytest.raises(ValueError, match="local_privacy"):
        FedSimulation(_cfg(optimizer="adam", local_privacy=PrivacyConfig(epsilon=8.0)))
    with pytest.raises(ValueError, match="optimizer"):
        FedSimulation(_cfg(optimizer="lion"))

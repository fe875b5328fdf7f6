"""Flagship grouped engine vs PyTorch autograd (fp32) on the same weights and batch."""

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _setup(c=3, b=16, steps=2, dropout=False, graph=False):
    from nanofed_b200 import ops
    from nanofed_b200.data.synthetic import make_client_shards
    from nanofed_b200.models import MNISTModel
    from nanofed_b200.models.mnist_cohort import MnistCohortNet
    from nanofed_b200.parallel.cohort import CohortConfig
    from nanofed_b200.runtime.arena import Arena
    from nanofed_b200.runtime.flat import FlatLayout

    torch.manual_seed(0)
    dev = torch.device("cuda")
    cfg = CohortConfig(clients_per_gpu=c, samples_per_client=b * steps, batch_size=b,
                       learning_rate=0.1, use_cuda_graph=graph)
    model = MNISTModel()
    layout = FlatLayout.from_module(model)
    arena = Arena(layout, c, dev, with_lowp=True, with_grads=False)
    arena.load_global(model.state_dict())
    shards = make_client_shards(c, b * steps, seed=3).to(dev)
    eng = MnistCohortNet(cfg, arena, shards, dropout=dropout)
    ops.bcast_global(arena.master, arena.slots, arena.slots_lowp, None)
    return cfg, model, layout, arena, shards, eng


def test_one_step_matches_autograd():
    from nanofed_b200.ops import tc

    cfg, model, layout, arena, shards, eng = _setup()
    eng.fuse_fc1_sgd = False               # keep the fc1 weight gradient observable
    c, b = cfg.clients_per_gpu, cfg.batch_size
    eng._step_group(0, 0)                  # forward + backward: gradients only
    torch.cuda.synchronize()
    tc.check_status(arena.device)
    grads_before_opt = eng.grads.clone()
    eng._apply_optimizer(0)                # fused SGD (and it zeroes the atomic gradient prefix)
    torch.cuda.synchronize()
    assert float(eng.grads[:, : eng._atomic_prefix].abs().sum()) == 0.0
    assert float(eng.hpre.abs().sum()) == 0.0      # split-K accumulator handed back clean
    ref = model.cuda().eval()
    for k in range(c):
        x, y = shards.images[k, :b], shards.labels[k, :b]
        # how far does PyTorch's own bf16 path (autocast, cuDNN/cuBLAS) sit from fp32?
        ref.zero_grad()
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out16 = ref(x)
        F.cross_entropy(out16.float(), y).backward()
        g16 = {n: p.grad.clone() for n, p in ref.named_parameters()}
        ref.zero_grad()
        out = ref(x)
        loss = F.cross_entropy(out, y)
        loss.backward()
        got = layout.views(grads_before_opt[k])
        for name, p in ref.named_parameters():
            g, r = got[name].float(), p.grad
            rel = (g - r).norm() / (r.norm() + 1e-12)
            rel16 = (g16[name] - r).norm() / (r.norm() + 1e-12)
            cos = F.cosine_similarity(g.flatten(), r.flatten(), dim=0)
            # bf16 activations: we must be at least as close to fp32 as torch's bf16 path
            assert rel < max(1.25 * rel16, 0.02), f"client {k} {name}: {rel:.4f} vs {rel16:.4f}"
            assert cos > 0.99, f"client {k} {name}: cosine {cos:.5f}"
        # loss / accuracy counters
        loss_sum, hits, seen = arena.stats[k].tolist()
        assert seen == b
        assert abs(loss_sum / b - loss.item()) < 0.02
        assert abs(hits - (out.argmax(1) == y).sum().item()) <= 1
        # SGD applied on the fp32 master, bf16 copy refreshed
        new = layout.views(arena.slots[k])
        for name, p in ref.named_parameters():
            want = p.data - cfg.learning_rate * got[name].float()
            torch.testing.assert_close(new[name], want, rtol=1e-5, atol=1e-6)
    assert torch.equal(arena.slots_lowp, arena.slots.bfloat16())


def test_fc1_update_fused_into_the_wgrad_epilogue_matches_separate_optimizer():
    from nanofed_b200.ops import tc

    results = []
    for fused in (True, False):
        cfg, model, layout, arena, shards, eng = _setup(c=3, b=16, steps=2)
        assert eng.fuse_fc1_sgd, "plain SGD must take the fused path by default"
        eng.fuse_fc1_sgd = fused
        for s in range(2):
            eng.step(s)
        torch.cuda.synchronize()
        tc.check_status(arena.device)
        assert torch.equal(arena.slots_lowp, arena.slots.bfloat16())
        if fused:   # the gradient of fc1.weight is never written
            lo, hi = eng._fc1_range
            assert float(eng.grads[:, lo:hi].abs().sum()) == 0.0
        results.append(arena.slots.clone())
    # same fp32 accumulators, same update formula; the atomically accumulated conv gradients
    # differ in summation order between runs
    diff = (results[0] - results[1]).norm() / results[1].norm()
    assert float(diff) < 1e-4, float(diff)
    torch.testing.assert_close(results[0], results[1], rtol=1e-3, atol=5e-5)
    moved = (results[0] - arena.master[None]).abs().max().item()
    assert moved > 1e-4                    # and the step did move the weights


def test_graph_replay_trains_and_dropout_masks_advance():
    cfg, model, layout, arena, shards, eng = _setup(c=2, b=16, steps=4, dropout=True, graph=True)
    losses = []
    for _ in range(6):
        arena.stats.zero_()
        eng.local_steps()
        torch.cuda.synchronize()
        s = arena.stats.sum(0).tolist()
        losses.append(s[0] / s[2])
    assert all(torch.isfinite(torch.tensor(losses)))
    assert losses[-1] < losses[0], losses          # it learns the synthetic labels
    assert int(eng.rng_state[0].item()) == 6 * cfg.local_steps  # one dropout stream step per local step


def test_full_round_through_fedsimulation():
    from nanofed_b200.parallel.cohort import CohortConfig, FedSimulation

    sim = FedSimulation(CohortConfig(clients_per_gpu=4, samples_per_client=64, batch_size=32,
                                     engine="grouped"))
    assert sim.engine_name == "grouped"
    first = sim.run_round().agg_metrics["loss"]
    for _ in range(4):
        last = sim.run_round().agg_metrics["loss"]
    assert last < first
    state = sim.global_state()
    assert set(state) == set(sim.template.state_dict())
    assert all(torch.isfinite(v).all() for v in state.values())


def test_prefetched_inputs_match_direct_staging():
    from nanofed_b200.parallel.cohort import CohortConfig, FedSimulation

    def run(pipelined: bool):
        sim = FedSimulation(CohortConfig(clients_per_gpu=2, samples_per_client=64, batch_size=32,
                                         engine="grouped", seed=5))
        gen = torch.Generator().manual_seed(9)
        rounds = [(torch.randn(2, 64, 1, 28, 28, generator=gen).pin_memory(),
                   torch.randint(0, 10, (2, 64), generator=gen).pin_memory()) for _ in range(3)]
        if pipelined:
            sim.prefetch_inputs(*rounds[0])
            for r in range(3):
                nxt = rounds[r + 1] if r + 1 < 3 else None
                sim.run_round(use_staged=True, prefetch_next=nxt)
        else:
            for r in range(3):
                sim.run_round(host_inputs=rounds[r])
        return sim.arena.master.clone()

    a, b = run(False), run(True)
    diff = (a - b).norm() / a.norm()            # atomics reorder sums between runs
    assert float(diff) < 2e-3, float(diff)


def test_rounds_submitted_ahead_of_their_metrics_match_sequential_rounds():
    from nanofed_b200.parallel.cohort import CohortConfig, FedSimulation

    def make():
        return FedSimulation(CohortConfig(clients_per_gpu=2, samples_per_client=64, batch_size=32,
                                          engine="grouped", seed=5))

    gen = torch.Generator().manual_seed(9)
    rounds = [(torch.randn(2, 64, 1, 28, 28, generator=gen).pin_memory(),
               torch.randint(0, 10, (2, 64), generator=gen).pin_memory()) for _ in range(4)]
    seq = make()
    want = [seq.run_round(host_inputs=r).agg_metrics["loss"] for r in rounds]
    sim = make()
    sim.prefetch_inputs(*rounds[0])
    got, pending = [], None
    for r in range(4):
        nxt = rounds[r + 1] if r + 1 < 4 else None
        h = sim.submit_round(use_staged=True, prefetch_next=nxt)   # round r runs ...
        if pending is not None:
            got.append(sim.collect(pending).agg_metrics["loss"])   # ... while r-1 is read
        pending = h
    got.append(sim.collect(pending).agg_metrics["loss"])
    # same data, same seeds; the two runs differ only in the summation order of the atomically
    # accumulated gradients (and in bf16 roundings that tip over because of it)
    assert got == pytest.approx(want, rel=1e-2, abs=1e-2)
    diff = (sim.arena.master - seq.arena.master).norm() / seq.arena.master.norm()
    assert float(diff) < 2e-3, float(diff)


def test_deadline_times_local_training_with_cuda_events():
    from nanofed_b200.parallel.cohort import CohortConfig, FedSimulation

    sim = FedSimulation(CohortConfig(clients_per_gpu=2, samples_per_client=64, batch_size=32,
                                     engine="grouped", round_deadline_ms=1e9))
    sim.run_round()                                 # includes graph capture: not representative
    sim.run_round()
    fast = sim.last_train_ms
    sim.inject_fault("delay", 0, delay_ms=60.0)
    sim.run_round()
    assert sim.last_train_ms > fast + 25.0          # the injected delay is seen by the event timer
    assert sim.stragglers_last_round == []          # deadline not exceeded
    sim.cfg.round_deadline_ms = 1e-3                # impossible deadline, single cohort: kept
    before = sim.arena.master.clone()
    sim.run_round()
    assert not torch.equal(before, sim.arena.master) and torch.isfinite(sim.arena.master).all()


def _dp_engine(fused: bool):
    from nanofed_b200 import ops
    from nanofed_b200.data.synthetic import make_client_shards
    from nanofed_b200.models import MNISTModel
    from nanofed_b200.models.mnist_cohort import MnistCohortNet
    from nanofed_b200.parallel.cohort import CohortConfig
    from nanofed_b200.privacy import PrivacyConfig
    from nanofed_b200.runtime.arena import Arena
    from nanofed_b200.runtime.flat import FlatLayout

    torch.manual_seed(0)
    dev = torch.device("cuda")
    c, b = 3, 16
    priv = PrivacyConfig(epsilon=10.0, noise_multiplier=0.2, max_gradient_norm=0.05)
    cfg = CohortConfig(clients_per_gpu=c, samples_per_client=b, batch_size=b, learning_rate=0.1,
                       use_cuda_graph=False, local_privacy=priv, dp_seed=77)
    model = MNISTModel()
    layout = FlatLayout.from_module(model)
    arena = Arena(layout, c, dev, with_lowp=True, with_grads=False)
    arena.load_global(model.state_dict())
    shards = make_client_shards(c, b, seed=3).to(dev)
    import os

    os.environ["NFB_FUSED_DP"] = "1" if fused else "0"
    try:
        eng = MnistCohortNet(cfg, arena, shards, dropout=False)
    finally:
        os.environ.pop("NFB_FUSED_DP", None)
    assert eng.local_dp is not None and not eng.fuse_fc1_sgd and eng.fuse_fc1_dp == fused
    ops.bcast_global(arena.master, arena.slots, arena.slots_lowp, None)
    return eng, arena, layout


def test_dp_sgd_mode_of_the_grouped_engine_matches_the_oracle_update():
    """`local_privacy`: per-client clip factor from the client's complete gradient norm, clip +
    Philox noise inside the optimizer kernels.  (a) materialised path: new = old - lr * (g * min(1,
    C / |g|) + sigma * C * z) with z from the same Philox stream; (b) fused path - the fc1 weight
    gradient never written, its norm from the ghost norm <dh dh^T, p p^T>, clip + noise in the
    wgrad epilogue - gives the same parameters."""
    from nanofed_b200.ops import philox, tc

    eng, arena, layout = _dp_engine(fused=False)
    dev = arena.device
    c = arena.num_clients
    before = arena.slots.clone()
    eng._step_group(0, 0)
    torch.cuda.synchronize()
    tc.check_status(dev)
    g = eng.grads.clone()
    step_ctr = int(eng.rng_state[0])
    eng._apply_optimizer(0)
    torch.cuda.synchronize()
    n = layout.total
    shift = step_ctr * ((c * n + 3) // 4) * 4
    z = philox.normal(c * n, 77, shift).view(c, n).to(dev)
    mask = arena.noise_mask.float()
    norms = g.norm(dim=1)
    assert float(norms.min()) > 0.05                       # the clip is active for every client
    clip = torch.clamp(0.05 / (norms + 1e-6), max=1.0).view(-1, 1)
    want = before - 0.1 * (g * clip + 0.2 * 0.05 * z * mask)
    torch.testing.assert_close(arena.slots, want, rtol=1e-4, atol=1e-6)
    assert torch.equal(arena.slots_lowp, arena.slots.bfloat16())
    # (b) the fused path on the same state
    eng2, arena2, _ = _dp_engine(fused=True)
    eng2.step(0)
    torch.cuda.synchronize()
    tc.check_status(dev)
    assert float(eng2.gram.abs().sum()) == 0.0             # Gram scratch handed back zeroed
    lo, hi = eng2._fc1_range
    assert float(eng2.grads[:, lo:hi].abs().sum()) == 0.0  # the fc1 gradient was never written
    # same noise stream, clip factors equal up to the summation order of the norm
    assert float((arena2.slots - arena.slots).abs().max()) < 2e-6
    assert torch.equal(arena2.slots_lowp, arena2.slots.bfloat16())



def test_central_dp_round_takes_the_update_norms_from_the_last_local_step(monkeypatch):
    """K-STATS folded into the last K-OPT step and the fused fc1 update: ``arena.sumsq`` after the
    local epoch equals the sum of squares of the trained client rows, no ``update_stats`` launch is
    part of the round, and the aggregated model equals the one of a round that runs K-STATS."""
    from nanofed_b200 import ops
    from nanofed_b200.parallel.cohort import CohortConfig, FedSimulation
    from nanofed_b200.privacy import PrivacyConfig

    def sim_of():
        cfg = CohortConfig(model="mnist_cnn", clients_per_gpu=4, samples_per_client=128, batch_size=64,
                           learning_rate=0.05, device="cuda", dp_seed=5, fuse_bcast=False,
                           privacy=PrivacyConfig(epsilon=8.0, max_gradient_norm=10.0,
                                                 noise_multiplier=0.5))
        return FedSimulation(cfg)

    calls = []
    real = ops.update_stats
    monkeypatch.setattr(ops, "update_stats", lambda x: (calls.append(1), real(x))[1])
    sim = sim_of()
    assert sim.engine.emits_sumsq
    sim.engine.p1 = sim.engine.p2 = 0.0          # same dropout-free training in both runs
    sim.run_round()
    assert not calls                              # no separate pass over the client rows
    rows = sim.arena.slots.double()
    torch.testing.assert_close(sim.arena.sumsq.double(), (rows * rows).sum(1), rtol=1e-5, atol=0)
    monkeypatch.setenv("NFB_NO_SUMSQ_SIDE_OUTPUT", "1")
    ref_sim = sim_of()
    assert not ref_sim.engine.emits_sumsq
    ref_sim.engine.p1 = ref_sim.engine.p2 = 0.0
    ref_sim.run_round()
    assert calls
    torch.testing.assert_close(sim.arena.master, ref_sim.arena.master, rtol=1e-3, atol=2e-4)


def test_adam_on_the_grouped_engine_applies_torch_adam_to_the_engines_gradient():
    """``optimizer="adam"`` on the hand-written MNIST engine: the step's materialised gradient goes
    through the fused Adam kernel of K-OPT.  Two steps, compared with the closed-form Adam update
    computed in torch from the SAME gradients (taken from the engine between backward and update),
    so the check is exact up to fp32 rounding and independent of bf16 activation noise."""
    from nanofed_b200 import ops
    from nanofed_b200.ops import tc
    from nanofed_b200.parallel.cohort import CohortConfig, FedSimulation

    lr, b1, b2, eps, wd = 1e-3, 0.9, 0.999, 1e-8, 0.01
    cfg = CohortConfig(model="mnist_cnn", clients_per_gpu=3, samples_per_client=128, batch_size=64,
                       learning_rate=lr, weight_decay=wd, optimizer="adam", device="cuda",
                       use_cuda_graph=False)
    sim = FedSimulation(cfg)
    assert sim.engine_name == "grouped" and sim.engine.adam and not sim.engine.fuse_fc1_sgd
    a, eng = sim.arena, sim.engine
    ops.bcast_global(a.master, a.slots, a.slots_lowp, a.momentum)
    eng.after_broadcast() if hasattr(eng, "after_broadcast") else None
    p = a.slots.clone().double()
    m = torch.zeros_like(p)
    v = torch.zeros_like(p)
    for step in range(2):
        eng._step_group(step, 0)
        torch.cuda.synchronize()
        tc.check_status(sim.device)
        g = eng.grads.clone().double() + wd * p
        eng._apply_optimizer(step)
        torch.cuda.synchronize()
        m = b1 * m + (1 - b1) * g
        v = b2 * v + (1 - b2) * g * g
        t = step + 1
        p = p - (lr / (1 - b1 ** t)) * m / (v.sqrt() / (1 - b2 ** t) ** 0.5 + eps)
        torch.testing.assert_close(a.slots.double(), p, rtol=1e-4, atol=1e-6)
        assert torch.equal(a.slots_lowp, a.slots.bfloat16())
        lo, hi = eng._fc1_range
        assert float(eng.grads[:, :lo].abs().sum()) == 0.0 and float(eng.grads[:, hi:].abs().sum()) == 0.0
        p = a.slots.clone().double()          # continue from the kernel's fp32 result
    # and a whole round through the graph trains
    sim2 = FedSimulation(CohortConfig(model="mnist_cnn", clients_per_gpu=4, samples_per_client=256,
                                      batch_size=64, learning_rate=1e-3, optimizer="adam",
                                      device="cuda"))
    losses = [sim2.run_round().agg_metrics["loss"] for _ in range(4)]
    assert losses[-1] < losses[0], losses

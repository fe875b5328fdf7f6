"""Trainers, callbacks, flat layout, logger."""

import asyncio
import json

import pytest
import torch
import torch.nn as nn
from torch.utils.data import DataLoader, TensorDataset

from nanofed_b200.models import MNISTModel, ResNet18
from nanofed_b200.privacy import PrivacyConfig
from nanofed_b200.runtime.flat import FlatLayout
from nanofed_b200.trainer import (MetricsLogger, PrivateTrainer, TorchTrainer, TrainingConfig,
                                  TrainingMetrics)
from nanofed_b200.utils import Logger, log_exec


def _loader(n=48, b=16):
    g = torch.Generator().manual_seed(0)
    return DataLoader(TensorDataset(torch.randn(n, 10, generator=g),
                                    torch.randint(0, 3, (n,), generator=g)), batch_size=b)


def test_train_epoch_returns_last_batch_and_exposes_averages(tmp_path):
    seen = []

    class Spy:
        def on_eopch_start(self, epoch): seen.append(("start", epoch))
        def on_epoch_end(self, epoch, m): seen.append(("end", m))
        def on_batch_end(self, batch, m): seen.append(("batch", batch, m.samples_processed))

    logger_cb = MetricsLogger(tmp_path, "exp")
    trainer = TorchTrainer(TrainingConfig(epochs=1, batch_size=16, learning_rate=0.1),
                           callbacks=[Spy(), logger_cb])
    model = nn.Sequential(nn.Linear(10, 3))
    opt = torch.optim.SGD(model.parameters(), lr=0.1)
    m = trainer.train_epoch(model, _loader(), opt, epoch=4)
    assert isinstance(m, TrainingMetrics) and m.batch == 2 and m.samples_processed == 48
    assert m.loss > 0 and 0 <= m.accuracy <= 1 and m.epoch == 4
    assert set(m.to_dict()) == {"loss", "accuracy", "samples_processed"}
    avg = trainer.last_epoch_metrics
    assert avg is not None and avg.samples_processed == 48
    assert [s[0] for s in seen] == ["start", "batch", "batch", "batch", "end"]
    assert [s[2] for s in seen if s[0] == "batch"] == [16, 32, 48]
    data = json.loads(next(tmp_path.glob("exp_*.json")).read_text())
    assert [d["type"] for d in data] == ["batch", "batch", "batch", "epoch"]
    m2 = TorchTrainer(TrainingConfig(1, 16, 0.1, max_batches=1)).train_epoch(model, _loader(), opt, 0)
    assert m2.samples_processed == 16


def test_private_trainer_clips_noises_and_accounts():
    cfg = PrivacyConfig(epsilon=1.0, max_gradient_norm=0.5, noise_multiplier=1.0)
    calls = []

    class Spy:
        def on_eopch_start(self, e): pass
        def on_epoch_end(self, e, m): calls.append("end")
        def on_batch_end(self, b, m): calls.append("batch")

    t = PrivateTrainer(TrainingConfig(1, 16, 0.1), cfg, callbacks=[Spy()])
    model = nn.Sequential(nn.Linear(10, 3))
    x, y = torch.randn(8, 10) * 100, torch.randint(0, 3, (8,))
    nn.functional.cross_entropy(model(x), y).backward()
    t._clip_gradients(model)
    total = torch.sqrt(sum(p.grad.pow(2).sum() for p in model.parameters()))
    assert float(total) <= 0.5 + 1e-5
    opt = torch.optim.SGD(model.parameters(), lr=0.01)
    m = t.train_batch(model, (x, y), opt)
    assert m.samples_processed == 8 and t._accountant._event_count == 1
    t.train_epoch(model, _loader(), opt, 0)                 # private epoch, callbacks forwarded
    assert t._accountant._event_count == 4 and calls.count("batch") == 3
    assert t.get_privacy_spent().epsilon_spent > 0
    tight = PrivateTrainer(TrainingConfig(1, 16, 0.1), PrivacyConfig(epsilon=0.01))
    tight.train_batch(model, (x, y), opt)
    assert tight.validate_privacy_budget() is False


@pytest.mark.parametrize("factory", [MNISTModel, ResNet18])
def test_flat_layout_roundtrip_keeps_dtypes(factory):
    model = factory()
    state = model.state_dict()
    layout = FlatLayout.from_module(model)
    assert layout.total % 64 == 0 and layout.num_values == sum(v.numel() for v in state.values())
    flat = layout.flatten(state)
    back = layout.unflatten(flat)
    for name, v in state.items():
        assert back[name].dtype == v.dtype and back[name].shape == v.shape
        assert torch.equal(back[name], v)
    mask = layout.float_mask()
    n_float = sum(v.numel() for v in state.values() if v.is_floating_point())
    assert int(mask.sum()) == n_float
    slab = layout.stack([state, state])
    assert slab.shape == (2, layout.total)
    other = factory()
    layout.load_into(other, flat)
    assert all(torch.equal(a, b) for a, b in zip(other.state_dict().values(), state.values()))
    if factory is MNISTModel:
        assert layout.num_values == 1_199_882 and len(layout) == 8


def test_logger_context_is_per_task_and_log_exec_times(caplog):
    log = Logger()
    assert Logger() is log
    seen = {}

    async def worker(name, delay):
        with log.context(name, "sub"):
            await asyncio.sleep(delay)
            seen[name] = str(log.current_context)

    async def main():
        await asyncio.gather(worker("a", 0.02), worker("b", 0.01))

    asyncio.run(main())
    assert seen == {"a": "a.sub", "b": "b.sub"}             # no cross-task mislabelling

    @log_exec
    def f(x):
        return x + 1

    @log_exec
    async def g(x):
        return x * 2

    assert f(1) == 2 and asyncio.run(g(2)) == 4


def test_per_sample_clipping_matches_explicit_per_sample_gradients():
    torch.manual_seed(0)
    model = nn.Sequential(nn.Conv2d(1, 4, 3), nn.ReLU(), nn.Flatten(), nn.Linear(4 * 6 * 6, 5))
    x, y = torch.randn(6, 1, 8, 8), torch.randint(0, 5, (6,))
    cfg = PrivacyConfig(max_gradient_norm=0.7, noise_multiplier=1e-9)       # ~no noise
    t = PrivateTrainer(TrainingConfig(1, 6, 0.1), cfg, per_sample_clipping=True)
    norms, _, _ = t.per_sample_grad_norms(model, x, y)
    # explicit per-sample gradients
    explicit, clipped_sum = [], [torch.zeros_like(p) for p in model.parameters()]
    for i in range(6):
        model.zero_grad()
        nn.functional.cross_entropy(model(x[i : i + 1]), y[i : i + 1]).backward()
        g = [p.grad.clone() for p in model.parameters()]
        n = torch.sqrt(sum(v.pow(2).sum() for v in g))
        explicit.append(n)
        c = min(0.7 / (float(n) + 1e-6), 1.0)
        for acc, v in zip(clipped_sum, g):
            acc += c * v
    torch.testing.assert_close(norms, torch.stack(explicit), rtol=1e-4, atol=1e-5)
    before = [p.detach().clone() for p in model.parameters()]
    opt = torch.optim.SGD(model.parameters(), lr=1.0)
    model.zero_grad()
    t.train_batch(model, (x, y), opt)
    for p, b0, g in zip(model.parameters(), before, clipped_sum):
        torch.testing.assert_close(b0 - p.detach(), g / 6, rtol=1e-3, atol=1e-5)
    with pytest.raises(NotImplementedError):
        t.per_sample_grad_norms(nn.Sequential(nn.Flatten(), nn.Linear(64, 5), nn.LayerNorm(5)), x, y)

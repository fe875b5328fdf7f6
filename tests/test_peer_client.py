"""In-box client API: the reference's client loop (HTTPClient-shaped) on the arena transport."""

import asyncio

import pytest
import torch

from nanofed_b200.communication import ClientSpec, PeerClient, PeerHub
from nanofed_b200.core.exceptions import NanoFedError
from nanofed_b200.models import MNISTModel
from nanofed_b200.trainer import TrainingConfig


def _data(n, seed):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(n, 1, 28, 28, generator=g), torch.randint(0, 10, (n,), generator=g)


def test_reference_shaped_client_loop_runs_unchanged_against_the_hub():
    """The loop of examples/mnist/run_experiment.py with PeerClient in HTTPClient's place."""
    torch.manual_seed(0)
    hub = PeerHub(MNISTModel(), [ClientSpec(f"client_{k}", n) for k, n in enumerate((16, 8, 24))],
                  num_rounds=2, device="cpu")
    submitted: dict[int, list[dict]] = {0: [], 1: [], 2: []}
    rounds_seen: dict[int, list[int]] = {0: [], 1: [], 2: []}

    async def run_client(k: int, n: int) -> None:
        x, y = _data(n, 10 + k)
        async with hub.client(k) as client:
            while not await client.check_server_status():
                state, rnd = await client.fetch_global_model()
                rounds_seen[k].append(rnd)
                model = MNISTModel()
                model.load_state_dict(state)
                model.train()
                opt = torch.optim.SGD(model.parameters(), lr=0.05)
                loss = torch.nn.functional.cross_entropy(model(x), y)
                loss.backward()
                opt.step()
                submitted[k].append({n_: t.clone() for n_, t in model.state_dict().items()})
                ok = await client.submit_update(model, {"loss": float(loss), "num_samples": n})
                assert ok

    async def main():
        await asyncio.gather(*[run_client(k, n) for k, n in enumerate((16, 8, 24))])

    asyncio.run(main())
    assert hub.round == 2 and all(v == [0, 1] for v in rounds_seen.values())
    # the global model after the last round is the sample-weighted mean of the last submissions
    w = torch.tensor([16.0, 8, 24]) / 48
    got = hub.model.state_dict()
    for name in ("fc1.weight", "conv1.bias"):
        want = sum(w[k] * submitted[k][-1][name] for k in range(3))
        torch.testing.assert_close(got[name], want, rtol=1e-5, atol=1e-6)
    rec = hub.history[-1]
    assert rec["num_clients"] == 3 and abs(sum(c["weight"] for c in rec["client_metrics"]) - 1) < 1e-6
    assert [c["client_id"] for c in rec["client_metrics"]] == ["client_0", "client_1", "client_2"]


def test_declarative_specs_with_own_configs_and_bound_models():
    hub = PeerHub(MNISTModel(), [
        ClientSpec("a", 16, data=_data(16, 1),
                   training=TrainingConfig(epochs=1, batch_size=8, learning_rate=0.05, device="cpu",
                                           log_interval=10**9)),
        ClientSpec("b", 8, data=_data(8, 2),
                   training=TrainingConfig(epochs=2, batch_size=4, learning_rate=0.01, device="cpu",
                                           log_interval=10**9),
                   optimizer=lambda m, lr: torch.optim.SGD(m.parameters(), lr=lr, momentum=0.9)),
    ], num_rounds=2, device="cpu")
    before = hub.arena.master.clone()
    history = asyncio.run(hub.run(MNISTModel))
    assert len(history) == 2 and history[0]["client_metrics"][0]["weight"] == pytest.approx(16 / 24)
    assert not torch.equal(before, hub.arena.master)
    # a model bound to its arena row uploads by training: submit copies nothing for parameters
    m = hub.bind_model(MNISTModel(), 0)
    assert m.fc1.weight.data_ptr() == hub.layout.views(hub.arena.slots[0])["fc1.weight"].data_ptr()


def test_client_errors_match_the_http_client():
    hub = PeerHub(MNISTModel(), 2, num_rounds=1, device="cpu")
    c = PeerClient(hub, 0)

    async def no_session():
        await c.fetch_global_model()

    with pytest.raises(NanoFedError, match="not initialized"):
        asyncio.run(no_session())
    with pytest.raises(NanoFedError):
        PeerClient(hub, 5)

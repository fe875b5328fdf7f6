"""End-to-end rounds over the real HTTP transport — the path the reference never tests
(and cannot complete at HEAD)."""

import asyncio
import json
import socket

import pytest
import torch
import torch.nn as nn
from torch.utils.data import DataLoader, TensorDataset

from nanofed_b200 import (Coordinator, CoordinatorConfig, FedAvgAggregator, HTTPClient, HTTPServer,
                          ModelManager, TorchTrainer, TrainingConfig, coordinate)
from nanofed_b200.communication.http import codec
from nanofed_b200.core import NanoFedError
from nanofed_b200.orchestration import RoundStatus
from nanofed_b200.server import FaultTolerantCoordinator


class Net(nn.Module):
    def __init__(self):
        super().__init__()
        self.fc = nn.Linear(8, 3)
        self.bn = nn.BatchNorm1d(3)

    def forward(self, x):
        return self.bn(self.fc(x))


def _port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


async def _client(url, cid, n, binary, rounds_seen):
    g = torch.Generator().manual_seed(hash(cid) % 1000)
    loader = DataLoader(TensorDataset(torch.randn(n, 8, generator=g),
                                      torch.randint(0, 3, (n,), generator=g)), batch_size=8)
    trainer = TorchTrainer(TrainingConfig(epochs=1, batch_size=8, learning_rate=0.05))
    async with HTTPClient(url, cid, binary=binary, timeout=30) as http:
        while True:
            if await http.check_server_status():
                return
            try:
                state, rnd = await http.fetch_global_model()
            except NanoFedError:
                await asyncio.sleep(0.02)
                continue
            if rnd in rounds_seen[cid]:
                await asyncio.sleep(0.02)
                continue
            model = Net()
            model.load_state_dict(state)
            opt = torch.optim.SGD(model.parameters(), lr=0.05)
            metrics = trainer.train_epoch(model, loader, opt, 0)
            if await http.submit_update(model, metrics):
                rounds_seen[cid].add(rnd)


@pytest.mark.parametrize("binary", [False, True])
def test_three_clients_two_rounds_files_on_disk(tmp_path, binary):
    async def main():
        server = HTTPServer("127.0.0.1", _port())
        await server.start()
        model = Net()
        coord = Coordinator(ModelManager(model), FedAvgAggregator(), server,
                            CoordinatorConfig(num_rounds=2, min_clients=3, min_completion_rate=1.0,
                                              round_timeout=30, base_dir=tmp_path),
                            fault_tolerance=FaultTolerantCoordinator(tmp_path))
        seen = {f"client_{i}": set() for i in range(3)}
        sizes = {"client_0": 32, "client_1": 16, "client_2": 8}
        try:
            await asyncio.wait_for(asyncio.gather(
                coordinate(coord),
                *[_client(server.url, cid, n, binary, seen) for cid, n in sizes.items()]), 60)
        finally:
            await server.stop()
        return coord, server, seen

    coord, server, seen = asyncio.run(main())
    assert coord.current_round == 2 and server.current_round == 2     # server round advances
    assert all(s == {0, 1} for s in seen.values())
    assert coord.training_progress["status"] == RoundStatus.COMPLETED.name
    assert coord.training_progress["active_clients"] == 3
    # checkpoint layout
    configs = sorted((tmp_path / "models" / "configs").glob("model_v_*.json"))
    weights = sorted((tmp_path / "models" / "models").glob("model_v_*.pt"))
    assert len(configs) == len(weights) == 3                          # initial + 2 rounds
    rec = json.loads(configs[-1].read_text())
    assert rec["config"]["round_id"] == 1 and rec["config"]["num_clients"] == 3
    assert abs(sum(rec["config"]["client_weights"].values()) - 1.0) < 1e-9
    assert rec["config"]["client_weights"]["client_0"] == pytest.approx(32 / 56)
    state = torch.load(weights[-1], weights_only=True)
    assert set(state) == set(Net().state_dict())
    m0 = json.loads((tmp_path / "metrics" / "metrics_round_0.json").read_text())
    assert set(m0) == {"round_id", "start_time", "end_time", "num_clients", "agg_metrics", "status",
                       "client_metrics"}
    assert m0["status"] == "COMPLETED" and len(m0["client_metrics"]) == 3
    assert (tmp_path / "checkpoints" / "round_1" / "state.pt").exists()
    assert (tmp_path / "data").is_dir()


def test_server_endpoints_errors_and_round_check(tmp_path):
    async def main():
        import aiohttp

        server = HTTPServer("127.0.0.1", _port())
        await server.start()
        out = {}
        try:
            async with aiohttp.ClientSession() as s:
                async with s.get(server.url + "/model") as r:
                    out["no_coord"] = (r.status, (await r.json())["message"])
                async with s.get(server.url + "/test") as r:
                    out["test"] = await r.text()
                Coordinator(ModelManager(Net()), FedAvgAggregator(), server,
                            CoordinatorConfig(1, 1, 1.0, 5, tmp_path))
                async with s.get(server.url + "/model") as r:
                    body = await r.json()
                    out["model"] = (r.status, body["status"], body["round_number"],
                                    sorted(body["model_state"]))
                async with s.post(server.url + "/update", json={"client_id": "x"}) as r:
                    out["missing"] = (r.status, (await r.json())["message"])
                good = {"client_id": "x", "round_number": 5, "model_state": {}, "metrics": {},
                        "timestamp": "2025-01-01T00:00:00+00:00"}
                async with s.post(server.url + "/update", json=good) as r:
                    out["stale"] = (r.status, (await r.json())["message"])
                good["round_number"] = 0
                async with s.post(server.url + "/update", json=good) as r:
                    out["ok"] = await r.json()
                async with s.get(server.url + "/status") as r:
                    out["status"] = await r.json()
                await server.stop_training()
                async with s.get(server.url + "/model") as r:
                    out["done"] = await r.json()
        finally:
            await server.stop()
        return out

    out = asyncio.run(main())
    assert out["no_coord"] == (500, "Server not initialized with coordinator")
    assert out["test"] == "Server is running"
    assert out["model"][:3] == (200, "success", 0) and "fc.weight" in out["model"][3]
    assert out["missing"][0] == 400 and "Missing keys" in out["missing"][1]
    assert out["stale"] == (400, "Invalid round number")
    assert out["ok"]["accepted"] is True and out["ok"]["update_id"] == "update_x_0"
    assert out["status"]["num_updates"] == 1 and out["status"]["is_training_done"] is False
    assert out["done"]["status"] == "terminated" and out["done"]["round_number"] == -1


def test_round_timeout_and_binary_codec(tmp_path):
    async def main():
        server = HTTPServer("127.0.0.1", _port())
        await server.start()
        coord = Coordinator(ModelManager(Net()), FedAvgAggregator(), server,
                            CoordinatorConfig(1, 2, 1.0, 1, tmp_path))
        try:
            with pytest.raises(TimeoutError):
                await coord.train_round()
        finally:
            await server.stop()
        return coord

    coord = asyncio.run(main())
    assert coord.training_progress["status"] == "FAILED"
    state = {"a": torch.randn(3, 5), "b": torch.tensor(7), "c": torch.randn(4).bfloat16(),
             "d": torch.zeros(0)}
    header, back = codec.decode_frame(codec.encode_frame({"round_number": 3}, state))
    assert header["round_number"] == 3
    for k, v in state.items():
        assert back[k].dtype == v.dtype and back[k].shape == v.shape and torch.equal(back[k], v)
    js = codec.state_to_jsonable(state)
    assert js["b"] == 7 and isinstance(js["c"], list)          # bf16 widened instead of failing

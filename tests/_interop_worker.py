"""Worker for tests/test_wire_interop.py: the UNMODIFIED reference (baseline/_ref, imported as
``nanofed``) and this package (``nanofed_b200``) in one process, talking over localhost HTTP.

  mode "ref_client":  reference HTTPClient / TorchTrainer  <->  our HTTPServer + Coordinator
                      (one full FedAvg round, two clients)
  mode "ref_server":  our HTTPClient  <->  reference HTTPServer (+ its ModelManager): fetch the
                      global model, submit an update, read /status
  mode "checkpoint":  a model version written by either ModelManager is loaded by the other

Prints INTEROP_OK on success, exits non-zero otherwise.
"""
import asyncio
import logging
import socket
import sys
import tempfile
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "baseline" / "_ref"))
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402
from torch.utils.data import DataLoader, TensorDataset  # noqa: E402

import nanofed as ref  # noqa: E402  (the reference)
import nanofed_b200 as ours  # noqa: E402
from nanofed.models import MNISTModel as RefModel  # noqa: E402
from nanofed_b200.models import MNISTModel as OurModel  # noqa: E402

logging.getLogger("nanofed").setLevel(logging.ERROR)
assert "baseline/_ref" in ref.__file__, ref.__file__


def free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def loader(seed: int, n: int = 32) -> DataLoader:
    g = torch.Generator().manual_seed(seed)
    return DataLoader(TensorDataset(torch.randn(n, 1, 28, 28, generator=g),
                                    torch.randint(0, 10, (n,), generator=g)), batch_size=16)


async def ref_client_vs_our_server() -> None:
    server = ours.HTTPServer("127.0.0.1", free_port(), max_request_size=1 << 30)
    await server.start()
    base = Path(tempfile.mkdtemp(prefix="nfb_interop_"))
    coord = ours.Coordinator(ours.ModelManager(OurModel()), ours.FedAvgAggregator(), server,
                             ours.CoordinatorConfig(1, 2, 1.0, 120, base))
    before = {k: v.clone() for k, v in coord.model_manager.model.state_dict().items()}
    sent = []

    async def client(cid: int) -> None:
        trainer = ref.TorchTrainer(ref.TrainingConfig(epochs=1, batch_size=16, learning_rate=0.1))
        async with ref.HTTPClient(server.url, f"client_{cid}") as http:   # reference client, JSON wire
            state, rnd = await http.fetch_global_model()
            assert rnd == 0 and set(state) == set(before)
            model = RefModel()
            model.load_state_dict(state)
            opt = torch.optim.SGD(model.parameters(), lr=0.1)
            metrics = trainer.train_epoch(model, loader(cid), opt, 0)
            ok = await http.submit_update(model, metrics)
            assert ok, "our server rejected the reference client's update"
            sent.append({k: v.detach().clone() for k, v in model.state_dict().items()})

    async def drive() -> list:
        return [m async for m in coord.start_training()]

    rounds, *_ = await asyncio.gather(drive(), client(0), client(1))
    await server.stop()
    assert len(rounds) == 1 and rounds[0].num_clients == 2
    after = coord.model_manager.model.state_dict()
    # equal sample counts -> plain mean of the two submitted models
    for k in after:
        want = (sent[0][k].float() + sent[1][k].float()) / 2
        torch.testing.assert_close(after[k].float(), want, rtol=1e-5, atol=1e-6)
    assert (base / "metrics" / "metrics_round_0.json").exists()


async def our_client_vs_ref_server() -> None:
    server = ref.HTTPServer("127.0.0.1", free_port())
    base = Path(tempfile.mkdtemp(prefix="nfb_interop_ref_"))
    manager = ref.ModelManager(RefModel())
    # the reference Coordinator constructor wires manager dirs + server.set_coordinator
    ref.Coordinator(manager, ref.FedAvgAggregator(), server,
                    ref.CoordinatorConfig(num_rounds=1, min_clients=1, min_completion_rate=1.0,
                                          round_timeout=60, base_dir=base))
    await server.start()
    try:
        async with ours.HTTPClient(server.url, "client_ours") as http:    # JSON wire (default)
            state, rnd = await http.fetch_global_model()
            assert rnd == 0
            want = manager.model.state_dict()
            assert set(state) == set(want)
            for k in want:
                torch.testing.assert_close(state[k].float(), want[k].float())
            model = OurModel()
            model.load_state_dict(state)
            trainer = ours.TorchTrainer(ours.TrainingConfig(epochs=1, batch_size=16,
                                                            learning_rate=0.1))
            opt = torch.optim.SGD(model.parameters(), lr=0.1)
            metrics = trainer.train_epoch(model, loader(7), opt, 0)
            assert await http.submit_update(model, metrics), "reference server rejected our update"
            assert await http.check_server_status() is False          # training not finished
        stored = server._updates["client_ours"]
        assert stored["round_number"] == 0 and set(stored["model_state"]) == set(want)
        assert stored["metrics"]["samples_processed"] == metrics.samples_processed
    finally:
        await server.stop()


def checkpoints_both_ways() -> None:
    """A version written by one ModelManager is loaded by the other (layout of SURVEY 5.4)."""
    for writer, reader, w_model, r_model in ((ours, ref, OurModel, RefModel),
                                             (ref, ours, RefModel, OurModel)):
        base = Path(tempfile.mkdtemp(prefix="nfb_ckpt_"))
        models, configs = base / "models", base / "configs"
        models.mkdir(), configs.mkdir()
        torch.manual_seed(3)
        src = w_model()
        wm = writer.ModelManager(src)
        wm.set_dirs(models, configs)                       # seeds version 001
        with torch.no_grad():
            for p_ in src.parameters():
                p_.add_(0.25)
        version = wm.save_model({"round_id": 7, "note": "interop"})
        rm = reader.ModelManager(r_model())
        rm.set_dirs(models, configs)                       # finds existing versions, saves nothing
        loaded = rm.load_model()                           # latest = the version saved above
        assert loaded.version_id == version.version_id, (loaded.version_id, version.version_id)
        assert loaded.config.get("round_id") == 7
        for (k, a), (_, b) in zip(src.state_dict().items(), rm.model.state_dict().items()):
            assert torch.equal(a, b), k
        assert len(rm.list_versions()) == len(wm.list_versions()) == 2

    # fault-tolerance store: <dir>/checkpoints/round_N/{metadata.json,state.pt}
    from datetime import datetime, timezone

    import nanofed.server.fault_tolerance as ref_ft
    import nanofed_b200.server.fault_tolerance as our_ft

    for w_ft, r_ft in ((our_ft, ref_ft), (ref_ft, our_ft)):
        base = Path(tempfile.mkdtemp(prefix="nfb_ft_"))
        update = {"model_state": {"w": [[1.0, 2.0], [3.0, 4.0]]}, "client_id": "c0",
                  "round_number": 3, "metrics": {"loss": 0.5, "samples_processed": 8},
                  "timestamp": datetime.now(timezone.utc).isoformat()}
        meta = w_ft.CheckpointMetadata(round_id=3, timestamp=datetime.now(timezone.utc),
                                       num_clients=1, client_updates={"c0": update},
                                       global_model_version="model_v_x_001",
                                       state=w_ft.RoundState.COMPLETED)
        w_ft.FileStateStore(base).save_checkpoint(meta, {"w": torch.ones(2, 2)})
        got = r_ft.FileStateStore(base).load_checkpoint(3)
        assert got is not None
        rmeta, rstate = got
        assert rmeta.round_id == 3 and rmeta.state.name == "COMPLETED" and rmeta.num_clients == 1
        assert torch.equal(rstate["w"], torch.ones(2, 2))
        assert torch.equal(torch.as_tensor(rmeta.client_updates["c0"]["model_state"]["w"]),
                           torch.tensor([[1.0, 2.0], [3.0, 4.0]]))
        assert [m.round_id for m in r_ft.FileStateStore(base).list_checkpoints()] == [3]


if __name__ == "__main__":
    mode = sys.argv[1]
    if mode == "checkpoint":
        checkpoints_both_ways()
    else:
        asyncio.run(ref_client_vs_our_server() if mode == "ref_client"
                    else our_client_vs_ref_server())
    print("INTEROP_OK", mode, flush=True)

"""MNIST federated learning, two ways.

    python examples/mnist/run_experiment.py http      # reference-style: HTTP server + 3 client coroutines
    python examples/mnist/run_experiment.py inbox     # B200-style: 32 simulated clients per GPU, fused kernels

``http`` mirrors the reference's example (``examples/mnist/run_experiment.py`` there): 3
clients with 12 000 / 8 000 / 4 000 samples, 2 local epochs, SGD lr 0.1, 2 rounds, checkpoints
under ``runs/`` — with synthetic MNIST-shaped data when the dataset cannot be downloaded.
``inbox`` runs the same FedAvg on the GPUs of this box through ``FedSimulation`` and writes
the same checkpoint layout through ``Coordinator(engine=...)``.
"""

import asyncio
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[2]))

from nanofed_b200 import (Coordinator, CoordinatorConfig, FedAvgAggregator, FedSimulation,  # noqa: E402
                          CohortConfig, HTTPClient, HTTPServer, ModelManager, TorchTrainer,
                          TrainingConfig, coordinate)
from nanofed_b200.data import load_mnist_data  # noqa: E402
from nanofed_b200.models import MNISTModel  # noqa: E402


async def run_client(client_id: str, coordinator: Coordinator, num_samples: int) -> None:
    loader = load_mnist_data(coordinator.data_dir, batch_size=64, train=True,
                             subset_fraction=num_samples / 60000, num_workers=0)
    cfg = TrainingConfig(epochs=2, batch_size=64, learning_rate=0.1,
                         device="cuda" if torch.cuda.is_available() else "cpu", log_interval=50)
    trainer = TorchTrainer(cfg)
    done_rounds: set[int] = set()
    async with HTTPClient(coordinator.server.url, client_id, binary=True) as client:
        while not await client.check_server_status():
            state, rnd = await client.fetch_global_model()
            if rnd in done_rounds:
                await asyncio.sleep(0.2)
                continue
            model = MNISTModel()
            model.load_state_dict(state)
            model.to(cfg.device)
            opt = torch.optim.SGD(model.parameters(), lr=cfg.learning_rate)
            metrics = None
            for epoch in range(cfg.epochs):
                metrics = await asyncio.to_thread(trainer.train_epoch, model, loader, opt, epoch)
            if metrics and await client.submit_update(model, metrics):
                done_rounds.add(rnd)


async def main_http() -> None:
    server = HTTPServer("127.0.0.1", 8080)
    await server.start()
    coordinator = Coordinator(
        ModelManager(MNISTModel()), FedAvgAggregator(), server,
        CoordinatorConfig(num_rounds=2, min_clients=3, min_completion_rate=1.0, round_timeout=300,
                          base_dir=Path("runs/")))
    try:
        await asyncio.gather(coordinate(coordinator),
                             run_client("client_1", coordinator, 12000),
                             run_client("client_2", coordinator, 8000),
                             run_client("client_3", coordinator, 4000))
    finally:
        await server.stop()


async def main_inbox() -> None:
    sim = FedSimulation(CohortConfig(model="mnist_cnn", clients_per_gpu=32, samples_per_client=512))
    coordinator = Coordinator(
        ModelManager(MNISTModel()), FedAvgAggregator(), None,
        CoordinatorConfig(num_rounds=5, min_clients=sim.num_clients, min_completion_rate=1.0,
                          round_timeout=300, base_dir=Path("runs_inbox/")), engine=sim)
    async for m in coordinator.start_training():
        print(f"round {m.round_id}: {m.agg_metrics}")


if __name__ == "__main__":
    mode = sys.argv[1] if len(sys.argv) > 1 else "inbox"
    asyncio.run(main_http() if mode == "http" else main_inbox())

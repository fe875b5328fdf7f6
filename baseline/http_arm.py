"""BASELINE config 1 on THIS framework: MNIST CNN FedAvg, 3 clients, 1 local epoch, CPU,
localhost HTTP — the reference's example shape, run through nanofed_b200's own HTTP stack.

    python baseline/http_arm.py [--binary] [--clients 3] [--samples 512] [--rounds 3]

Prints one JSON line (rounds/s, client-samples/s, wall clock between completed rounds).
The reference side of the same config: ``python bench.py --impl reference --clients-per-gpu 3
--samples-per-client 512 --steps 2 --warmup 1`` on a machine without a GPU.
"""
import argparse, asyncio, json, socket, sys, tempfile, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from torch.utils.data import DataLoader, TensorDataset
from nanofed_b200 import (Coordinator, CoordinatorConfig, FedAvgAggregator, HTTPClient, HTTPServer,
                          ModelManager, TorchTrainer, TrainingConfig)
from nanofed_b200.models import MNISTModel

ap = argparse.ArgumentParser()
ap.add_argument("--binary", action="store_true")
ap.add_argument("--clients", type=int, default=3)
ap.add_argument("--samples", type=int, default=512)
ap.add_argument("--rounds", type=int, default=3)
ap.add_argument("--device", default="cpu")
args = ap.parse_args()
import logging
logging.getLogger("nanofed").setLevel(logging.ERROR)


def port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


async def client(url, cid, done):
    g = torch.Generator().manual_seed(1000 + cid)
    loader = DataLoader(TensorDataset(torch.randn(args.samples, 1, 28, 28, generator=g),
                                      torch.randint(0, 10, (args.samples,), generator=g)),
                        batch_size=64, shuffle=True)
    cfg = TrainingConfig(epochs=1, batch_size=64, learning_rate=0.1, device=args.device,
                         log_interval=10**9)
    trainer = TorchTrainer(cfg)
    seen = set()
    async with HTTPClient(url, f"client_{cid}", binary=args.binary) as http:
        while not done["stop"]:
            try:
                if await http.check_server_status():
                    break
                state, rnd = await http.fetch_global_model()
                if rnd in seen:
                    await asyncio.sleep(0.01)
                    continue
                model = MNISTModel()
                model.load_state_dict(state)
                model.to(cfg.device)
                opt = torch.optim.SGD(model.parameters(), lr=0.1)
                m = await asyncio.to_thread(trainer.train_epoch, model, loader, opt, 0)
                if await http.submit_update(model, m):
                    seen.add(rnd)
            except Exception:
                await asyncio.sleep(0.05)


async def main():
    server = HTTPServer("127.0.0.1", port(), max_request_size=1 << 30)
    await server.start()
    coord = Coordinator(ModelManager(MNISTModel()), FedAvgAggregator(), server,
                        CoordinatorConfig(args.rounds, args.clients, 1.0, 3600,
                                          Path(tempfile.mkdtemp(prefix="nfb_http_"))))
    marks = [time.perf_counter()]
    done = {"stop": False}

    async def drive():
        async for _ in coord.start_training():
            marks.append(time.perf_counter())
        done["stop"] = True

    await asyncio.gather(drive(), *[client(server.url, i, done) for i in range(args.clients)])
    await server.stop()
    d = [b - a for a, b in zip(marks, marks[1:])][1:] or [marks[-1] - marks[0]]
    sec = sum(d) / len(d)
    print(json.dumps({"impl": "ours-http", "binary": args.binary, "clients": args.clients,
                      "samples_per_client": args.samples, "device": args.device,
                      "sec_per_round": sec, "rounds_per_sec": 1 / sec,
                      "client_samples_per_sec": args.clients * args.samples / sec,
                      "timed_rounds": len(d)}))


asyncio.run(main())

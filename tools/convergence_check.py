"""Does the bf16 grouped engine train like fp32?  The SAME federation twice:

  A. ``FedSimulation(engine="grouped")`` - hand-written sm_100a kernels, bf16 operands, fp32
     accumulation and master weights, fused FedAvg kernel;
  B. the reference's recipe in fp32 with TF32 off - ``TorchTrainer.train_epoch`` per client on a
     fresh copy of the global model + ``FedAvgAggregator`` (reference
     ``nanofed/trainer/base.py:134-160``, ``nanofed/server/aggregator/fedavg.py:53-66``).

Same initial model, same client shards (learnable: every image is its class prototype at a low
signal-to-noise ratio plus unit Gaussian noise), same batch order, dropout off in both (its random
stream differs between the two implementations).  After every round the global model is evaluated
in fp32 on a held-out set drawn the same way.

    python tools/convergence_check.py [--rounds 30] [--out profiles/convergence_mnist.json]
"""

from __future__ import annotations

import argparse
import copy
import json
import sys
from datetime import datetime, timezone
from pathlib import Path

import torch
import torch.nn.functional as F

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False


def evaluate(model, x, y) -> tuple[float, float]:
    model.eval()
    with torch.no_grad():
        out = model(x)
        return float(F.cross_entropy(out, y)), float((out.argmax(1) == y).float().mean())


def prototype_shards(clients: int, samples: int, shape: tuple[int, ...], seed: int, snr: float = 0.25):
    """Class-prototype data: x = snr * prototype[y] + N(0, 1).  Learnable in a few dozen steps."""
    from nanofed_b200.data.synthetic import ClientShards

    protos = torch.randn(10, *shape, generator=torch.Generator().manual_seed(4711))
    gen = torch.Generator().manual_seed(seed)
    labels = torch.randint(0, 10, (clients, samples), generator=gen)
    images = snr * protos[labels] + torch.randn(clients, samples, *shape, generator=gen)
    return ClientShards(images, labels, torch.full((clients,), samples, dtype=torch.int32))


def run(rounds: int = 30, clients: int = 8, samples: int = 256, batch: int = 64, lr: float = 0.03,
        model_name: str = "mnist_cnn", verbose: bool = True, snr: float = 0.25) -> dict:
    from torch.utils.data import DataLoader, TensorDataset

    from nanofed_b200 import FedAvgAggregator, TorchTrainer, TrainingConfig
    from nanofed_b200.parallel.cohort import CohortConfig, FedSimulation, build_model

    import logging

    logging.getLogger("nanofed").setLevel(logging.WARNING)
    dev = torch.device("cuda")
    template = build_model(model_name)
    shape = tuple(getattr(template, "INPUT_SHAPE", (1, 28, 28)))
    shards = prototype_shards(clients, samples, shape, seed=77, snr=snr)
    held = prototype_shards(1, 2048, shape, seed=991, snr=snr)
    hx, hy = held.images[0].to(dev), held.labels[0].to(dev)

    cfg = CohortConfig(model=model_name, clients_per_gpu=clients, samples_per_client=samples,
                       batch_size=batch, learning_rate=lr, engine="grouped", device="cuda")
    sim = FedSimulation(cfg, shards=copy.deepcopy(shards))
    if hasattr(sim.engine, "p1"):
        sim.engine.p1 = sim.engine.p2 = 0.0          # before the first round captures the graph
    init = {k: v.clone() for k, v in sim.global_state().items()}

    # ---- B: the reference recipe in fp32 ---------------------------------------------------
    global_model = build_model(model_name).to(dev)
    global_model.load_state_dict(init)
    for m in global_model.modules():
        if isinstance(m, torch.nn.Dropout):
            m.p = 0.0
    trainer = TorchTrainer(TrainingConfig(epochs=1, batch_size=batch, learning_rate=lr,
                                          device="cuda", log_interval=10**9))
    aggregator = FedAvgAggregator()
    loaders = [DataLoader(TensorDataset(shards.images[k], shards.labels[k]), batch_size=batch,
                          shuffle=False) for k in range(clients)]
    eval_model = build_model(model_name).to(dev)
    curve = []
    for r in range(rounds):
        sim.run_round()
        eval_model.load_state_dict(sim.global_state())
        la, aa = evaluate(eval_model, hx, hy)
        updates = []
        for k in range(clients):
            local = copy.deepcopy(global_model)
            opt = torch.optim.SGD(local.parameters(), lr=lr)
            metrics = trainer.train_epoch(local, loaders[k], opt, 0)
            md = metrics.to_dict()
            md["num_samples"] = samples
            updates.append({"model_state": {n: t.detach().clone() for n, t in local.state_dict().items()},
                            "client_id": f"client_{k}", "round_number": r, "metrics": md,
                            "timestamp": datetime.now(timezone.utc)})
        aggregator.aggregate(global_model, updates)
        lb, ab = evaluate(global_model, hx, hy)
        curve.append({"round": r, "grouped_bf16": {"loss": la, "accuracy": aa},
                      "reference_fp32": {"loss": lb, "accuracy": ab}})
        if verbose:
            print(f"round {r:2d}: grouped bf16 loss {la:.4f} acc {aa:.4f} | fp32 TorchTrainer+FedAvg "
                  f"loss {lb:.4f} acc {ab:.4f}", flush=True)
    # the last three rounds, averaged: on the steep part of the curve a single round's value moves
    # by several per cent with the summation order of one fp32 atomic
    tail = curve[-3:]
    last = {arm: {m: sum(c[arm][m] for c in tail) / len(tail) for m in ("loss", "accuracy")}
            for arm in ("grouped_bf16", "reference_fp32")}
    summary = {
        "model": model_name, "rounds": rounds, "clients": clients, "samples_per_client": samples,
        "batch": batch, "lr": lr, "snr": snr, "engine": sim.engine_name,
        "final_loss_rel_diff": abs(last["grouped_bf16"]["loss"] - last["reference_fp32"]["loss"])
        / last["reference_fp32"]["loss"],
        "final_accuracy_diff": abs(last["grouped_bf16"]["accuracy"] - last["reference_fp32"]["accuracy"]),
        "initial_loss": curve[0]["reference_fp32"]["loss"], "curve": curve,
    }
    return summary


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--rounds", type=int, default=30)
    ap.add_argument("--model", default="mnist_cnn")
    ap.add_argument("--lr", type=float, default=0.03)
    ap.add_argument("--snr", type=float, default=0.25)
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    s = run(rounds=args.rounds, model_name=args.model, lr=args.lr, snr=args.snr)
    print(json.dumps({k: v for k, v in s.items() if k != "curve"}))
    if args.out:
        Path(args.out).parent.mkdir(parents=True, exist_ok=True)
        Path(args.out).write_text(json.dumps(s, indent=1))
    ok = s["final_loss_rel_diff"] < 0.02 and s["final_accuracy_diff"] < 0.01
    print("CONVERGENCE OK" if ok else "CONVERGENCE MISMATCH")
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())

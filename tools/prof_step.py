"""One eager local step of the flagship engine at the bench shape (32 clients x batch 64),
for `ncu --metrics gpu__time_duration.sum` launch lists and per-kernel captures."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from nanofed_b200 import ops
from nanofed_b200.parallel.cohort import CohortConfig, FedSimulation

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 4
sim = FedSimulation(CohortConfig(clients_per_gpu=32, samples_per_client=512, batch_size=64,
                                 engine="grouped", use_cuda_graph=False))
a = sim.arena
ops.bcast_global(a.master, a.slots, a.slots_lowp, a.momentum)
for s in range(steps):
    sim.engine.step(s)
torch.cuda.synchronize()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
ev[0].record()
for s in range(8):
    sim.engine.step(s)
ev[1].record()
torch.cuda.synchronize()
print(f"eager: {ev[0].elapsed_time(ev[1]) / 8:.3f} ms per local step (32 clients x 64)")

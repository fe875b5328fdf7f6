"""Three eager local steps of the ResNet-18 cohort engine at the bench shape (8 clients x batch 64):
the target of `ncu` launch lists and per-kernel captures (a step is 123 launches after the one-off
weight conversion; the third step starts at launch 247)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import os
os.environ.setdefault("NFB_RN_STREAMS", "1")   # one lane: per-launch event times stay meaningful
import torch
from nanofed_b200 import ops
from nanofed_b200.ops import native, tc
from nanofed_b200.parallel.cohort import CohortConfig, FedSimulation

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
model = sys.argv[2] if len(sys.argv) > 2 else "resnet18"      # or resnet18_gn
sim = FedSimulation(CohortConfig(model=model, clients_per_gpu=8, samples_per_client=128,
                                 batch_size=64, engine="grouped", use_cuda_graph=False, device="cuda"))
a = sim.arena
ops.bcast_global(a.master, a.slots, a.slots_lowp, a.momentum)
native.reset_launch_count()
sim.engine.after_broadcast()
for s in range(steps):
    before = native.LAUNCHES
    sim.engine.step(s % 2)
    per_step = native.LAUNCHES - before
torch.cuda.synchronize()
tc.check_status(sim.device)
print(f"ok: {steps} steps, {per_step} launches per step, {native.LAUNCHES} in total")

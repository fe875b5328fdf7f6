"""Per-kernel device times of one eager local step of the ResNet-18 cohort engine (CUDA events around
every launch, warm caches, bench shape: 8 clients x batch 64), grouped by kernel and shape, with the
tensor-core efficiency of each convolution against MEASURED_PEAKS.json (bf16_tflops_sustained).

    python tools/rn_prof.py [--clients 8] [--batch 64] [--reps 3]
"""

from __future__ import annotations

import argparse
import json
import sys
from collections import defaultdict
from pathlib import Path

import os
os.environ.setdefault("NFB_RN_STREAMS", "1")   # one lane: per-launch event times stay meaningful
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--clients", type=int, default=8)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--reps", type=int, default=3)
    args = ap.parse_args()
    from nanofed_b200 import ops
    from nanofed_b200.ops import rn, tc
    from nanofed_b200.parallel.cohort import CohortConfig, FedSimulation

    peak = 1361.0
    try:
        peak = float(json.loads((ROOT / "MEASURED_PEAKS.json").read_text())["bf16_tflops_sustained"])
    except Exception:
        pass
    cfg = CohortConfig(model="resnet18", clients_per_gpu=args.clients,
                       samples_per_client=args.batch * 2, batch_size=args.batch,
                       use_cuda_graph=False, device="cuda", engine="grouped")
    sim = FedSimulation(cfg)
    a = sim.arena
    ops.bcast_global(a.master, a.slots, a.slots_lowp, a.momentum)
    sim.engine.after_broadcast()
    records: list[tuple[str, float, torch.cuda.Event, torch.cuda.Event]] = []
    recording = {"on": False}

    def wrap(name, label_fn):
        fn = getattr(rn, name)

        def timed(*p, **kw):
            if not recording["on"]:
                return fn(*p, **kw)
            label, flops = label_fn(*p, **kw)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            out = fn(*p, **kw)
            e1.record()
            records.append((label, flops, e0, e1))
            return out

        setattr(rn, name, timed)

    def conv_label(dgrad, a_, w, out, **kw):
        taps = kw["taps"]
        k_total = sum((a_.shape[3] if t[5] == 0 else kw["a1"].shape[3]) for t in taps)
        h, wd = kw["grid_hw"]
        pos = a_.shape[0] * h * wd
        flops = 2.0 * pos * kw["n_out"] * k_total
        kind = "dgrad" if dgrad else "fwd"
        par = "" if kw.get("out_parity") is None else f" cls{kw['out_parity']}"
        return (f"conv_{kind} {a_.shape[3]}->{kw['n_out']} @{h}x{wd} taps{len(taps)}{par} "
                f"tile{kw.get('tile')}"), flops

    def wgrad_label(x, dy, dwt, t_off, **kw):
        pos = dy.shape[0] * dy.shape[1] * dy.shape[2]
        flops = 2.0 * pos * dy.shape[3] * x.shape[3] * len(kw["taps"])
        return (f"conv_wgrad {x.shape[3]}->{dy.shape[3]} @{dy.shape[1]} taps{len(kw['taps'])} "
                f"split{kw.get('k_split')}{' +sgd' if kw.get('sgd') is not None else ''}"), flops

    wrap("conv_gemm", conv_label)
    wrap("conv_wgrad", wgrad_label)
    wrap("im2col3", lambda *p, **kw: ("im2col3", 0.0))
    wrap("bn_apply", lambda x, *p, **kw: (f"bn_apply C{x.shape[3]} @{x.shape[1]}"
                                          f"{' +res' if kw.get('res') is not None else ''}", 0.0))
    wrap("bn_bwd", lambda apply, dout, out, x, *p, **kw:
         (f"bn_bwd_{'apply' if apply else 'reduce'} C{x.shape[3]} @{x.shape[1]}"
          f"{' x2' if kw.get('x2') is not None else ''}", 0.0))
    wrap("head", lambda *p, **kw: ("head", 0.0))
    wrap("tap_sgd", lambda *p, **kw: ("tap_sgd", 0.0))
    wrap("small_sgd", lambda *p, **kw: ("small_sgd", 0.0))

    for _ in range(2):
        sim.engine.step(0)
    torch.cuda.synchronize()
    tc.check_status(sim.device)
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(args.reps):
        sim.engine.step(0)
    t1.record()
    torch.cuda.synchronize()
    plain_ms = t0.elapsed_time(t1) / args.reps
    recording["on"] = True
    for _ in range(args.reps):
        sim.engine.step(1)
    torch.cuda.synchronize()
    agg: dict[str, list[float]] = defaultdict(lambda: [0.0, 0.0, 0])
    for label, flops, e0, e1 in records:
        s = agg[label]
        s[0] += e0.elapsed_time(e1) * 1e3 / args.reps
        s[1] += flops / args.reps
        s[2] += 1
    total = sum(v[0] for v in agg.values())
    print(f"local step, {args.clients} clients x batch {args.batch}: {plain_ms:.3f} ms eager "
          f"(sum of per-launch event times {total / 1e3:.3f} ms); peak = {peak:.0f} TFLOP/s sustained (measured)")
    kinds: dict[str, float] = defaultdict(float)
    for label, (us, flops, n) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
        eff = f"{flops / (us * 1e-6) / 1e12:7.1f} TF/s {100 * flops / (us * 1e-6) / 1e12 / peak:5.1f}%" if flops else ""
        print(f"{us:9.1f} us {100 * us / total:5.1f}%  x{n // args.reps:<3d} {label:70s} {eff}")
        kinds[label.split()[0]] += us
    print("by kernel:")
    for k, us in sorted(kinds.items(), key=lambda kv: -kv[1]):
        print(f"{us:9.1f} us {100 * us / total:5.1f}%  {k}")
    return 0


if __name__ == "__main__":
    sys.exit(main())

"""Microbenchmark of the round-level kernels at the flagship shape (K=32, N=1.2M) and
the aggregation-sweep shapes.  Prints achieved bandwidth vs MEASURED_PEAKS.json.

    python tools/prof_agg.py [--big]        (CUDA events, warm-up 3, L2 flushed between iters)
"""
import argparse
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from nanofed_b200 import ops  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--big", action="store_true")
ap.add_argument("--iters", type=int, default=10)
args = ap.parse_args()
peaks = {}
try:
    peaks = json.loads((Path(__file__).resolve().parent.parent / "MEASURED_PEAKS.json").read_text())
except Exception:
    pass
hbm = peaks.get("hbm_gbs", 6650.0)
dev = "cuda"
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(fn, iters):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


shapes = [(32, 1_199_936, torch.float32), (32, 1_199_936, torch.bfloat16)]
if args.big:
    shapes += [(8, 350_000_000 // 64 * 64, torch.bfloat16), (16, 350_000_000 // 64 * 64, torch.bfloat16),
               (8, 1_000_000_000 // 64 * 64, torch.bfloat16)]
rows = []
for k, n, dt in shapes:
    x = torch.randn(k, n, device=dev, dtype=dt) if k * n < 2**31 else torch.empty(k, n, device=dev, dtype=dt).normal_()
    counts = torch.rand(k, device=dev) + 0.5
    out = torch.empty(n, device=dev)
    lowp = torch.empty(n, device=dev, dtype=torch.bfloat16)
    eb = x.element_size()
    ms = timeit(lambda: ops.fedavg_reduce(x, counts, out=out, out_lowp=lowp), args.iters)
    byt = k * n * eb + n * 6
    rows.append(("fedavg_reduce", k, n, str(dt), ms, byt / ms / 1e6, byt / ms / 1e6 / hbm))
    ss = torch.empty(k, device=dev)
    ms = timeit(lambda: ops.update_stats(x), args.iters)
    rows.append(("update_stats", k, n, str(dt), ms, k * n * eb / ms / 1e6, k * n * eb / ms / 1e6 / hbm))
    ss, _ = ops.update_stats(x)
    nm = torch.ones(n, dtype=torch.uint8, device=dev)
    ms = timeit(lambda: ops.fedavg_reduce(x, counts, out=out, out_lowp=lowp, sumsq=ss, clip_norm=1.0,
                                          noise_std=0.1, noise_mask=nm, seed=1), args.iters)
    byt = k * n * eb + n * 7
    rows.append(("fedavg_reduce+dp", k, n, str(dt), ms, byt / ms / 1e6, byt / ms / 1e6 / hbm))
    if dt == torch.float32 and k * n * 10 < 60e9:
        slots = torch.empty(k, n, device=dev)
        sl = torch.empty(k, n, device=dev, dtype=torch.bfloat16)
        mom = torch.empty(k, n, device=dev)
        ms = timeit(lambda: ops.bcast_global(out, slots, sl, mom), args.iters)
        byt = n * 4 + k * n * 10
        rows.append(("bcast_global", k, n, "f32+bf16+mom", ms, byt / ms / 1e6, byt / ms / 1e6 / hbm))
        g = torch.randn(k * n, device=dev)
        ms = timeit(lambda: ops.fused_sgd_(slots.view(-1), g, 0.1, lowp_out=sl.view(-1)), args.iters)
        byt = k * n * (4 + 4 + 4 + 2)
        rows.append(("fused_sgd", k, n, "f32", ms, byt / ms / 1e6, byt / ms / 1e6 / hbm))
    del x
print(f"{'kernel':18s} {'K':>4s} {'N':>12s} {'dtype':>16s} {'ms':>9s} {'GB/s':>9s} {'of measured HBM':>16s}")
for r in rows:
    print(f"{r[0]:18s} {r[1]:4d} {r[2]:12d} {r[3]:>16s} {r[4]:9.4f} {r[5]:9.1f} {r[6]:16.3f}")

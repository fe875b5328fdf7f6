"""BASELINE config 5: aggregation sweep across GPUs — fused peer kernel vs NCCL baseline.

    torchrun --nproc-per-node G tools/agg_sweep.py [--params 350000000] [--clients-per-gpu 8]

Per rank: K client updates of N fp32 parameters (random), sample-weighted FedAvg over all
G*K clients, result replicated on every rank.  Timed on the device (CUDA events), max over
ranks, L2 flushed between iterations.  Roofline = slower of HBM (K*N*4 B at measured copy
bandwidth) and NVLink (2 * 4N(G-1)/G B at the measured 770 GB/s per direction).
"""
import argparse, json, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch, torch.distributed as dist
from nanofed_b200.parallel.peer import NcclReducer, PeerReducer
from nanofed_b200.runtime import dist as rdist
from nanofed_b200.runtime.arena import Arena
from nanofed_b200.runtime.flat import FlatLayout

ap = argparse.ArgumentParser()
ap.add_argument("--params", type=int, nargs="+", default=[1_199_936, 350_000_000])
ap.add_argument("--clients-per-gpu", type=int, default=8)
ap.add_argument("--iters", type=int, default=5)
ap.add_argument("--nvls", action="store_true")
ap.add_argument("--both", action="store_true", help="time the P2P and the NVLS path in one run")
args = ap.parse_args()
info = rdist.init()
dev = info.device
G = info.world_size
peaks = {}
try:
    peaks = json.loads((Path(__file__).resolve().parent.parent / "MEASURED_PEAKS.json").read_text())
except Exception:
    pass
hbm = peaks.get("hbm_gbs", 6650.0) * 1e9
link = 770e9
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

def timed(fn, iters):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        torch.cuda.synchronize(); dist.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(rdist.all_reduce_max(a.elapsed_time(b), dev))
    ts.sort()
    return ts[len(ts) // 2]

for n in args.params:
    n = n // 64 * 64
    k = args.clients_per_gpu
    layout = FlatLayout.from_state({"w": torch.empty(0)})  # placeholder; build arena by hand
    layout.total = n
    layout.entries = ()
    arena = Arena.__new__(Arena)
    arena.layout, arena.num_clients, arena.device = layout, k, dev
    f32 = dict(dtype=torch.float32, device=dev)
    arena.master = torch.zeros(n, **f32)
    arena.master_lowp = torch.zeros(n, dtype=torch.bfloat16, device=dev)
    arena.slots = torch.empty(k, n, **f32).normal_(0, 0.02)
    arena.counts = torch.randint(1, 500, (k,), device=dev).float()
    arena.mask = torch.ones(k, **f32)
    nccl = NcclReducer(arena, info)
    t_nccl = timed(lambda: nccl.reduce(arena), args.iters)
    ref = arena.master.clone()
    t_hbm = k * n * 4 / hbm
    t_link = 2 * 4 * n * (G - 1) / G / link
    roof = max(t_hbm, t_link)
    for nvls in ([False, True] if args.both else [args.nvls]):
        peer = PeerReducer(arena, info, use_multicast=nvls)
        t_peer = timed(lambda: peer.reduce(arena), args.iters)
        peer.check()
        err = (arena.master - ref).abs().max().item()
        if info.is_main:
            print(json.dumps({"n_gpus": G, "params": n, "clients_per_gpu": k, "clients": k * G,
                              "fused_ms": round(t_peer, 4), "nccl_baseline_ms": round(t_nccl, 4),
                              "speedup_vs_nccl": round(t_nccl / t_peer, 3), "path": peer.path,
                              "roofline_ms": round(roof * 1e3, 4),
                              "bound": "hbm" if t_hbm >= t_link else "nvlink",
                              "fraction_of_roofline": round(roof * 1e3 / t_peer, 3),
                              "nccl_fraction_of_roofline": round(roof * 1e3 / t_nccl, 3),
                              "max_abs_diff_vs_nccl": err}), flush=True)
        del peer
        torch.cuda.empty_cache()
    del arena, nccl, ref
    torch.cuda.empty_cache()
rdist.shutdown()

"""BASELINE config 5: aggregation sweep across GPUs — fused peer kernel vs NCCL baseline.

    torchrun --nproc-per-node G tools/agg_sweep.py [--params 350000000] [--clients-per-gpu 8]

Per rank: K client updates of N fp32 parameters (random), sample-weighted FedAvg over all
G*K clients, result replicated on every rank.  Timed on the device (CUDA events), max over
ranks.  Two timings: "isolated" (host barrier + L2 flush before every call: includes the launch
skew between the G processes, tens of microseconds) and "b2b" (R calls back to back between one
pair of events: the steady-state cost of a call, which is what a training round pays).

Roofline = the slower of
  HBM     K * N * b bytes of client rows at the measured copy bandwidth, and
  NVLink  bytes through ONE direction of a GPU's links at the measured 770 GB/s:
          P2P  : partial shards pulled from G-1 peers (4N(G-1)/G) + fp32 and bf16 result shards
                 received from G-1 peers (6N(G-1)/G)                      = 10 N (G-1)/G
          NVLS : reduced shard returned by the switch (4N/G) + the peers' replicated fp32 + bf16
                 stores (6N(G-1)/G)                                       = (4 + 6(G-1)) N/G
--waves W feeds the K clients of a rank in W streaming waves through ONE row buffer of K/W rows
(accumulate / local-only calls; the cross-GPU phase runs once, with the last wave).
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
ap.add_argument("--dtype", default="fp32", choices=["fp32", "bf16"], help="dtype of the client rows")
ap.add_argument("--waves", type=int, default=1)
ap.add_argument("--b2b", type=int, default=10, help="calls per back-to-back measurement")
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

def timed_b2b(fn, reps):
    for _ in range(2):
        fn()
    torch.cuda.synchronize(); dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record(); torch.cuda.synchronize()
    return rdist.all_reduce_max(a.elapsed_time(b), dev) / reps


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
    assert k % args.waves == 0
    kw_rows = k // args.waves                       # rows resident at a time
    dt = torch.float32 if args.dtype == "fp32" else torch.bfloat16
    esz = 4 if args.dtype == "fp32" else 2
    layout = FlatLayout.from_state({"w": torch.empty(0)})  # placeholder; build arena by hand
    layout.total = n
    layout.entries = ()
    arena = Arena.__new__(Arena)
    arena.layout, arena.num_clients, arena.device = layout, kw_rows, dev
    f32 = dict(dtype=torch.float32, device=dev)
    arena.master = torch.zeros(n, **f32)
    arena.master_lowp = torch.zeros(n, dtype=torch.bfloat16, device=dev)
    arena.slots = torch.empty(kw_rows, n, dtype=dt, device=dev).normal_(0, 0.02)
    arena.counts = torch.randint(1, 500, (kw_rows,), device=dev).float()
    arena.mask = torch.ones(kw_rows, **f32)
    arena.stats = torch.zeros(kw_rows, 3, **f32)

    def waves_of(red):
        def call():
            for w in range(args.waves):
                red.reduce(arena, accumulate=w > 0, local_only=w < args.waves - 1)
        return call

    nccl = NcclReducer(arena, info)
    t_nccl = timed(waves_of(nccl), args.iters)
    t_nccl_b2b = timed_b2b(waves_of(nccl), args.b2b)
    ref = arena.master.clone()
    t_hbm = k * n * esz / hbm
    for nvls in ([False, True] if args.both else [args.nvls]):
        peer = PeerReducer(arena, info, use_multicast=nvls)
        t_peer = timed(waves_of(peer), args.iters)
        t_peer_b2b = timed_b2b(waves_of(peer), args.b2b)
        peer.check()
        err = (arena.master - ref).abs().max().item()
        if peer.dbg is not None:       # NANOFED_B200_AGG_TRACE=1: phase timestamps of one call
            torch.cuda.synchronize(); dist.barrier()
            for _ in range(3):
                peer.reduce(arena)
            torch.cuda.synchronize()
            t = peer.dbg.cpu().tolist()
            rel = [round((v - t[0]) / 1e3, 1) if v else None for v in t[:7]]
            gathered = [None] * G
            dist.all_gather_object(gathered, rel)
            if info.is_main:
                print(json.dumps({"trace_us_since_kernel_start": {
                    "fields": ["start", "local_done+flag", "peer_flags_seen", "scalars_read",
                               "shard_done", "completion_seen", "bcast_tail_done"],
                    "per_rank": gathered, "params": n, "path": peer.path}}), flush=True)
        is_nvls = peer.path.endswith("nvls")
        link_bytes = (4 + 6 * (G - 1)) * n / G if is_nvls else 10 * n * (G - 1) / G
        t_link = link_bytes / link
        roof = max(t_hbm, t_link)
        if info.is_main:
            print(json.dumps({"n_gpus": G, "params": n, "dtype": args.dtype, "clients_per_gpu": k,
                              "clients": k * G, "waves": args.waves,
                              "fused_ms": round(t_peer_b2b, 4),
                              "nccl_baseline_ms": round(t_nccl_b2b, 4),
                              "speedup_vs_nccl": round(t_nccl_b2b / t_peer_b2b, 3),
                              "fused_isolated_ms": round(t_peer, 4),
                              "nccl_isolated_ms": round(t_nccl, 4), "path": peer.path,
                              "roofline_ms": round(roof * 1e3, 4),
                              "bound": "hbm" if t_hbm >= t_link else "nvlink",
                              "hbm_ms": round(t_hbm * 1e3, 4), "nvlink_ms": round(t_link * 1e3, 4),
                              "fraction_of_roofline": round(roof * 1e3 / t_peer_b2b, 3),
                              "nccl_fraction_of_roofline": round(roof * 1e3 / t_nccl_b2b, 3),
                              "max_abs_diff_vs_nccl": err}), flush=True)
        del peer
        torch.cuda.empty_cache()
    del arena, nccl, ref
    torch.cuda.empty_cache()
rdist.shutdown()

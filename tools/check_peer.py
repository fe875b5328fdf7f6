"""Multi-GPU correctness of the fused peer kernel (run under torchrun with >= 2 ranks):
PeerReducer (P2P, and NVLS multimem if NANOFED_B200_NVLS=1) vs the NCCL baseline reducer."""
import os, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch, torch.distributed as dist
from nanofed_b200 import ops
from nanofed_b200.models import MNISTModel
from nanofed_b200.parallel.peer import NcclReducer, PeerReducer
from nanofed_b200.runtime import dist as rdist
from nanofed_b200.runtime.arena import Arena
from nanofed_b200.runtime.flat import FlatLayout

info = rdist.init()
dev = info.device
layout = FlatLayout.from_module(MNISTModel())
c = 8
def make():
    torch.manual_seed(100 + info.rank)
    a = Arena(layout, c, dev, with_grads=False)
    a.slots.normal_(0, 0.02)
    a.slots[:, ::64] = 0  # keep some structure
    a.counts.copy_(torch.randint(1, 500, (c,), device=dev).float())
    a.mask[info.rank % c] = 0
    return a
ref_arena, peer_arena = make(), make()
nccl = NcclReducer(ref_arena, info)
results = {}
for nvls in ([False, True] if os.environ.get("NANOFED_B200_NVLS") == "1" else [False]):
    try:
        peer = PeerReducer(peer_arena, info, use_multicast=nvls)
    except Exception as e:
        results[f"nvls={nvls}"] = f"setup failed: {e!r}"
        continue
    for dp in (False, True):
        kw = {}
        if dp:
            ss, _ = ops.update_stats(ref_arena.slots)
            kw = dict(sumsq=ss, clip_norm=1.0, noise_std=0.01, noise_mask=ref_arena.noise_mask, seed=5, offset=64)
        nccl.reduce(ref_arena, **kw)
        for _ in range(3):   # repeated calls exercise barrier reuse
            peer.reduce(peer_arena, **kw)
        torch.cuda.synchronize()
        peer.check()
        err = (peer_arena.master - ref_arena.master).abs().max().item()
        scale = ref_arena.master.abs().max().item()
        gathered = [torch.zeros_like(peer_arena.master) for _ in range(info.world_size)]
        dist.all_gather(gathered, peer_arena.master)
        same = all(torch.equal(g, gathered[0]) for g in gathered)
        results[f"path={peer.path} dp={dp}"] = f"max_err={err:.3e} (scale {scale:.3e}) all_ranks_identical={same} lowp_ok={torch.equal(peer_arena.master_lowp, peer_arena.master.bfloat16())}"
    # timing: 20 calls
    torch.cuda.synchronize(); dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20): peer.reduce(peer_arena)
    b.record(); torch.cuda.synchronize()
    t_peer = a.elapsed_time(b) / 20
    a.record()
    for _ in range(20): nccl.reduce(ref_arena)
    b.record(); torch.cuda.synchronize()
    results[f"time path={peer.path}"] = f"fused {t_peer*1000:.1f} us vs nccl-baseline {a.elapsed_time(b)/20*1000:.1f} us (K={c}/gpu, N={layout.total})"
if info.is_main:
    for k, v in results.items(): print(k, "::", v)
rdist.shutdown()

"""In-box transport: cross-GPU aggregation + broadcast over peer memory.

Replaces, for clients that live on the GPUs of one box, the reference's whole
HTTP round trip (``POST /update`` → ``Coordinator`` list→tensor rebuild →
``FedAvgAggregator`` → ``GET /model``).  Two reducers share one
interface, ``reduce(arena, **dp_kwargs)``:

* ``PeerReducer`` — the product path.  One launch of
  ``csrc/fedavg_allreduce.cu`` per rank does local weighted reduce → flag
  barrier → in-kernel reduce-scatter over NVLink (P2P loads or NVLS
  ``multimem.ld_reduce``) → FedAvg epilogue (1/Σn, DP noise, bf16 cast) →
  all-gather by storing the shard into every rank's global-model buffers (P2P
  stores or ``multimem.st``) → flag barrier.  Buffers live in a symmetric
  allocation obtained from ``torch.distributed._symmetric_memory`` (allocation +
  handle exchange only; every byte is moved by our kernel).
* ``NcclReducer`` — the baseline the north star names: per-rank weighted partial
  sum, ``dist.all_reduce``, then separate scale / noise / cast kernels.  Also the
  path used on CPU (gloo) for the host-logic tests.

``emulate_allreduce`` runs G emulated ranks of the fused kernel as ONE
cooperative launch on a single GPU, which is how it is unit-tested without
co-scheduling kernels that wait on each other.
"""

from __future__ import annotations

import ctypes as C
import os
from typing import Any

import torch
import torch.distributed as dist

from nanofed_b200 import ops
from nanofed_b200.runtime.arena import Arena
from nanofed_b200.runtime.dist import RankInfo
from nanofed_b200.utils import Logger

_ALIGN = 256
_MAX_CHUNKS = 64
_CHUNK_TARGET = 4 << 20  # elements per pipeline chunk (16 MB of fp32 partials)
_SPIN_NS = 10_000_000_000  # 10 s: far above any legitimate wait, far below a gpurun limit


def _round_up(v: int, a: int = _ALIGN) -> int:
    return (v + a - 1) // a * a


def chunk_elems_for(n: int) -> int:
    """Pipeline chunk: ~4 M elements, at most 64 chunks, multiple of 1024 elements."""
    chunk = max(_CHUNK_TARGET, -(-n // _MAX_CHUNKS))
    return min(_round_up(chunk, 1024), _round_up(n, 1024))


def _declare(lib: C.CDLL) -> None:
    if getattr(lib, "_nfb_allreduce_declared", False):
        return
    vp, i32, i64, f32, u64, u32 = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_uint64, C.c_uint32
    lib.nfb_fedavg_allreduce.restype = i32
    lib.nfb_fedavg_allreduce.argtypes = [vp, i32, i32, i32, i32, i64, i64, i64, f32, f32, vp, u64,
                                         u64, u32, i32, i64, i32, vp, vp, vp, vp, u64, i32, vp,
                                         i32, vp, vp, vp]
    lib.nfb_allreduce_max_blocks.restype = i32
    lib.nfb_allreduce_max_blocks.argtypes = [i32, i64]
    lib._nfb_allreduce_declared = True


class _Layout:
    """Byte offsets of the symmetric buffer's regions."""

    def __init__(self, n: int, g: int, stats_len: int = 0) -> None:
        self.n, self.g, self.stats_len = n, g, stats_len
        self.partial = 0
        self.master = _round_up(self.partial + 4 * n)
        self.lowp = _round_up(self.master + 4 * n)
        self.scal = _round_up(self.lowp + 2 * n)
        self.flags = self.scal + 2 * _ALIGN          # scal: 4 + 2g floats (g <= 32)
        self.stats = self.flags + _round_up(4 * (1 + _MAX_CHUNKS) * g)   # [g][stats_len] fp32
        self.total = self.stats + _round_up(4 * g * max(stats_len, 1))


class NcclReducer:
    """Baseline: torch partial sum + ``all_reduce`` + separate elementwise kernels."""

    path = "nccl_allreduce"

    status = None

    def __init__(self, arena: Arena, rank: RankInfo) -> None:
        self.rank = rank
        self.partial = torch.zeros_like(arena.master)
        self.scal = torch.zeros(2, dtype=torch.float32, device=arena.device)
        self.stats_all = torch.zeros(rank.world_size, arena.num_clients, 3, dtype=torch.float32,
                                     device=arena.device)

    def check(self) -> None:
        pass

    def all_stats(self) -> torch.Tensor:
        """Per-client round metrics of every rank, ``[G, C, 3]`` (filled by ``reduce``)."""
        return self.stats_all

    def reduce(self, arena: Arena, *, sumsq=None, clip_norm: float = 0.0, noise_std: float = 0.0,
               noise_mask=None, seed: int = 0, offset: int = 0, accumulate: bool = False,
               local_only: bool = False, bcast: tuple | None = None) -> None:
        ops.fedavg_reduce(arena.slots, arena.counts, out=self.partial, mask=arena.mask,
                          sumsq=sumsq, clip_norm=clip_norm, normalize=False, accumulate=accumulate)
        eff = arena.counts * arena.mask
        if accumulate:
            self.scal[0] += eff.sum()
            self.scal[1] += (eff * eff).sum()
        else:
            self.scal[0] = eff.sum()
            self.scal[1] = (eff * eff).sum()
        if local_only:
            return
        if dist.is_initialized():
            dist.all_reduce(self.partial)
            dist.all_reduce(self.scal)
            dist.all_gather_into_tensor(self.stats_all.view(-1), arena.stats.reshape(-1).contiguous())
        else:
            self.stats_all[0].copy_(arena.stats)
        alive = self.scal[0] > 0            # every client masked out: keep the previous model
        new = self.partial / torch.where(alive, self.scal[0], torch.ones_like(self.scal[0]))
        if noise_std > 0:
            z = torch.empty_like(arena.master)
            ops.philox_normal_(z, 1.0, seed, offset)
            if noise_mask is not None:
                z.mul_(noise_mask.to(z.dtype))
            new.add_(z * (noise_std * self.scal[1].sqrt() / self.scal[0].clamp_min(1e-30)))
        arena.master.copy_(torch.where(alive, new, arena.master))
        arena.master_lowp.copy_(arena.master)
        if bcast is not None:
            ops.bcast_global(arena.master, *bcast)


class PeerReducer:
    """Product path: the fused kernel over symmetric (peer-mapped) memory."""

    path = "fused_peer_kernel"

    def __init__(self, arena: Arena, rank: RankInfo, use_multicast: bool | None = None) -> None:
        import torch.distributed._symmetric_memory as symm

        from nanofed_b200.ops import native

        self.lib = native.lib()
        _declare(self.lib)
        self.rank = rank
        g, n, dev = rank.world_size, arena.layout.total, arena.device
        self.stats_len = arena.num_clients * 3
        self.lay = _Layout(n, g, self.stats_len)
        self.buf = symm.empty(self.lay.total, dtype=torch.uint8, device=dev)
        self.buf.zero_()
        group = dist.group.WORLD
        self.handle = symm.rendezvous(self.buf, group.group_name)
        ptrs = [int(p) for p in self.handle.buffer_ptrs]
        mc = int(getattr(self.handle, "multicast_ptr", 0) or 0)
        if use_multicast is None:
            # NVLS (in-switch reduction, replicated stores) is the faster path from 4 GPUs up
            # (8 x 350 M: 5.31 vs 5.88 ms); with 2 GPUs plain P2P loads win.  NANOFED_B200_NVLS=0/1
            # overrides.
            env = os.environ.get("NANOFED_B200_NVLS")
            use_multicast = (env == "1") if env in ("0", "1") else g >= 4
        self.mc_base = mc if (use_multicast and mc) else 0
        if self.mc_base:
            self.path = "fused_peer_kernel_nvls"

        lay = self.lay
        self.master = self.buf[lay.master : lay.master + 4 * n].view(torch.float32)
        self.lowp = self.buf[lay.lowp : lay.lowp + 2 * n].view(torch.bfloat16)
        self.master.copy_(arena.master)
        self.lowp.copy_(arena.master_lowp)
        arena.master, arena.master_lowp = self.master, self.lowp

        self.grid_ctr = torch.zeros(2 + _MAX_CHUNKS, dtype=torch.int32, device=dev)
        self.status = torch.zeros(1, dtype=torch.int32, device=dev)
        # NANOFED_B200_AGG_TRACE=1: the kernel records %globaltimer at its phase boundaries
        self.dbg = (torch.zeros(8, dtype=torch.int64, device=dev)
                    if os.environ.get("NANOFED_B200_AGG_TRACE") == "1" else None)
        self.sumsq_dummy = None
        rows = []
        for r in range(g):
            base = ptrs[r]
            mine = r == rank.rank
            rows.append([
                arena.slots.data_ptr() if mine else 0,
                arena.counts.data_ptr() if mine else 0,
                arena.mask.data_ptr() if mine else 0,
                0,  # sumsq: filled per call
                base + lay.partial, base + lay.scal, base + lay.flags,
                base + lay.master, base + lay.lowp,
                self.grid_ctr.data_ptr() if mine else 0,
                base + lay.stats,
            ])
        self.stats_all = self.buf[lay.stats : lay.stats + 4 * g * self.stats_len].view(
            torch.float32).view(g, arena.num_clients, 3)
        self.views_host = torch.tensor(rows, dtype=torch.int64)
        self.views = self.views_host.to(dev)
        self.epoch = 0
        self.chunk = chunk_elems_for(n)
        max_blocks = int(self.lib.nfb_allreduce_max_blocks(native._dt(arena.slots),
                                                           arena.num_clients))
        if max_blocks <= 0:
            raise RuntimeError("fedavg_allreduce: occupancy query failed")
        # one wave, at most 4 CTAs per SM: enough loads in flight to saturate HBM/NVLink
        self.grid = max(2, min(max_blocks, 4 * torch.cuda.get_device_properties(dev).multi_processor_count))
        dist.barrier()
        torch.cuda.synchronize(dev)

    def all_stats(self) -> torch.Tensor:
        """Per-client round metrics of every rank, ``[G, C, 3]``: written into this rank's symmetric
        buffer by the peers' kernels before their completion flag."""
        return self.stats_all

    def reduce(self, arena: Arena, *, sumsq=None, clip_norm: float = 0.0, noise_std: float = 0.0,
               noise_mask=None, seed: int = 0, offset: int = 0, accumulate: bool = False,
               local_only: bool = False, bcast: tuple | None = None) -> None:
        """One fused launch.  ``local_only`` (streaming waves, ``K·N·b`` beyond HBM): only add this
        wave's clients to the unnormalised partial; the last wave (``accumulate=True,
        local_only=False``) then does the single cross-GPU reduce-scatter / all-gather.
        ``bcast=(slots, slots_lowp, momentum)``: K-BCAST as the tail of the same kernel — once this
        rank owns the complete new model, all CTAs replicate it into the client rows."""
        from nanofed_b200.ops import native

        target = None
        if bcast is not None and not local_only:
            b_slots, b_lowp, b_mom = bcast
            rows = next(t.shape[0] for t in (b_slots, b_lowp, b_mom) if t is not None)
            target = native._BcastTarget(
                native._ptr(b_slots), b_slots.stride(0) if b_slots is not None else 0,
                native._ptr(b_lowp), b_lowp.stride(0) if b_lowp is not None else 0,
                native._ptr(b_mom), b_mom.stride(0) if b_mom is not None else 0, rows)

        r = self.rank.rank
        want = sumsq.data_ptr() if sumsq is not None else 0
        if int(self.views_host[r, 3]) != want:
            self.views_host[r, 3] = want
            self.views[r, 3] = want
        self.epoch += 1
        lay, mc = self.lay, self.mc_base
        code = self.lib.nfb_fedavg_allreduce(
            self.views.data_ptr(), r, self.rank.world_size, 1, native._dt(arena.slots),
            arena.slots.shape[0], arena.layout.total, arena.slots.stride(0),
            float(clip_norm), float(noise_std),
            noise_mask.data_ptr() if noise_mask is not None else None,
            seed & (2**64 - 1), offset, self.epoch, self.grid // 2, self.chunk, self.grid,
            (mc + lay.partial) if mc else None, (mc + lay.master) if mc else None,
            (mc + lay.lowp) if mc else None, self.status.data_ptr(), _SPIN_NS,
            (1 if accumulate else 0) | (2 if local_only else 0),
            arena.stats.data_ptr(), self.stats_len,
            C.byref(target) if target is not None else None,
            self.dbg.data_ptr() if self.dbg is not None else None,
            torch.cuda.current_stream(arena.device).cuda_stream,
        )
        native._check(code, "fedavg_allreduce")
        native._count()

    def check(self) -> None:
        """Raise if any barrier timed out (reads one int from the device)."""
        if int(self.status.item()) != 0:
            raise RuntimeError("fedavg_allreduce: cross-GPU barrier timed out")


def make_reducer(transport: str, arena: Arena, rank: RankInfo, allow_fallback: bool = False) -> Any:
    """``transport="peer"`` is the product path and fails loudly: if the symmetric allocation or the
    fused kernel cannot be set up, this raises unless the caller explicitly opted into the NCCL
    baseline (``allow_fallback=True`` / ``CohortConfig.allow_nccl_fallback``)."""
    log = Logger()
    if transport == "nccl" or arena.device.type != "cuda":
        return NcclReducer(arena, rank)
    try:
        return PeerReducer(arena, rank)
    except Exception as exc:  # symmetric memory unavailable on this box
        if not allow_fallback:
            raise RuntimeError(
                f"peer transport unavailable ({exc!r}); pass transport='nccl' or "
                "allow_nccl_fallback=True to run the NCCL baseline instead") from exc
        log.error(f"peer transport unavailable ({exc!r}); FALLING BACK to the NCCL baseline path")
        red = NcclReducer(arena, rank)
        red.path = "nccl_allreduce(fallback)"
        return red


def emulate_allreduce(
    slots: torch.Tensor,      # [G, K, N] on one GPU
    counts: torch.Tensor,     # [G, K]
    *,
    mask: torch.Tensor | None = None,
    sumsq: torch.Tensor | None = None,
    clip_norm: float = 0.0,
    noise_std: float = 0.0,
    noise_mask: torch.Tensor | None = None,
    seed: int = 0,
    offset: int = 0,
    calls: int = 1,
    chunk_elems: int = 0,
    waves: int = 1,
    stats: torch.Tensor | None = None,
    spin_ns: int = 2_000_000_000,
    master_init: torch.Tensor | None = None,
    return_stats: bool = False,
    bcast: tuple | None = None,
) -> tuple[torch.Tensor, ...]:
    """Run G emulated ranks of the fused kernel in one cooperative launch.

    Returns ``(master [G, N] fp32, lowp [G, N] bf16)`` — every row must be equal.  ``waves`` > 1
    feeds each rank's K clients in that many streaming waves (all but the last ``local_only``);
    ``stats`` ``[G, L]``: per-rank metric rows that every rank must end up holding
    (``return_stats=True`` appends the gathered ``[G, G, L]``).
    """
    from nanofed_b200.ops import native

    lib = native.lib()
    _declare(lib)
    g, k, n = slots.shape
    dev = slots.device
    stats_len = 0 if stats is None else int(stats.shape[1])
    lay = _Layout(n, g, stats_len)
    bufs = torch.zeros(g, lay.total, dtype=torch.uint8, device=dev)
    if master_init is not None:
        for r in range(g):
            bufs[r, lay.master : lay.master + 4 * n].view(torch.float32).copy_(master_init)
    ctr = torch.zeros(g, 2 + _MAX_CHUNKS, dtype=torch.int32, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    target = None
    if bcast is not None:        # (slots [G, R, N], lowp [G, R, N] | None, momentum [G, R, N] | None)
        b_slots, b_lowp, b_mom = bcast
        rows = next(t.shape[1] for t in (b_slots, b_lowp, b_mom) if t is not None)
        target = native._BcastTarget(
            native._ptr(b_slots), b_slots.stride(1) if b_slots is not None else 0,
            native._ptr(b_lowp), b_lowp.stride(1) if b_lowp is not None else 0,
            native._ptr(b_mom), b_mom.stride(1) if b_mom is not None else 0, rows)
    counts = counts.to(torch.float32).contiguous()
    mask = torch.ones_like(counts) if mask is None else mask.to(torch.float32).contiguous()
    assert k % waves == 0, "waves must divide the client count"
    kw = k // waves
    wave_views = []
    for w in range(waves):
        rows = []
        for r in range(g):
            base = bufs[r].data_ptr()
            rows.append([
                slots[r, w * kw].data_ptr(), counts[r, w * kw :].data_ptr(),
                mask[r, w * kw :].data_ptr(),
                sumsq[r, w * kw :].data_ptr() if sumsq is not None else 0,
                base + lay.partial, base + lay.scal, base + lay.flags, base + lay.master,
                base + lay.lowp, ctr[r].data_ptr(), base + lay.stats,
            ])
        wave_views.append(torch.tensor(rows, dtype=torch.int64).to(dev))
    if stats is not None:
        # emulation runs all ranks from one launch: each emulated rank reads its own row through
        # the per-rank pointer, so the rows must be contiguous and the kernel gets row 0's address
        stats = stats.to(torch.float32).contiguous()
    max_blocks = int(lib.nfb_allreduce_max_blocks(native._dt(slots), kw))
    grid = max(2, min(max_blocks // g, 64))
    chunk = chunk_elems if chunk_elems else chunk_elems_for(n)
    for call in range(1, calls + 1):
        for w in range(waves):
            flags = (1 if w > 0 else 0) | (2 if w < waves - 1 else 0)
            code = lib.nfb_fedavg_allreduce(
                wave_views[w].data_ptr(), 0, g, g, native._dt(slots), kw, n, slots.stride(1),
                float(clip_norm), float(noise_std),
                noise_mask.data_ptr() if noise_mask is not None else None,
                seed & (2**64 - 1), offset, call, grid // 2, chunk, grid, None, None, None,
                status.data_ptr(), spin_ns, flags,
                stats.data_ptr() if stats is not None else None, stats_len,
                C.byref(target) if (target is not None and w == waves - 1) else None, None,
                torch.cuda.current_stream(dev).cuda_stream,
            )
            native._check(code, "fedavg_allreduce(emulated)")
            native._count()
    torch.cuda.synchronize(dev)
    if int(status.item()) != 0:
        raise RuntimeError("emulated fedavg_allreduce: barrier timed out")
    master = torch.stack([bufs[r, lay.master : lay.master + 4 * n].view(torch.float32)
                          for r in range(g)])
    lowp = torch.stack([bufs[r, lay.lowp : lay.lowp + 2 * n].view(torch.bfloat16)
                        for r in range(g)])
    if return_stats:
        got = torch.stack([bufs[r, lay.stats : lay.stats + 4 * g * stats_len].view(torch.float32)
                           .view(g, stats_len) for r in range(g)])
        return master, lowp, got
    return master, lowp

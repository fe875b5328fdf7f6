"""ResNet-18 (CIFAR shape, BatchNorm) for a whole cohort of clients on hand-written kernels.

Same math as ``models.resnet.ResNet18(norm="batch")`` trained by the reference's client loop
(``nanofed/trainer/base.py:134-160`` with ``torch.optim.SGD``), expressed for all C clients of a
GPU at once: the client is the group dimension of every launch, a local step is ~125 launches
whatever the client count, and a local epoch is one CUDA graph.  No autograd, no cuDNN, no cuBLAS.

    forward   im2col(3x3x3 stem) → [conv (tcgen05, BN batch sums in the epilogue) → BN+ReLU
              (+ residual | + BN(projection))]* → avg-pool + fc + cross-entropy (one kernel)
    backward  per BatchNorm: reduce (dγ, dβ) + apply; every ReLU backward is a mask in the epilogue
              of the kernel that PRODUCES the gradient (head, dgrad), never a separate operand;
              per conv: wgrad (tcgen05, both operands MN-major, tap-major fp32 result) and dgrad
              (tcgen05; stride 2 = four parity-class launches, the 1×1 projection's gradient rides
              in the k-loop of class (0,0); identity shortcuts are added in the dgrad epilogue)
    update    conv weights live in a tap-major fp32 master + bf16 working copy during local
              training (converted from / to the arena's torch layout once per round); where the
              weight gradient is complete in one CTA (94 % of the parameters) SGD is applied in the
              wgrad epilogue and the gradient is never written; split-K layers, BatchNorm and the
              classifier take two small elementwise launches

Activations are bf16 NHWC, accumulation / BatchNorm statistics / master weights fp32.  BatchNorm
running statistics and ``num_batches_tracked`` are updated in the client's arena row, so FedAvg
averages them exactly like the reference does for every ``state_dict`` entry
(``nanofed/server/aggregator/fedavg.py:53-66``).
"""

from __future__ import annotations

import os
import struct
from dataclasses import dataclass, field
from typing import Any

import torch

from nanofed_b200 import ops
from nanofed_b200.ops import native, rn, tc

AVAILABLE = True
_WIDTHS, _STRIDES = (64, 128, 256, 512), (1, 2, 2, 2)


@dataclass
class _Conv:
    name: str          # state_dict key of the weight
    cin: int           # channels of the A tensor (64 patch entries for the stem)
    cout: int
    k: int
    stride: int
    hin: int
    w_off: int = 0     # fp32 master inside a client row
    t_off: int = 0     # tap-major copy / gradient inside a wt / dwt row
    stem: bool = False

    @property
    def taps(self) -> int:
        return self.k * self.k

    @property
    def hout(self) -> int:
        return self.hin // self.stride


@dataclass
class _Bn:
    prefix: str
    c: int
    offs: tuple[int, ...] = ()       # weight, bias, running_mean, running_var, num_batches_tracked
    s_off: int = 0                   # batch sums [clients][2][c] inside the stats buffer
    g_off: tuple[int, int] = (0, 0)  # dgamma, dbeta inside a sgrad row


@dataclass
class _Block:
    conv1: _Conv
    bn1: _Bn
    conv2: _Conv
    bn2: _Bn
    conv_sc: _Conv | None = None
    bn_sc: _Bn | None = None
    bufs: dict[str, torch.Tensor] = field(default_factory=dict)


class _Rows:
    """Clients ``[lo, hi)`` of an arena (or of the resident shards): row views, nothing copied."""

    def __init__(self, arena: Any, lo: int, hi: int) -> None:
        self.device, self.layout, self.num_clients = arena.device, arena.layout, hi - lo
        self.client0 = lo
        self.slots, self.stats = arena.slots[lo:hi], arena.stats[lo:hi]
        self.momentum = None if arena.momentum is None else arena.momentum[lo:hi]


class _ShardRows:
    def __init__(self, shards: Any, lo: int, hi: int) -> None:
        self.images, self.labels = shards.images[lo:hi], shards.labels[lo:hi]


class _Lane:
    """Lock-step local training of ``arena.num_clients`` ResNet-18s on the current stream."""

    def __init__(self, cfg: Any, arena: Any, shards: Any, client_base: int = 0,
                 clients_total: int | None = None) -> None:
        assert arena.device.type == "cuda", "ResNetCohortNet needs a CUDA device"
        self.cfg, self.arena, self.shards = cfg, arena, shards
        self.device = dev = arena.device
        g, b = arena.num_clients, cfg.batch_size
        if b % 8:
            raise ValueError("ResNetCohortNet: batch_size must be a multiple of 8")
        self.g, self.b = g, b
        lay = arena.layout
        self.n_total = lay.total
        off = {name: lay[name].offset for name in lay.names}
        # "batch": BatchNorm2d (statistics per client and channel, from the conv epilogue);
        # "group": GroupNorm(8) (statistics per image and group, csrc/rn_gn.cu) - model resnet18_gn
        self.norm = "batch" if "bn1.running_mean" in off else "group"

        def bn(prefix: str, c: int) -> _Bn:
            return _Bn(prefix, c, tuple(off.get(f"{prefix}.{n}", -1) for n in
                                        ("weight", "bias", "running_mean", "running_var",
                                         "num_batches_tracked")))

        # ---- layer table ------------------------------------------------------------------
        self.stem = _Conv("conv1.weight", 64, 64, 1, 1, 32, stem=True)
        self.stem_bn = bn("bn1", 64)
        self.blocks: list[_Block] = []
        cin, h = 64, 32
        for li, (width, stride) in enumerate(zip(_WIDTHS, _STRIDES)):
            for bi in range(2):
                p = f"layer{li + 1}.{bi}"
                s = stride if bi == 0 else 1
                blk = _Block(_Conv(f"{p}.conv1.weight", cin, width, 3, s, h), bn(f"{p}.bn1", width),
                             _Conv(f"{p}.conv2.weight", width, width, 3, 1, h // s),
                             bn(f"{p}.bn2", width))
                if s != 1 or cin != width:
                    blk.conv_sc = _Conv(f"{p}.shortcut.0.weight", cin, width, 1, s, h)
                    blk.bn_sc = bn(f"{p}.shortcut.1", width)
                self.blocks.append(blk)
                cin, h = width, h // s
        self.convs = [self.stem] + [c for blk in self.blocks
                                    for c in (blk.conv1, blk.conv2, blk.conv_sc) if c is not None]
        self.bns = [self.stem_bn] + [n for blk in self.blocks
                                     for n in (blk.bn1, blk.bn2, blk.bn_sc) if n is not None]
        t = 0
        for c in self.convs:
            c.w_off, c.t_off = off[c.name], t
            t += c.taps * c.cout * c.cin
        self.t_stride = t
        s_cursor = g_cursor = 0
        small: list[tuple[int, int, int]] = []           # (p_off, g_off, len)
        for n in self.bns:
            n.s_off = s_cursor
            s_cursor += g * 2 * n.c if self.norm == "batch" else g * b * 16
            n.g_off = (g_cursor, g_cursor + n.c)
            small += [(n.offs[0], g_cursor, n.c), (n.offs[1], g_cursor + n.c, n.c)]
            g_cursor += 2 * n.c
        self.fc_w, self.fc_b = off["fc.weight"], off["fc.bias"]
        self.fc_gw, self.fc_gb = g_cursor, g_cursor + 10 * 512
        small += [(self.fc_w, self.fc_gw, 10 * 512), (self.fc_b, self.fc_gb, 10)]
        g_cursor += 10 * 512 + 64
        self.n_small = len(small)

        # ---- device state -----------------------------------------------------------------
        f32 = dict(dtype=torch.float32, device=dev)
        bf = dict(dtype=torch.bfloat16, device=dev)
        self.wt = torch.zeros(g, self.t_stride, **bf)       # tap-major bf16 weights
        self.wm = torch.zeros(g, self.t_stride, **f32)      # tap-major fp32 master weights
        self.dwt = torch.zeros(g, self.t_stride, **f32)     # tap-major fp32 gradients (split-K layers)
        self.mom = torch.zeros(g, self.t_stride, **f32) if cfg.momentum != 0.0 else None
        self.sgrad = torch.zeros(g, g_cursor, **f32)        # BatchNorm / classifier gradients
        # BatchNorm batch sums: [0, s_cursor) forward (sum x, sum x^2), [s_cursor, 2 s_cursor)
        # backward (sum g, sum g x) left by the dgrad epilogues; zeroed together after every step
        self.stats = torch.zeros(2 * s_cursor, **f32)
        self._s_half = s_cursor
        cols, rows = 0, []
        for c in self.convs:
            rows.append(struct.pack("<qqqiiii", c.w_off, c.t_off, cols, c.cout, c.cin, c.taps,
                                    1 if c.stem else 0))
            cols += c.cout * c.cin
        self.total_cols = cols
        self.conv_table = torch.frombuffer(bytearray(b"".join(rows)), dtype=torch.uint8).to(dev)
        self.small_table = torch.tensor(small, dtype=torch.int64, device=dev)

        def act(h: int, c: int) -> torch.Tensor:
            return torch.zeros(g * b, h, h, c, **bf)

        self.xcol, self.raw0, self.a0 = act(32, 64), act(32, 64), act(32, 64)
        for blk in self.blocks:
            ho, w = blk.conv1.hout, blk.conv1.cout
            blk.bufs = {"raw1": act(ho, w), "a": act(ho, w), "raw2": act(ho, w), "out": act(ho, w)}
            if blk.conv_sc is not None:
                blk.bufs["rawsc"] = act(ho, w)
        big = g * b * 32 * 32 * 64
        self._scratch = [torch.zeros(big, **bf) for _ in range(5)]   # D0, D1, S1, S2, S3
        self.sms = torch.cuda.get_device_properties(dev).multi_processor_count
        # which convolutions get their SGD step inside the wgrad epilogue (complete gradient in one
        # CTA, momentum-free); the others go through dwt + tap_sgd
        # client-side DP-SGD (CohortConfig.local_privacy; reference nanofed/trainer/private.py:54-86):
        # the clip factor needs the norm of the COMPLETE gradient, so no update rides in a wgrad
        # epilogue; K-STATS over the two gradient buffers, clip + Philox noise inside tap_sgd /
        # small_sgd.  dp_steps counts the local steps of earlier rounds (advanced once per round,
        # outside the graph) so a replayed graph keeps drawing fresh noise.
        self.local_dp = getattr(cfg, "local_privacy", None)
        seed = getattr(cfg, "dp_seed", None)
        self.dp_seed = int(seed) if seed is not None else int.from_bytes(os.urandom(8), "little") >> 1
        self.dp_steps = torch.zeros(1, dtype=torch.int64, device=dev)
        self.dp_client0 = int(getattr(arena, "client0", 0)) + int(client_base)
        self.dp_clients_total = int(clients_total) if clients_total else self.dp_client0 + g
        self.fuse_sgd = (cfg.momentum == 0.0 and self.local_dp is None
                         and os.environ.get("NFB_RN_NO_FUSED_SGD") is None)
        self._ksplit = {c.name: self._pick_ksplit(c) for c in self.convs}
        # tap-major ranges updated by tap_sgd, cut into pieces of 8192 elements: one (piece, client)
        # per block, so a 295 k-element layer and a 4 k-element one do not share one grid shape
        ranges = []
        for c in self.convs:
            if self.fuse_sgd and self._ksplit[c.name] == 1:
                continue
            n = c.taps * c.cout * c.cin
            ranges += [(c.t_off + lo, min(8192, n - lo)) for lo in range(0, n, 8192)]
        self.tap_ranges = torch.tensor(ranges or [(0, 0)], dtype=torch.int64, device=dev)
        self.tap_max = max((r[1] for r in ranges), default=0)
        # BatchNorm backward sums from the dgrad epilogue instead of a reduction pass
        self.fuse_bn_sums = (self.norm == "batch"
                             and os.environ.get("NFB_RN_NO_FUSED_BN_SUMS") is None)
        # GroupNorm statistics from the forward conv epilogue instead of a pass over the output
        self.fuse_gn_stats = os.environ.get("NFB_RN_GN_STATS_KERNEL") is None
        self._tile_override = os.environ.get("NFB_RN_TILE")   # "bn,halves" for experiments
        self._wgrad_posk = int(os.environ.get("NFB_RN_WGRAD_K", "0"))   # 0 = per-layer choice

    # ---- helpers ---------------------------------------------------------------------------
    def _view(self, i: int, h: int, c: int) -> torch.Tensor:
        n = self.g * self.b * h * h * c
        return self._scratch[i][:n].view(self.g * self.b, h, h, c)

    def _wref(self, c: _Conv) -> rn.WeightRef:
        return rn.WeightRef(self.wt, c.t_off, c.taps * c.cout, c.cin)

    def _bn_stats(self, n: _Bn) -> torch.Tensor:
        if self.norm == "group":      # [images][8 groups][sum, sum of squares]
            return self.stats[n.s_off : n.s_off + self.g * self.b * 16].view(self.g * self.b, 8, 2)
        return self.stats[n.s_off : n.s_off + self.g * 2 * n.c].view(self.g, 2, n.c)

    def _bn_bsum(self, n: _Bn) -> torch.Tensor:
        lo = self._s_half + n.s_off
        if self.norm == "group":      # [images][8 groups][sum gamma g, sum gamma g xhat]
            return self.stats[lo : lo + self.g * self.b * 16].view(self.g * self.b, 8, 2)
        return self.stats[lo : lo + self.g * 2 * n.c].view(self.g, 2, n.c)

    def _tile(self, positions: int, n_out: int) -> tuple[int, int]:
        if self._tile_override:
            bn, halves = (int(v) for v in self._tile_override.split(","))
            if n_out % bn == 0 and positions % (128 * halves) == 0:
                return bn, halves
        return rn.pick_tile(positions, n_out, self.g, self.sms)

    def _fwd_conv(self, c: _Conv, x: torch.Tensor, raw: torch.Tensor, n: _Bn) -> None:
        batch = self.norm == "batch"
        rn.conv_gemm(False, x, self._wref(c), raw, clients=self.g, grid_hw=(c.hout, c.hout),
                     taps=rn.fwd_taps(c.k, c.stride, c.cout), a_parity=c.stride == 2,
                     stats=self._bn_stats(n) if batch else None, n_out=c.cout,
                     gn_stats=self._bn_stats(n) if not batch and self.fuse_gn_stats else None,
                     tile=self._tile(self.b * c.hout * c.hout, c.cout))
        if not batch and not self.fuse_gn_stats:
            rn.gn_stats(raw, self._bn_stats(n))

    def _norm_apply(self, n: _Bn, raw: torch.Tensor, y: torch.Tensor, *, res=None,
                    res_n: "_Bn | None" = None) -> None:
        """y = relu(norm(raw) [+ res | + norm_sc(res)])."""
        fn = rn.bn_apply if self.norm == "batch" else rn.gn_apply
        fn(raw, y, self.arena.slots, self._bn_stats(n), n.offs, clients=self.g, relu=True, res=res,
           res_stats=self._bn_stats(res_n) if res_n is not None else None,
           res_offs=res_n.offs if res_n is not None else None)

    def _norm_bwd(self, n: _Bn, dout: torch.Tensor, x: torch.Tensor, dx: torch.Tensor, *,
                  n2: "_Bn | None" = None, x2=None, dx2=None, have_sums: bool = False) -> None:
        """Normalisation backward: parameter gradients into ``sgrad``, dx (and dx2 for the
        projection shortcut's layer, which receives the same ``dout``).  ``have_sums``: the dgrad
        that produced ``dout`` already left the BatchNorm sums in ``_bn_bsum(n)``."""
        slots = self.arena.slots
        offs = n.g_off + (n2.g_off if n2 is not None else (0, 0))
        if self.norm == "group":
            kw = dict(clients=self.g, x2=x2, stats2=self._bn_stats(n2) if n2 is not None else None,
                      gn2_offs=n2.offs if n2 is not None else None,
                      bsums2=self._bn_bsum(n2) if n2 is not None else None)
            args = (dout, x, slots, self.sgrad, offs, self._bn_stats(n), n.offs, self._bn_bsum(n))
            rn.gn_bwd(False, *args, **kw)
            rn.gn_bwd(True, *args, dx=dx, dx2=dx2, **kw)
            return
        kw = dict(clients=self.g, x2=x2, stats2=self._bn_stats(n2) if n2 is not None else None,
                  bn2_offs=n2.offs if n2 is not None else None)
        if not have_sums:
            rn.bn_bwd(False, dout, None, x, slots, self.sgrad, offs, self._bn_stats(n), n.offs, **kw)
        rn.bn_bwd(True, dout, None, x, slots, self.sgrad, offs, self._bn_stats(n), n.offs, dx=dx,
                  dx2=dx2, raw_sums=self._bn_bsum(n) if have_sums else None, **kw)

    def _pick_ksplit(self, c: _Conv) -> int:
        bn = min(c.cout, 256)
        kb_total = self.b * c.hout * c.hout // 128
        tiles = (c.taps * (c.cin // 64) + 1) // 2 * (c.cout // bn) * self.g
        if tiles * 2 > self.sms:          # at least half a wave without splitting: keep K whole
            return 1
        return max(1, min(kb_total // 4 if kb_total >= 4 else 1, self.sms // tiles))

    def _posk(self, c: _Conv) -> int:
        """Positions per k-block of the wgrad kernel.  64 keeps two CTAs on an SM, which pays where
        the k-loop is short and the epilogue (fused SGD over 128 x 256 weights) is not: the
        512-channel layers (246 vs 339 us for the three 512->512 convs); elsewhere 128 wins."""
        if self._wgrad_posk:
            return self._wgrad_posk
        return 64 if c.cout >= 512 else 128

    def _wgrad(self, c: _Conv, x: torch.Tensor, dy: torch.Tensor) -> None:
        """Weight gradient of ``c`` — call AFTER the data gradient that reads the old weights."""
        k_split = self._ksplit[c.name]
        cfg = self.cfg
        sgd = ((self.wm, self.wt, cfg.learning_rate, cfg.weight_decay)
               if self.fuse_sgd and k_split == 1 else None)
        rn.conv_wgrad(x, dy, self.dwt, c.t_off, clients=self.g,
                      taps=rn.fwd_taps(c.k, c.stride, c.cout), x_parity=c.stride == 2,
                      k_split=k_split, tile=(min(c.cout, 256), self._posk(c)), sgd=sgd)

    def _dgrad(self, c: _Conv, dy: torch.Tensor, dx: torch.Tensor, *, relu_of: torch.Tensor,
               residual=None, sc: _Conv | None = None, dy_sc: torch.Tensor | None = None,
               bn: "tuple[_Bn, torch.Tensor] | None" = None) -> None:
        """Gradient w.r.t. the PRE-ReLU value of the conv's input: ``relu_of`` is that (activated)
        input, its ReLU mask is applied in the epilogue; ``residual`` (an already masked gradient
        arriving over an identity shortcut) is added before the mask.  ``bn = (layer, raw)``: the
        input is ``relu(bn(raw) [+ shortcut])`` - the epilogue also leaves that BatchNorm's backward
        sums (sum g, sum g * raw) in ``_bn_bsum(layer)``, so no reduction pass over dx is needed."""
        kw = dict(stats=self._bn_bsum(bn[0]), bn_x=bn[1]) if bn is not None else {}
        if c.stride == 1:
            rn.conv_gemm(True, dy, self._wref(c), dx, clients=self.g, grid_hw=(c.hin, c.hin),
                         taps=rn.dgrad_taps(c.k, 1, c.cout), residual=residual, out_mask=relu_of,
                         n_out=c.cin, tile=self._tile(self.b * c.hin * c.hin, c.cin), **kw)
            return
        assert residual is None
        # the four output parity classes in ONE launch (grid.z = class x client); the 1x1 projection
        # shortcut's gradient rides in the k-loop of class (0, 0)
        rn.conv_gemm(True, dy, self._wref(c), dx, clients=self.g, grid_hw=(c.hout, c.hout),
                     taps=rn.dgrad_taps_all_classes(c.k, c.cout, sc.cout if sc is not None else None),
                     out_parity=(0, 0), all_classes=True, a1=dy_sc, w1=self._wref(sc) if sc else None,
                     out_mask=relu_of, n_out=c.cin,
                     tile=self._tile(self.b * c.hout * c.hout, c.cin), **kw)

    # ---- hooks used by FedSimulation ---------------------------------------------------------
    def after_broadcast(self) -> None:
        """Fresh global weights arrived in the client rows: rebuild the bf16 tap-major copies."""
        rn.wm_convert(self.conv_table, len(self.convs), self.total_cols, self.arena.slots, self.wm,
                      self.wt, to_slots=False)
        if self.mom is not None:
            self.mom.zero_()             # fresh optimizer per round, like K-BCAST does for the arena
        if self.local_dp is not None:
            self.dp_steps += self.cfg.local_steps      # new noise for this round's replay

    def before_aggregate(self) -> None:
        """Trained conv weights back into the client rows (torch layout) that K-AGG reads."""
        rn.wm_convert(self.conv_table, len(self.convs), self.total_cols, self.arena.slots, self.wm,
                      self.wt, to_slots=True)

    # ---- one local step for every client -------------------------------------------------------
    def step(self, step: int) -> None:
        cfg, a, g, b = self.cfg, self.arena, self.g, self.b
        lo = (step % cfg.steps_per_epoch) * b
        x = self.shards.images[:, lo : lo + b]
        y = self.shards.labels[:, lo : lo + b]
        slots = a.slots
        # ---------------- forward ----------------
        rn.im2col3(x, self.xcol)
        self._fwd_conv(self.stem, self.xcol, self.raw0, self.stem_bn)
        self._norm_apply(self.stem_bn, self.raw0, self.a0)
        cur = self.a0
        for blk in self.blocks:
            u = blk.bufs
            self._fwd_conv(blk.conv1, cur, u["raw1"], blk.bn1)
            self._norm_apply(blk.bn1, u["raw1"], u["a"])
            self._fwd_conv(blk.conv2, u["a"], u["raw2"], blk.bn2)
            if blk.conv_sc is not None:
                self._fwd_conv(blk.conv_sc, cur, u["rawsc"], blk.bn_sc)
                self._norm_apply(blk.bn2, u["raw2"], u["out"], res=u["rawsc"], res_n=blk.bn_sc)
            else:
                self._norm_apply(blk.bn2, u["raw2"], u["out"], res=cur)
            u["in"] = cur
            cur = u["out"]
        # ---------------- head: loss, metrics, gradient of the last feature map ----------------
        d0, d1 = 0, 1
        dout = self._view(d0, 4, 512)
        rn.head(cur, y, slots, self.fc_w, self.fc_b, self.sgrad, self.fc_gw, self.fc_gb, dout,
                a.stats, clients=g)
        # ---------------- backward ----------------
        fuse = self.fuse_bn_sums
        have_sums = False          # did dout's producer leave this block's bn2 backward sums?
        for bi in range(len(self.blocks) - 1, -1, -1):
            blk = self.blocks[bi]
            u = blk.bufs
            ho, w, hin, cin = blk.conv1.hout, blk.conv1.cout, blk.conv1.hin, blk.conv1.cin
            dout = self._view(d0, ho, w)
            dc2 = self._view(2, ho, w)
            proj = blk.conv_sc is not None
            dcs = self._view(3, ho, w) if proj else None
            # dout is already the gradient w.r.t. the pre-ReLU block output (its producer - the
            # head or the next block's dgrad epilogue - applied the mask): no mask operand here
            self._norm_bwd(blk.bn2, dout, u["raw2"], dc2, n2=blk.bn_sc if proj else None,
                           x2=u["rawsc"] if proj else None, dx2=dcs, have_sums=have_sums)
            da = self._view(4, ho, w)
            # reads the old weights: before the wgrad
            self._dgrad(blk.conv2, dc2, da, relu_of=u["a"],
                        bn=(blk.bn1, u["raw1"]) if fuse else None)
            self._wgrad(blk.conv2, u["a"], dc2)
            dc1 = dc2                                     # dc2 is dead after the two launches above
            self._norm_bwd(blk.bn1, da, u["raw1"], dc1, have_sums=fuse)
            dx = self._view(d1, hin, cin)
            # dx is the gradient at the producer of this block's input: the stem BatchNorm or the
            # previous block's bn2 (+ shortcut).  A projection block there needs sum(g * x) for two
            # layers - it keeps its own reduction pass.
            prev = self.blocks[bi - 1] if bi else None
            nxt = None
            if fuse and prev is None:
                nxt = (self.stem_bn, self.raw0)
            elif fuse and prev.conv_sc is None:
                nxt = (prev.bn2, prev.bufs["raw2"])
            if proj:
                self._dgrad(blk.conv1, dc1, dx, relu_of=u["in"], sc=blk.conv_sc, dy_sc=dcs, bn=nxt)
                self._wgrad(blk.conv_sc, u["in"], dcs)
            else:
                self._dgrad(blk.conv1, dc1, dx, relu_of=u["in"], residual=dout, bn=nxt)
            self._wgrad(blk.conv1, u["in"], dc1)
            have_sums = nxt is not None
            d0, d1 = d1, d0
        dout = self._view(d0, 32, 64)
        dc0 = self._view(2, 32, 64)
        self._norm_bwd(self.stem_bn, dout, self.raw0, dc0, have_sums=have_sums)
        self._wgrad(self.stem, self.xcol, dc0)
        # ---------------- update ----------------
        dp = None
        if self.local_dp is not None:
            ss = ops.update_stats(self.dwt)[0] + ops.update_stats(self.sgrad)[0]
            c_norm = float(self.local_dp.max_gradient_norm)
            dp = dict(sumsq=ss, clip_norm=c_norm,
                      noise_std=float(self.local_dp.noise_multiplier) * c_norm, seed=self.dp_seed,
                      step_base=self.dp_steps, step=step, client0=self.dp_client0,
                      clients_total=self.dp_clients_total,
                      row_groups=(self.t_stride + self.sgrad.shape[1] + 3) // 4)
        if self.tap_max:
            rn.tap_sgd(self.tap_ranges, self.tap_max, self.wm, self.dwt, self.mom, self.wt,
                       lr=cfg.learning_rate, mu=cfg.momentum, wd=cfg.weight_decay, dp=dp)
        rn.small_sgd(self.small_table, self.n_small, slots, a.momentum, self.sgrad,
                     lr=cfg.learning_rate, mu=cfg.momentum, wd=cfg.weight_decay, zero=self.stats,
                     dp=dp, dp_idx_base=self.t_stride)



class ResNetCohortNet:
    """Local training of ``arena.num_clients`` ResNet-18s on one GPU: the clients are dealt to
    ``NFB_RN_STREAMS`` lanes, each a lock-step sub-cohort on its own CUDA stream; the whole local
    epoch of all lanes is one CUDA graph with one fork and one join.

    Default: ONE lane.  Within a lane every kernel depends on the one before it, so the HBM-bound
    BatchNorm passes and the tails of the persistent convolution kernels (512 tiles on 148 SMs =
    3.46 waves) leave the tensor pipe idle, and two independent lanes should fill those holes with
    each other's work.  Measured (8 clients x batch 64, one B200): 1 lane 142.6 k samples/s,
    2 lanes 140.3 k, 4 lanes 129.2 k.  The persistent convolution CTA owns its SM outright (384
    threads x 168 registers = the whole register file, 225 KB of shared memory), so no BatchNorm
    block can run next to it: the lanes only alternate, and every half-size launch pays its own
    prologue and its own last wave."""

    def __init__(self, cfg: Any, arena: Any, shards: Any, client_base: int = 0,
                 clients_total: int | None = None) -> None:
        """``client_base`` / ``clients_total``: position of this GPU's clients in the whole job
        (distinct DP noise streams per client across ranks that share a seed)."""
        self.cfg, self.arena, self.shards = cfg, arena, shards
        self.device = arena.device
        g = arena.num_clients
        clients_total = clients_total or client_base + g
        n = max(1, int(os.environ.get("NFB_RN_STREAMS", "1")))
        while g % n:
            n -= 1
        per = g // n
        self.lanes = [_Lane(cfg, _Rows(arena, k * per, (k + 1) * per),
                            _ShardRows(shards, k * per, (k + 1) * per), client_base,
                            clients_total) for k in range(n)]
        self.streams = [torch.cuda.Stream(device=self.device) for _ in self.lanes[1:]]
        self._graph: "torch.cuda.CUDAGraph | None" = None
        self.launches_per_round = 0
        self.fuse_bn_sums = self.lanes[0].fuse_bn_sums

    def after_broadcast(self) -> None:
        for lane in self.lanes:
            lane.after_broadcast()

    def before_aggregate(self) -> None:
        for lane in self.lanes:
            lane.before_aggregate()

    def _run(self, steps: "range | list[int]") -> None:
        """``steps`` of every lane: lane 0 on the current stream, the others forked from it and
        joined back (also what stream capture records)."""
        main = torch.cuda.current_stream(self.device)
        if not self.streams:
            for s in steps:
                self.lanes[0].step(s)
            return
        fork = torch.cuda.Event()
        fork.record(main)
        for st in self.streams:
            st.wait_event(fork)
        for s in steps:      # interleaved issue order: eager launches of the lanes overlap too
            self.lanes[0].step(s)
            for lane, st in zip(self.lanes[1:], self.streams):
                with torch.cuda.stream(st):
                    lane.step(s)
        for st in self.streams:
            done = torch.cuda.Event()
            done.record(st)
            main.wait_event(done)

    def step(self, step: int) -> None:
        self._run([step])

    def _all_steps(self) -> None:
        self._run(range(self.cfg.local_steps))

    def local_steps(self) -> None:
        if not self.cfg.use_cuda_graph:
            self._all_steps()
            return
        if self._graph is None:
            self._capture()
        self._graph.replay()

    def _capture(self) -> None:
        a = self.arena
        state = [a.slots, a.stats] + ([a.momentum] if a.momentum is not None else [])
        for lane in self.lanes:
            state += [lane.wt, lane.wm] + ([lane.mom] if lane.mom is not None else [])
        keep = [t.clone() for t in state]

        def restore() -> None:
            for t, k in zip(state, keep):
                t.copy_(k)

        side = torch.cuda.Stream(device=self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):
            self.step(0)      # eager warm-up: tensor maps, function attributes
        torch.cuda.current_stream(self.device).wait_stream(side)
        torch.cuda.synchronize(self.device)
        tc.check_status(self.device)
        restore()
        torch.cuda.synchronize(self.device)
        before = native.LAUNCHES
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=side):
            self._all_steps()
        self.launches_per_round = native.LAUNCHES - before
        restore()
        torch.cuda.synchronize(self.device)
        self._graph = graph

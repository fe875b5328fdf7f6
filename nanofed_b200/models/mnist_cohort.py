"""Flagship engine: the MNIST CNN for a whole cohort of clients, on hand-written kernels.

Same math as ``MNISTModel`` (reference ``nanofed/models/mnist.py:6-28``) and the same
training step as the reference's client loop (``trainer/base.py:134-160`` with
``torch.optim.SGD``, ``examples/mnist/run_experiment.py:70-78``), but expressed
for C clients at once — the client index is the *group* dimension of every launch,
so a local step is 13 launches no matter how many clients share the GPU:

    prep        conv2 weights fp32 [co][ci][3][3] → bf16 [co][tap·32+ci]            (K-ELT)
    conv1       direct 3×3, 1→32, +bias +ReLU → NHWC bf16 on a 26×26 grid             (K-CONV, CUDA cores: K=9)
    conv2+pool  implicit GEMM (9 taps = 9 descriptors into one TMA slab) → tcgen05 → TMEM,
                epilogue: +bias, ReLU, 2×2 max-pool, dropout(.25), 1 B/elem argmax code   (K-CONV)
    fc1         grouped tcgen05 GEMM, split-K, fp32 accumulate                        (K-GEMM)
    head        +bias ReLU dropout(.5) → fc2 → log-softmax/NLL/accuracy → dlogits
                → fc2 grads → dh → fc1 bias grad; device-side loss/acc counters      (K-XENT)
    fc1 wgrad   tcgen05 GEMM, both operands MN-major                                  (K-GEMM)
    fc1 dgrad   tcgen05 GEMM, A MN-major                                              (K-GEMM)
    pool bwd    scatter through argmax/ReLU/dropout → grid gradient + conv2 bias grad (K-ELT)
    conv2 wgrad 9 TMEM accumulators sharing one A tile, fp32 atomics into OIHW        (K-CONV)
    conv2 dgrad implicit GEMM with negative shifts, ReLU mask in the epilogue         (K-CONV)
    conv1 wgrad direct reduction + bias grad                                          (K-CONV)
    K-OPT       fused SGD over the whole [C·N] slab, bf16 working copy written back

No autograd, no cuDNN/cuBLAS, no host sync; the local epoch is captured in a CUDA
graph.  Activations are bf16, accumulation and master weights fp32.
"""

from __future__ import annotations

import ctypes as C
import os
from typing import Any

import torch

from nanofed_b200 import ops
from nanofed_b200.ops import native, tc

AVAILABLE = True
_GRID = 26 * 26

_declared = False


def _lib() -> C.CDLL:
    global _declared
    lib = native.lib()
    if not _declared:
        vp, i32, i64, f32, u64 = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_uint64
        sigs = {
            "nfb_mnist_prep_conv2": [vp, i64, i64, i32, vp, vp],
            "nfb_mnist_conv1_fwd": [vp, i64, vp, i64, i64, i64, i32, i32, vp, vp],
            "nfb_mnist_pool_fwd": [vp, i32, i32, f32, vp, u64, vp, vp, vp],
            "nfb_mnist_pool_bwd": [vp, vp, i32, i32, f32, vp, vp, i64, i64, vp],
            "nfb_mnist_fc_head": [vp, vp, i64, i32, i32, vp, vp, i64, i64, i64, i64, f32, vp,
                                  i32, u64, vp, vp, vp],
            "nfb_mnist_conv1_wgrad": [vp, vp, i64, i32, i32, vp, i64, i64, i64, vp],
        }
        for name, args in sigs.items():
            fn = getattr(lib, name)
            fn.restype = i32
            fn.argtypes = args
        _declared = True
    return lib


class MnistCohortNet:
    """Lock-step local training of ``arena.num_clients`` MNIST CNNs on one GPU."""

    def __init__(self, cfg: Any, arena: Any, shards: Any, dropout: bool = True) -> None:
        assert arena.device.type == "cuda", "MnistCohortNet needs a CUDA device"
        assert arena.slots_lowp is not None, "arena must be created with with_lowp=True"
        self.cfg, self.arena, self.shards = cfg, arena, shards
        self.lib = _lib()
        dev = arena.device
        self.device = dev
        c_all, b = arena.num_clients, cfg.batch_size
        # Clients are processed in groups of `group` per launch, forward AND backward of a group
        # before the next one, reusing ONE set of activation buffers: the producer->consumer
        # distance of every activation stays within the 126 MB L2 instead of round-tripping
        # through HBM (activations are 0.53 GB per step for 32 clients).
        group = int(os.environ.get("NFB_COHORT_GROUP", "0")) or getattr(cfg, "cohort_group", 0) or c_all
        group = max(1, min(group, c_all))
        while c_all % group:
            group -= 1
        self.group = group
        c = group
        self.c_all, self.c, self.b = c_all, c, b
        self.rows = b * _GRID
        lay = arena.layout
        self.off = {name: lay[name].offset for name in lay.names}
        self.n_total = lay.total
        bf = dict(dtype=torch.bfloat16, device=dev)
        self.w2p = torch.zeros(c, 64, 288, **bf)
        self.a1 = torch.zeros(c, self.rows, 32, **bf)
        self.p = torch.zeros(c, b, 9216, **bf)
        self.pidx = torch.zeros(c, b, 144, 64, dtype=torch.uint8, device=dev)
        self.hpre = torch.zeros(c, b, 128, dtype=torch.float32, device=dev)
        self.dh = torch.zeros(c, b, 128, **bf)
        self.dp = torch.zeros(c, b, 9216, **bf)
        self.da2 = torch.zeros(c, self.rows, 64, **bf)
        self.da1 = torch.zeros(c, self.rows, 32, **bf)
        self.grads = torch.zeros(c_all, self.n_total, dtype=torch.float32, device=dev)
        # [0] dropout step counter, [1] fc_head's block-arrival counter (always back at zero)
        self.rng_state = torch.zeros(2, dtype=torch.int64, device=dev)
        self.p1 = 0.25 if dropout else 0.0
        self.p2 = 0.5 if dropout else 0.0
        self.seed = int(cfg.seed) * 7919 + 13
        # fc1 weights: bf16 working copy, viewed [C, 128, 9216] inside the arena rows
        o = self.off["fc1.weight"]
        self.fc1_w = arena.slots_lowp[:, o : o + 128 * 9216].unflatten(1, (128, 9216))
        self.fc1_g = self.grads[:, o:]
        self.fc1_master = arena.slots[:, o : o + 128 * 9216].unflatten(1, (128, 9216))
        self._fc1_range = (o, o + 128 * 9216)
        self.fuse_conv1_wgrad = os.environ.get("NFB_NO_FUSED_WGRAD1") is None
        self.tc_conv1 = os.environ.get("NFB_NO_TC_CONV1") is None
        # K-OPT also writes the permuted bf16 conv2 weights (needs one w2p buffer for all clients)
        self.opt_writes_w2p = (self.group == c_all and os.environ.get("NFB_NO_OPT_W2P") is None
                               and getattr(cfg, "optimizer", "sgd") == "sgd")
        # DP-SGD on the client (reference PrivateTrainer semantics: clip the batch gradient to C, add
        # N(0, (sigma C)^2) per parameter): every client's gradient norm comes from K-STATS over the
        # gradient slab, the clip factor and the Philox noise are applied inside K-OPT
        self.local_dp = getattr(cfg, "local_privacy", None)
        self.dp_seed = int(getattr(cfg, "dp_seed", None) if getattr(cfg, "dp_seed", None) is not None
                           else int.from_bytes(os.urandom(8), "little") >> 1)
        # momentum needs the gradient history in a buffer: only plain SGD is fused; DP needs the
        # complete gradient (its norm) before any of it is applied
        # optimizer="adam": the complete gradient is materialised and the fused Adam kernel of K-OPT
        # runs over the arena (the SGD-specific fusions - fc1 update in its GEMM epilogue, permuted
        # conv2 weights out of K-OPT - are off)
        self.adam = getattr(cfg, "optimizer", "sgd") == "adam"
        self.fuse_fc1_sgd = (cfg.momentum == 0.0 and self.local_dp is None and not self.adam
                             and os.environ.get("NFB_NO_FUSED_SGD") is None)
        # DP with the fc1 update still inside its GEMM: the norm of the (never written) fc1 weight
        # gradient comes from a ghost norm - |sum_b dh_b (x) p_b|^2 = <dh dh^T, p p^T> - so the clip
        # factor is known before the wgrad kernel runs and is applied, with the noise, in its epilogue
        # MEASURED: slower than the materialised path (round 6.52 vs 5.83 ms; 3.98 ms without DP) -
        # 38 M Box-Muller normals per step generated by the GEMM's four epilogue warps per CTA take
        # longer than the K-OPT pass they replace (which spreads them over the whole GPU: +37 us).
        # Kept behind NFB_FUSED_DP=1 with its test; the default is the materialised path.
        self.fuse_fc1_dp = (self.local_dp is not None and cfg.momentum == 0.0 and self.group == c_all
                            and os.environ.get("NFB_FUSED_DP") == "1")
        # Central DP clips every client's update by its L2 norm.  The last local step's update
        # kernels (K-OPT and the fused fc1 update) leave each client's sum of squared parameters in
        # arena.sumsq, so the round runs no K-STATS pass over the client rows.
        # Not with non-finite screening (needs the scan anyway) nor with NFB_FUSED_DP's epilogue.
        self.emits_sumsq = (not self.adam and getattr(cfg, "privacy", None) is not None
                            and not getattr(cfg, "screen_nonfinite", False)
                            and not self.fuse_fc1_dp
                            and os.environ.get("NFB_NO_SUMSQ_SIDE_OUTPUT") is None)
        self.gram = torch.zeros(c_all, b, b, dtype=torch.float32, device=dev) if self.fuse_fc1_dp else None
        self.conv2_g = self.grads[:, self.off["conv2.weight"] :]
        # atomic-accumulated gradient entries (conv1.*, conv2.*) sit in front of fc1.weight
        self._atomic_prefix = self.off["fc1.weight"]
        self._conv2_bias = arena.slots[:, self.off["conv2.bias"] : self.off["conv2.bias"] + 64]
        self._graph: "torch.cuda.CUDAGraph | None" = None
        self.launches_per_round = 0
        self.k_split_fc = 8
        self.k_split_wgrad = max(1, min(16, (4 * 148) // max(c, 1)))
        self.k_split_fc = max(1, min(16, 256 // max(c, 1)))

    # -- hooks used by FedSimulation -------------------------------------------------
    def after_broadcast(self) -> None:
        pass

    def before_aggregate(self) -> None:
        pass

    # -- one local step for every client ------------------------------------------------
    def _check(self, code: int, what: str) -> None:
        native._check(code, what)
        native._count()

    def step(self, step: int, last: bool = False) -> None:
        """``last``: the final step of the local epoch - with central DP configured its update
        kernels also leave every client's sum of squared parameters in ``arena.sumsq`` (K-STATS
        folded into K-OPT and the fused fc1 update), so the round needs no extra pass over the rows."""
        cfg, a = self.cfg, self.arena
        emit = last and self.emits_sumsq
        if emit:
            a.sumsq.zero_()
        for g0 in range(0, self.c_all, self.group):
            self._step_group(step, g0, emit)
        self._apply_optimizer(step, emit)

    def _apply_optimizer(self, step: int, emit_sumsq: bool = False) -> None:
        """One fused K-OPT launch over all clients; also hands the atomically accumulated
        gradient entries (conv1.*, conv2.*) back zeroed for the next step."""
        cfg, a = self.cfg, self.arena
        if self.adam:
            if step == 0:                 # fresh optimizer every round
                a.exp_avg.zero_()
                a.exp_avg_sq.zero_()
            ops.fused_adam_(a.slots.view(-1), self.grads.view(-1), a.exp_avg.view(-1),
                            a.exp_avg_sq.view(-1), step + 1, cfg.learning_rate,
                            beta1=cfg.adam_betas[0], beta2=cfg.adam_betas[1], eps=cfg.adam_eps,
                            weight_decay=cfg.weight_decay, lowp_out=a.slots_lowp.view(-1))
            # entries the next step accumulates into with atomics (conv1.*, conv2.*, fc1.bias, fc2.*)
            self.grads[:, : self._atomic_prefix].zero_()
            self.grads[:, self._fc1_range[1] :].zero_()
            return
        dp_kw: dict[str, Any] = {}
        if emit_sumsq:
            dp_kw["sumsq_out"] = a.sumsq
        skip = self._fc1_range if self.fuse_fc1_sgd else None
        if self.local_dp is not None:
            if self.fuse_fc1_dp:
                lo, hi = self._fc1_range
                ss_a, _ = ops.update_stats(self.grads[:, :lo])       # conv1.*, conv2.*
                ss_b, _ = ops.update_stats(self.grads[:, hi:])       # fc1.bias, fc2.*
                tc.linear_fwd(self.p, self.p, self.gram, k_split=self.k_split_fc)   # p p^T per client
                ss = native.batch_ghost_norm(self.gram, self.dh, ss_a, ss_b)
                tc.linear_wgrad_sgd(self.dh, self.p, self.fc1_master, self.fc1_w, cfg.learning_rate,
                                    cfg.weight_decay,
                                    dp=dict(row_sumsq=ss, clip_norm=self.local_dp.max_gradient_norm,
                                            noise_std=self.local_dp.noise_multiplier
                                            * self.local_dp.max_gradient_norm, seed=self.dp_seed,
                                            step_counter=self.rng_state[0:1], row_len=self.n_total,
                                            w_off=lo))
                skip = self._fc1_range
            else:
                ss, _ = ops.update_stats(self.grads)      # per-client squared gradient norm
            dp_kw.update(row_sumsq=ss, clip_norm=float(self.local_dp.max_gradient_norm),
                         noise_std=float(self.local_dp.noise_multiplier
                                         * self.local_dp.max_gradient_norm),
                         noise_mask=a.noise_mask, seed=self.dp_seed,
                         step_counter=self.rng_state[0:1])   # bumped by fc_head once per step
        ops.fused_sgd_(a.slots.view(-1), self.grads.view(-1), cfg.learning_rate,
                       momentum_buf=None if a.momentum is None else a.momentum.view(-1),
                       momentum=cfg.momentum, weight_decay=cfg.weight_decay,
                       first_step=(step == 0), lowp_out=a.slots_lowp.view(-1),
                       row_len=self.n_total, zero_grad_prefix=self._atomic_prefix,
                       skip_range=skip, **dp_kw,
                       zero_grad_from=self._fc1_range[1],   # fc1.bias / fc2.*: fc_head's atomics
                       # the tap-major bf16 conv2 weights of the NEXT step come out of this pass
                       perm_out=(self.w2p.view(self.c_all, -1), self.off["conv2.weight"], 64, 32, 9)
                       if self.opt_writes_w2p else None)

    def _step_group(self, step: int, g0: int, emit_sumsq: bool = False) -> None:
        cfg, a, lib = self.cfg, self.arena, self.lib
        c, b, n = self.c, self.b, self.n_total
        g1 = g0 + c
        st = torch.cuda.current_stream(self.device).cuda_stream
        lo = (step % cfg.steps_per_epoch) * b
        x = self.shards.images[g0:g1, lo : lo + b]
        y = self.shards.labels[g0:g1, lo : lo + b]
        slots, grads = a.slots[g0:g1], self.grads[g0:g1]
        stats = a.stats[g0:g1]
        off = self.off
        xs = x.stride(0)   # x / y are [c, b, ...] slices of [C, S, ...]: client stride stays S*784
        # ---- forward ----
        if step == 0 or not self.opt_writes_w2p:
            # first step after a broadcast; later steps get w2p from K-OPT's permuted side output
            self._check(lib.nfb_mnist_prep_conv2(slots.data_ptr(), n, off["conv2.weight"], c,
                                                 self.w2p.data_ptr(), st), "prep_conv2")
        if self.tc_conv1:
            tc.conv1_fwd(x, slots, off["conv1.weight"], off["conv1.bias"], self.a1, 26)
        else:
            self._check(lib.nfb_mnist_conv1_fwd(x.data_ptr(), xs, slots.data_ptr(), n,
                                                off["conv1.weight"], off["conv1.bias"], c, b,
                                                self.a1.data_ptr(), st), "conv1_fwd")
        # conv2 + bias + ReLU + max-pool + dropout in one kernel: the 24x24x64 conv output
        # (177 MB per step for 32 clients) never exists in memory
        tc.conv3x3_pool_fwd(self.a1, self.w2p, self._conv2_bias[g0:g1], self.p, self.pidx, b, 26,
                            self.p1, self.rng_state, self.seed + 7 * g0)
        # (hpre and the atomic gradient prefix arrive zeroed: fc_head / K-OPT clear them after use)
        tc.linear_fwd(self.fc1_w[g0:g1], self.p, self.hpre, k_split=self.k_split_fc)
        last = 1 if g1 >= self.c_all else 0
        self._check(lib.nfb_mnist_fc_head(self.hpre.data_ptr(), y.data_ptr(), y.stride(0), c, b,
                                          slots.data_ptr(), grads.data_ptr(), n, off["fc1.bias"],
                                          off["fc2.weight"], off["fc2.bias"], self.p2,
                                          self.rng_state.data_ptr(), last,
                                          self.seed + 7 * g0 + 3, self.dh.data_ptr(),
                                          stats.data_ptr(), st), "fc_head")
        # ---- backward ----
        # the dropout scale 1/(1-p1) of the pooled activations is applied to dp in the dgrad
        # epilogue (fp32, before the bf16 rounding); pool_bwd then only masks
        pre = self.p1 > 0.0 and tc.fc_dgrad_is_persistent(128, b, 9216)
        tc.linear_dgrad(self.fc1_w[g0:g1], self.dh, self.dp,
                        scale=1.0 / (1.0 - self.p1) if pre else 1.0)
        if self.fuse_fc1_dp:
            pass       # DP: the fc1 update needs the client's total gradient norm: _apply_optimizer
        elif self.fuse_fc1_sgd:
            # 96% of the parameters are fc1.weight: its gradient is applied inside the GEMM
            # epilogue (after the data gradient above has read the old weights) and never
            # written; K-OPT then skips that range
            tc.linear_wgrad_sgd(self.dh, self.p, self.fc1_master[g0:g1], self.fc1_w[g0:g1],
                                cfg.learning_rate, cfg.weight_decay,
                                sumsq_out=a.sumsq[g0:g1] if emit_sumsq else None)
        else:
            tc.linear_wgrad(self.dh, self.p, self.fc1_g[g0:g1], n)
        self._check(lib.nfb_mnist_pool_bwd(self.dp.data_ptr(), self.pidx.data_ptr(), c, b,
                                           0.0 if pre else self.p1, self.da2.data_ptr(),
                                           grads.data_ptr(), n,
                                           off["conv2.bias"], st), "pool_bwd")
        tc.conv3x3_wgrad_persistent(self.da2, self.a1, self.conv2_g[g0:g1], 26, n)
        if self.fuse_conv1_wgrad:
            # the masked conv1 output gradient goes from the dgrad epilogue straight into a second
            # MMA against the input windows: da1 (88 MB per step) is never written or re-read
            tc.conv3x3_dgrad_wgrad1(self.da2, self.w2p, self.a1, x, grads, off["conv1.weight"],
                                    off["conv1.bias"], 26)
        else:
            tc.conv3x3_dgrad_persistent(self.da2, self.w2p, self.da1, 26, relu_mask=self.a1)
            self._check(lib.nfb_mnist_conv1_wgrad(self.da1.data_ptr(), x.data_ptr(), xs, c, b,
                                                  grads.data_ptr(), n, off["conv1.weight"],
                                                  off["conv1.bias"], st), "conv1_wgrad")

    def _all_steps(self) -> None:
        n = self.cfg.local_steps
        for s in range(n):
            self.step(s, last=s == n - 1)

    def local_steps(self) -> None:
        if not self.cfg.use_cuda_graph:
            self._all_steps()
            return
        if self._graph is None:
            self._capture()
        self._graph.replay()

    def _capture(self) -> None:
        a = self.arena
        keep = (a.slots.clone(), a.slots_lowp.clone(), a.stats.clone(), self.rng_state.clone(),
                None if a.momentum is None else a.momentum.clone())

        def restore() -> None:
            a.slots.copy_(keep[0])
            a.slots_lowp.copy_(keep[1])
            a.stats.copy_(keep[2])
            self.rng_state.copy_(keep[3])
            if keep[4] is not None:
                a.momentum.copy_(keep[4])

        side = torch.cuda.Stream(device=self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):
            self.step(0)  # eager warm-up: tensor maps, function attributes, lazy inits
        torch.cuda.current_stream(self.device).wait_stream(side)
        torch.cuda.synchronize(self.device)
        tc.check_status(self.device)
        restore()
        before = native.LAUNCHES
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=side):
            self._all_steps()
        self.launches_per_round = native.LAUNCHES - before
        restore()
        torch.cuda.synchronize(self.device)
        self._graph = graph

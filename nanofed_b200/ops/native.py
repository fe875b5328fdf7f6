"""ctypes binding of ``libnanofed_b200.so`` (C ABI: ``include/nanofed_b200.h``).

The library is plain ``nvcc -shared`` output — no libtorch dependency — so it
builds in seconds, ships inside the source tree (``nanofed_b200/_native/``), and
any host language can call it.  Tensors cross the boundary as raw device
pointers + sizes; every launch goes onto the *current PyTorch stream* of the
tensor's device, so ops compose with torch work, CUDA graphs capture them, and
``torch.cuda.Event`` timing brackets them.

All wrappers count launches in ``LAUNCHES`` (read by ``bench.py`` for its
``gpu_launches`` field).

Reference spans each kernel replaces are listed per entry point in the header; in short:
FedAvg reduce ``nanofed/server/aggregator/fedavg.py:53-63``, DP clip + noise
``privacy/mechanisms.py:85-129``, optimizer step / DP-SGD ``trainer/private.py:54-86``,
loss + accuracy + ``.item()`` syncs ``trainer/torch.py:14-22`` / ``trainer/base.py:153-159``.
"""

from __future__ import annotations

import ctypes as C
import os
from pathlib import Path
from typing import Sequence

import torch

_LIB_DIR = Path(__file__).resolve().parent.parent / "_native"
_LIB_NAME = "libnanofed_b200.so"
_lib: C.CDLL | None = None

LAUNCHES = 0  # kernels launched through this module since import / last reset


class NativeLibraryError(RuntimeError):
    """The sm_100a library is missing or failed to load."""


def lib_path() -> Path:
    return Path(os.environ.get("NANOFED_B200_LIB", _LIB_DIR / _LIB_NAME))


def is_available() -> bool:
    return lib_path().exists()


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        path = lib_path()
        if not path.exists():
            raise NativeLibraryError(
                f"{path} not found. Build it with `python -m nanofed_b200.build` "
                "(or `python -c 'import __graft_entry__ as g; g.build()'`). CUDA tensors "
                "are never routed to a PyTorch fallback."
            )
        try:
            _lib = C.CDLL(str(path))
        except OSError as exc:  # pragma: no cover
            raise NativeLibraryError(f"failed to load {path}: {exc}") from exc
        _declare(_lib)
    return _lib


def reset_launch_count() -> None:
    global LAUNCHES
    LAUNCHES = 0


def _count(n: int = 1) -> None:
    global LAUNCHES
    LAUNCHES += n


# ---------------------------------------------------------------------------
# helpers
# ---------------------------------------------------------------------------
_vp = C.c_void_p
_i64 = C.c_int64
_i32 = C.c_int
_f32 = C.c_float
_u64 = C.c_uint64
_sz = C.c_size_t


def _ptr(t: torch.Tensor | None) -> int | None:
    return None if t is None else t.data_ptr()


def _stream(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def _check(code: int, what: str) -> None:
    if code != 0:
        msg = lib().nfb_last_error()
        raise RuntimeError(f"{what} failed ({code}): {msg.decode() if msg else '?'}")


_DT = {torch.float32: 0, torch.bfloat16: 1, torch.float16: 2}


def _dt(t: torch.Tensor) -> int:
    try:
        return _DT[t.dtype]
    except KeyError:
        raise TypeError(f"unsupported dtype {t.dtype}; expected fp32/bf16/fp16") from None


def _declare(L: C.CDLL) -> None:
    L.nfb_last_error.restype = C.c_char_p
    L.nfb_last_error.argtypes = []
    L.nfb_version.restype = C.c_int
    L.nfb_stats_scratch_bytes.restype = C.c_size_t
    L.nfb_stats_scratch_bytes.argtypes = [_i64]
    sigs = {
        "nfb_philox_fill": [_vp, _i32, _i64, _i32, _f32, _u64, _u64, _vp],
        "nfb_update_stats": [_vp, _i32, _i64, _i64, _i64, _vp, _vp, _vp, _sz, _vp],
        "nfb_multi_sumsq": [_vp, _vp, _i32, _i32, _vp, _vp, _vp, _sz, _vp],
        "nfb_segment_sumsq": [_vp, _i32, _i64, _i64, _vp, _i32, _vp, _vp, _sz, _vp],
        "nfb_fedavg_reduce": [_vp, _i32, _i64, _i64, _i64, _vp, _vp, _vp, _f32, _f32, _vp,
                              _vp, _u64, _u64, _vp, _vp, _i32, _i32, _vp, _vp],
        "nfb_bcast_global": [_vp, _i64, _i64, _vp, _i64, _vp, _i64, _vp, _i64, _vp],
        "nfb_fused_sgd": [_vp, _vp, _i32, _vp, _vp, _i64, _f32, _f32, _f32, _f32, _i32, _i32,
                          _vp, _f32, _u64, _u64, _i64, _i64, _vp],
        "nfb_fused_sgd_skip": [_vp, _vp, _i32, _vp, _vp, _i64, _f32, _f32, _f32, _f32, _i32, _i32,
                          _vp, _f32, _u64, _u64, _i64, _i64, _i64, _i64, _i64, _vp],
        "nfb_fused_sgd_perm": [_vp, _vp, _i32, _vp, _vp, _i64, _f32, _f32, _f32, _f32, _i32, _i32,
                               _vp, _f32, _u64, _u64, _i64, _i64, _i64, _i64, _i64, _vp, _i64,
                               _i64, _i64, _i32, _i32, _vp],
        "nfb_fused_sgd_dp": [_vp, _vp, _i32, _vp, _vp, _i64, _f32, _f32, _f32, _f32, _i32, _i32,
                             _f32, _u64, _u64, _i64, _i64, _i64, _i64, _i64, _vp, _f32, _vp, _vp,
                             _vp, _i64, _i64, _i64, _i32, _i32, _vp, _vp],
        "nfb_batch_ghost_norm": [_vp, _vp, _i32, _i32, _i32, _vp, _vp, _vp, _vp],
        "nfb_fused_adam": [_vp, _vp, _i32, _vp, _vp, _vp, _i64, _f32, _f32, _f32, _f32, _f32,
                           _f32, _f32, _i32, _vp, _vp],
        "nfb_xent_fused": [_vp, _i32, _vp, _i32, _i64, _i32, _i32, _vp, _i32, _f32, _vp, _vp,
                           _vp, _vp],
        "nfb_linear_psgn": [_vp, _i32, _vp, _i32, _i64, _i64, _i64, _i32, _vp, _vp],
        "nfb_mask_add": [_vp, _vp, _i64, _u64, _u64, _vp],
    }
    for name, args in sigs.items():
        fn = getattr(L, name, None)
        if fn is None:
            continue
        fn.restype = _i32
        fn.argtypes = args


# ---------------------------------------------------------------------------
# K-RNG
# ---------------------------------------------------------------------------
def philox_fill_(out: torch.Tensor, kind: int, scale: float, seed: int, offset: int) -> torch.Tensor:
    assert out.is_contiguous()
    _check(lib().nfb_philox_fill(out.data_ptr(), _dt(out), out.numel(), kind, scale,
                                 seed & (2**64 - 1), offset, _stream(out)), "philox_fill")
    _count()
    return out


# ---------------------------------------------------------------------------
# K-STATS
# ---------------------------------------------------------------------------
_scratch: dict[tuple[int, int], torch.Tensor] = {}


def stats_scratch(device: torch.device, rows: int) -> torch.Tensor:
    """Zero-initialised reduction scratch for ``rows`` rows, cached per (device, stream).

    The kernels leave it zeroed ("self-cleaning"), so it is zero-filled only when it is
    (re)allocated.  Keyed by stream because two launches that overlap must not share it.
    """
    need = int(lib().nfb_stats_scratch_bytes(rows))
    key = (device.index or 0, torch.cuda.current_stream(device).cuda_stream)
    buf = _scratch.get(key)
    if buf is None or buf.numel() < need:
        buf = torch.zeros(max(need, 1 << 16), dtype=torch.uint8, device=device)
        _scratch[key] = buf
    return buf


def update_stats(x: torch.Tensor) -> tuple[torch.Tensor, torch.Tensor]:
    assert x.dim() == 2 and x.stride(1) == 1
    k, n = x.shape
    sumsq = torch.empty(k, dtype=torch.float32, device=x.device)
    bad = torch.empty(k, dtype=torch.int32, device=x.device)
    scratch = stats_scratch(x.device, k)
    _check(lib().nfb_update_stats(x.data_ptr(), _dt(x), k, n, x.stride(0), sumsq.data_ptr(),
                                  bad.data_ptr(), scratch.data_ptr(), scratch.numel(),
                                  _stream(x)), "update_stats")
    _count()
    return sumsq, bad


def _ptr_table(tensors: Sequence[torch.Tensor]) -> tuple[torch.Tensor, torch.Tensor, int]:
    dev = tensors[0].device
    dts = {t.dtype for t in tensors}
    if len(dts) != 1:
        tensors = [t.to(torch.float32) for t in tensors]
    tensors = [t if t.is_contiguous() else t.contiguous() for t in tensors]
    ptrs = torch.tensor([t.data_ptr() for t in tensors], dtype=torch.int64)
    lens = torch.tensor([t.numel() for t in tensors], dtype=torch.int64)
    table = torch.stack([ptrs, lens]).pin_memory().to(dev, non_blocking=True)
    table._keepalive = tensors  # type: ignore[attr-defined]
    return table[0], table[1], _dt(tensors[0])


def _multi(ptrs, lens, dt, count, dev):
    out = torch.empty(count, dtype=torch.float32, device=dev)
    bad = torch.empty(count, dtype=torch.int32, device=dev)
    scratch = stats_scratch(dev, count)
    _check(lib().nfb_multi_sumsq(ptrs.data_ptr(), lens.data_ptr(), count, dt, out.data_ptr(),
                                 bad.data_ptr(), scratch.data_ptr(), scratch.numel(),
                                 torch.cuda.current_stream(dev).cuda_stream), "multi_sumsq")
    _count()
    return out, bad


def sumsq_multi(tensors: Sequence[torch.Tensor]) -> torch.Tensor:
    ptrs, lens, dt = _ptr_table(tensors)
    dev = tensors[0].device
    out, _ = _multi(ptrs, lens, dt, len(tensors), dev)
    return out.sum()


def per_tensor_stats(tensors: Sequence[torch.Tensor]) -> tuple[torch.Tensor, torch.Tensor]:
    ptrs, lens, dt = _ptr_table(tensors)
    dev = tensors[0].device
    return _multi(ptrs, lens, dt, len(tensors), dev)


def segment_sumsq(x: torch.Tensor, segments: torch.Tensor) -> torch.Tensor:
    assert x.dim() == 2 and x.stride(1) == 1
    k, _ = x.shape
    seg = segments.to(device=x.device, dtype=torch.int64).contiguous()
    s = seg.shape[0]
    out = torch.zeros(k, s, dtype=torch.float32, device=x.device)
    if s:
        scratch = stats_scratch(x.device, k * s)
        _check(lib().nfb_segment_sumsq(x.data_ptr(), _dt(x), k, x.stride(0), seg.data_ptr(), s,
                                       out.data_ptr(), scratch.data_ptr(), scratch.numel(),
                                       _stream(x)), "segment_sumsq")
        _count()
    return out


# ---------------------------------------------------------------------------
# K-AGG
# ---------------------------------------------------------------------------
class _BcastTarget(C.Structure):
    _fields_ = [("slots", C.c_void_p), ("slots_stride", C.c_int64), ("lowp", C.c_void_p),
                ("lowp_stride", C.c_int64), ("momentum", C.c_void_p),
                ("momentum_stride", C.c_int64), ("rows", C.c_int64)]


def fedavg_reduce(
    x: torch.Tensor,
    counts: torch.Tensor,
    *,
    out: torch.Tensor | None = None,
    out_lowp: torch.Tensor | None = None,
    sumsq: torch.Tensor | None = None,
    mask: torch.Tensor | None = None,
    clip_norm: float = 0.0,
    noise_std: float = 0.0,
    noise_mask: torch.Tensor | None = None,
    seed: int = 0,
    offset: int = 0,
    unit_noise: torch.Tensor | None = None,
    normalize: bool = True,
    accumulate: bool = False,
    bcast: tuple | None = None,
) -> torch.Tensor:
    """``bcast=(slots, slots_lowp, momentum)`` (each optional): K-BCAST fused into the same launch —
    the result is replicated into the client rows / bf16 copies and the momentum rows are zeroed."""
    assert x.dim() == 2 and x.stride(1) == 1, "x must be [K, N] with unit inner stride"
    k, n = x.shape
    dev = x.device
    if out is None:
        out = torch.empty(n, dtype=torch.float32, device=dev)
        accumulate = False
    assert out.dtype == torch.float32 and out.is_contiguous() and out.numel() == n
    counts = counts.to(device=dev, dtype=torch.float32).contiguous()
    if sumsq is not None:
        sumsq = sumsq.to(device=dev, dtype=torch.float32).contiguous()
    if mask is not None:
        mask = mask.to(device=dev, dtype=torch.float32).contiguous()
    if noise_mask is not None:
        noise_mask = noise_mask.to(device=dev, dtype=torch.uint8).contiguous()
    if unit_noise is not None:
        unit_noise = unit_noise.to(device=dev, dtype=torch.float32).contiguous()
    lowp_dt = _dt(out_lowp) if out_lowp is not None else 1
    target = None
    if bcast is not None:
        b_slots, b_lowp, b_mom = bcast
        rows = next(t.shape[0] for t in (b_slots, b_lowp, b_mom) if t is not None)
        target = _BcastTarget(
            _ptr(b_slots), b_slots.stride(0) if b_slots is not None else 0,
            _ptr(b_lowp), b_lowp.stride(0) if b_lowp is not None else 0,
            _ptr(b_mom), b_mom.stride(0) if b_mom is not None else 0, rows)
    _check(
        lib().nfb_fedavg_reduce(
            x.data_ptr(), _dt(x), k, n, x.stride(0),
            counts.data_ptr(), _ptr(sumsq), _ptr(mask),
            float(clip_norm), float(noise_std), _ptr(noise_mask), _ptr(unit_noise),
            seed & (2**64 - 1), offset,
            out.data_ptr(), _ptr(out_lowp), lowp_dt,
            (1 if normalize else 0) | (2 if accumulate else 0),
            C.byref(target) if target is not None else None, _stream(x),
        ),
        "fedavg_reduce",
    )
    _count()
    return out


# ---------------------------------------------------------------------------
# K-BCAST
# ---------------------------------------------------------------------------
def bcast_global(master, slots=None, slots_lowp=None, momentum=None) -> None:
    assert master.dtype == torch.float32 and master.is_contiguous()
    n = master.numel()

    def rows(t):
        if t is None:
            return 0, 0
        assert t.dim() == 2 and t.stride(1) == 1 and t.shape[1] == n
        return t.shape[0], t.stride(0)

    k1, s1 = rows(slots)
    k2, s2 = rows(slots_lowp)
    k3, s3 = rows(momentum)
    k = max(k1, k2, k3)
    for kk in (k1, k2, k3):
        assert kk in (0, k)
    if slots is not None:
        assert slots.dtype == torch.float32
    if slots_lowp is not None:
        assert slots_lowp.dtype == torch.bfloat16
    if momentum is not None:
        assert momentum.dtype == torch.float32
    _check(lib().nfb_bcast_global(master.data_ptr(), n, k, _ptr(slots), s1, _ptr(slots_lowp), s2,
                                  _ptr(momentum), s3, _stream(master)), "bcast_global")
    _count()


# ---------------------------------------------------------------------------
# K-OPT
# ---------------------------------------------------------------------------
def fused_sgd_(
    p: torch.Tensor,
    g: torch.Tensor,
    lr: float,
    *,
    momentum_buf: torch.Tensor | None = None,
    momentum: float = 0.0,
    dampening: float = 0.0,
    weight_decay: float = 0.0,
    nesterov: bool = False,
    first_step: bool = False,
    grad_scale: torch.Tensor | None = None,
    noise_std: float = 0.0,
    seed: int = 0,
    offset: int = 0,
    lowp_out: torch.Tensor | None = None,
    row_len: int = 0,
    zero_grad_prefix: int = 0,
    skip_range: tuple[int, int] | None = None,
    zero_grad_from: int = 0,
    perm_out: tuple[torch.Tensor, int, int, int, int] | None = None,
    row_sumsq: torch.Tensor | None = None,
    clip_norm: float = 0.0,
    noise_mask: torch.Tensor | None = None,
    step_counter: torch.Tensor | None = None,
    sumsq_out: torch.Tensor | None = None,
) -> torch.Tensor:
    """``sumsq_out`` (fp32 ``[rows]``, accumulated): K-STATS side output — Σ of the squares of the
    updated parameters per row over the columns this call touches.
    ``row_sumsq`` / ``clip_norm`` / ``noise_mask`` / ``step_counter``: DP-SGD for a cohort — row
    ``r`` of the ``[rows, row_len]`` buffers is clipped by ``min(1, clip_norm / sqrt(row_sumsq[r]))``
    and gets ``N(0, noise_std²)`` on the masked entries; ``step_counter`` (device int64) advances
    the noise stream inside a replayed CUDA graph."""
    assert p.dtype == torch.float32 and p.is_contiguous() and g.is_contiguous()
    assert p.numel() == g.numel()
    if momentum == 0.0:
        momentum_buf = None
    if (row_sumsq is not None or noise_mask is not None or step_counter is not None
            or sumsq_out is not None):
        assert grad_scale is None and row_len > 0
        if sumsq_out is not None:
            assert sumsq_out.dtype == torch.float32 and sumsq_out.numel() == p.numel() // row_len
        lo, hi = skip_range if skip_range is not None else (0, 0)
        if noise_mask is not None:
            noise_mask = noise_mask.to(device=p.device, dtype=torch.uint8).contiguous()
        pout = (None, 0, 0, 0, 0, 0)
        if perm_out is not None:
            out_t, plo, prows, pinner, ptaps = perm_out
            pout = (out_t.data_ptr(), out_t.stride(0), int(plo), int(prows), int(pinner), int(ptaps))
        _check(lib().nfb_fused_sgd_dp(
            p.data_ptr(), g.data_ptr(), _dt(g), _ptr(momentum_buf), _ptr(lowp_out), p.numel(),
            float(lr), float(momentum), float(dampening), float(weight_decay),
            1 if nesterov else 0, 1 if first_step else 0, float(noise_std), seed & (2**64 - 1),
            offset, int(row_len), int(zero_grad_prefix), int(lo), int(hi), int(zero_grad_from),
            _ptr(row_sumsq),
            float(clip_norm), _ptr(noise_mask), _ptr(step_counter), *pout, _ptr(sumsq_out),
            _stream(p)),
            "fused_sgd_dp")
        _count()
        return p
    if grad_scale is not None:
        grad_scale = grad_scale.to(device=p.device, dtype=torch.float32).reshape(1)
    common = (p.data_ptr(), g.data_ptr(), _dt(g), _ptr(momentum_buf), _ptr(lowp_out), p.numel(),
              float(lr), float(momentum), float(dampening), float(weight_decay),
              1 if nesterov else 0, 1 if first_step else 0, _ptr(grad_scale), float(noise_std),
              seed & (2**64 - 1), offset, int(row_len), int(zero_grad_prefix))
    if perm_out is not None:
        # (out [rows_of_p, >= n_perm] bf16, lo, rows, inner, taps): permuted copy of one parameter
        out_t, plo, prows, pinner, ptaps = perm_out
        assert out_t.dtype == torch.bfloat16 and out_t.is_contiguous()
        lo, hi = skip_range if skip_range is not None else (0, 0)
        _check(lib().nfb_fused_sgd_perm(*common, int(lo), int(hi), int(zero_grad_from),
                                        out_t.data_ptr(), out_t.stride(0), int(plo), int(prows),
                                        int(pinner), int(ptaps), _stream(p)), "fused_sgd_perm")
    elif skip_range is None and not zero_grad_from:
        _check(lib().nfb_fused_sgd(*common, _stream(p)), "fused_sgd")
    else:  # columns [lo, hi) of every row were already updated by the kernel that made their gradient
        lo, hi = skip_range if skip_range is not None else (0, 0)
        _check(lib().nfb_fused_sgd_skip(*common, int(lo), int(hi), int(zero_grad_from), _stream(p)),
               "fused_sgd_skip")
    _count()
    return p


def batch_ghost_norm(gram_x: torch.Tensor, d: torch.Tensor, add_a=None, add_b=None) -> torch.Tensor:
    """``out[c] = ‖Σ_b d[c,b] ⊗ x[c,b]‖_F² + add_a[c] + add_b[c]`` from ``gram_x[c] = x xᵀ`` (fp32
    ``[C, B, B]``, consumed and zeroed) and ``d`` (bf16 ``[C, B, J]``)."""
    c, b, j = d.shape
    assert gram_x.shape == (c, b, b) and gram_x.dtype == torch.float32 and gram_x.is_contiguous()
    assert d.dtype == torch.bfloat16 and d.is_contiguous()
    out = torch.empty(c, dtype=torch.float32, device=d.device)
    _check(lib().nfb_batch_ghost_norm(gram_x.data_ptr(), d.data_ptr(), c, b, j, _ptr(add_a),
                                      _ptr(add_b), out.data_ptr(), _stream(d)), "batch_ghost_norm")
    _count()
    return out


def fused_adam_(
    p, g, exp_avg, exp_avg_sq, step: int, lr: float, *, beta1=0.9, beta2=0.999, eps=1e-8,
    weight_decay=0.0, decoupled=False, grad_scale=None, lowp_out=None,
) -> torch.Tensor:
    assert p.dtype == torch.float32 and p.is_contiguous() and g.is_contiguous()
    if grad_scale is not None:
        grad_scale = grad_scale.to(device=p.device, dtype=torch.float32).reshape(1)
    bc1 = 1.0 - beta1**step
    bc2 = 1.0 - beta2**step
    _check(
        lib().nfb_fused_adam(
            p.data_ptr(), g.data_ptr(), _dt(g), exp_avg.data_ptr(), exp_avg_sq.data_ptr(),
            _ptr(lowp_out), p.numel(), float(lr), float(beta1), float(beta2), float(eps),
            float(weight_decay), float(bc1), float(bc2), 1 if decoupled else 0,
            _ptr(grad_scale), _stream(p),
        ),
        "fused_adam",
    )
    _count()
    return p


# ---------------------------------------------------------------------------
# K-XENT
# ---------------------------------------------------------------------------
def xent_fused(logits, targets, *, dlogits=None, stats=None, grad_scale=None):
    assert logits.is_contiguous()
    b, c = logits.shape[-2], logits.shape[-1]
    groups = logits.numel() // (b * c)
    lead = logits.shape[:-2]
    if targets.dtype not in (torch.int64, torch.int32):
        targets = targets.to(torch.int64)
    targets = targets.contiguous()
    if dlogits is None:
        dlogits = torch.empty_like(logits)
    loss = torch.empty(lead, dtype=torch.float32, device=logits.device)
    correct = torch.empty(lead, dtype=torch.float32, device=logits.device)
    scale = (1.0 / b) if grad_scale is None else float(grad_scale)
    _check(
        lib().nfb_xent_fused(
            logits.data_ptr(), _dt(logits), targets.data_ptr(),
            1 if targets.dtype == torch.int64 else 0, groups, b, c, dlogits.data_ptr(),
            _dt(dlogits), scale, loss.data_ptr(), correct.data_ptr(), _ptr(stats),
            _stream(logits),
        ),
        "xent_fused",
    )
    _count()
    return loss, correct, dlogits


# ---------------------------------------------------------------------------
# K-PSGN / K-MASK
# ---------------------------------------------------------------------------
def linear_psgn(activations, grad_outputs, bias=True) -> torch.Tensor:
    a = activations.flatten(1).contiguous()
    d = grad_outputs.flatten(1).contiguous()
    out = torch.empty(a.shape[0], dtype=torch.float32, device=a.device)
    _check(lib().nfb_linear_psgn(a.data_ptr(), _dt(a), d.data_ptr(), _dt(d), a.shape[0],
                                 a.shape[1], d.shape[1], 1 if bias else 0, out.data_ptr(),
                                 _stream(a)), "linear_psgn")
    _count()
    return out


def mask_add_(x, cumulative, seed: int, offset: int) -> torch.Tensor:
    assert x.dtype == torch.float32 and cumulative.dtype == torch.float32
    assert x.is_contiguous() and cumulative.is_contiguous()
    _check(lib().nfb_mask_add(x.data_ptr(), cumulative.data_ptr(), x.numel(),
                              seed & (2**64 - 1), offset, _stream(x)), "mask_add")
    _count()
    return x

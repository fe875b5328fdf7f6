"""Python face of the generic tcgen05 convolution kernels and the BatchNorm / head / optimizer
kernels of the ResNet cohort engine (``csrc/rn_conv.cu``, ``csrc/rn_elt.cu``).

Conventions: activations are bf16 NHWC ``[clients·batch, H, W, C]`` (C a multiple of 64);
convolution weights live per client in tap-major bf16 rows ``wt[client][t_off + (tap·Cout + co)·Cin
+ ci]`` (written by ``wm_convert`` / the SGD kernels), weight gradients in the same layout in fp32.

Replaces (reference): the torch library calls a ResNet-18 makes inside
``nanofed/trainer/base.py:134-160`` (``model(data)``, ``loss.backward()``, ``optimizer.step()``).
"""

from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import torch

from . import native
from .tc import status_word


class _Dp(C.Structure):
    _fields_ = [("row_sumsq", C.c_void_p), ("clip_norm", C.c_float), ("noise_std", C.c_float),
                ("seed", C.c_ulonglong), ("step_base", C.c_void_p), ("step", C.c_int),
                ("client0", C.c_int), ("clients_total", C.c_int), ("row_groups", C.c_longlong),
                ("idx_base", C.c_longlong)]


class _Act(C.Structure):
    _fields_ = [("ptr", C.c_void_p), ("imgs", C.c_int), ("H", C.c_int), ("W", C.c_int),
                ("C", C.c_int), ("parity", C.c_int)]


_declared = False


def _lib() -> C.CDLL:
    global _declared
    lib = native.lib()
    if not _declared:
        vp, i32, i64, f32, ll = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_longlong
        pa, pi, pll = C.POINTER(_Act), C.POINTER(C.c_int), C.POINTER(C.c_longlong)
        sigs = {
            "nfb_rn_conv_gemm": [i32, pa, pa, vp, vp, pi, i32, i32, i32, i32, i32, i32, pi, pa, i32,
                                 i32, vp, vp, vp, vp, vp, vp, i32, i32, i32, vp, vp],
            "nfb_rn_conv_wgrad": [pa, pa, i32, i32, i32, pi, vp, ll, i32, i32, i32, vp, vp, f32, f32,
                                  vp, vp],
            "nfb_rn_im2col3": [vp, ll, i32, i32, vp, vp],
            "nfb_rn_bn_apply": [vp, vp, vp, vp, ll, vp, pll, vp, pll, i32, i32, i32, i32, f32, f32, vp],
            "nfb_rn_bn_bwd": [i32, vp, vp, vp, vp, vp, vp, vp, ll, vp, ll, pll, vp, pll, vp, pll,
                              i32, i32, i32, f32, vp, vp],
            "nfb_rn_gn_stats": [vp, i32, i32, i32, vp, vp],
            "nfb_rn_gn_apply": [vp, vp, vp, vp, ll, vp, pll, vp, pll, i32, i32, i32, i32, i32, f32, vp],
            "nfb_rn_gn_bwd": [i32, vp, vp, vp, vp, vp, vp, ll, vp, ll, pll, vp, pll, vp, pll, vp, vp,
                              i32, i32, i32, i32, f32, vp],
            "nfb_rn_head": [vp, i32, i32, i32, vp, ll, vp, ll, ll, ll, vp, ll, ll, ll, vp, vp, vp],
            "nfb_rn_wm_convert": [vp, i32, ll, vp, ll, vp, vp, ll, i32, i32, vp],
            "nfb_rn_tap_sgd": [vp, i32, ll, vp, vp, vp, vp, ll, i32, f32, f32, f32, C.POINTER(_Dp), vp],
            "nfb_rn_small_sgd": [vp, i32, vp, ll, vp, vp, ll, i32, f32, f32, f32, vp, ll,
                                 C.POINTER(_Dp), vp],
        }
        for name, args in sigs.items():
            fn = getattr(lib, name)
            fn.restype = i32
            fn.argtypes = args
        _declared = True
    return lib


def _stream(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def _act(t: torch.Tensor, parity: bool = False) -> _Act:
    assert t.dtype == torch.bfloat16 and t.dim() == 4 and t.is_contiguous(), (t.dtype, t.shape)
    n, h, w, c = t.shape
    return _Act(t.data_ptr(), n, h, w, c, 1 if parity else 0)


def _ptr(t: torch.Tensor | None) -> int | None:
    return None if t is None else t.data_ptr()


def _check(code: int, what: str) -> None:
    native._check(code, what)
    native._count()


# ---------------------------------------------------------------------------------------------
# tap tables: rows of (dh, dw, ph, pw, wrow, src)
# ---------------------------------------------------------------------------------------------
def fwd_taps(ksize: int, stride: int, cout: int, src: int = 0) -> list[tuple[int, ...]]:
    """A-side taps of a forward convolution (also the x-side taps of its weight gradient)."""
    if ksize == 1:
        return [(0, 0, 0, 0, 0, src)]
    taps = []
    for kh in range(3):
        for kw in range(3):
            t = kh * 3 + kw
            if stride == 1:
                taps.append((kh - 1, kw - 1, 0, 0, t * cout, src))
            else:   # input pixel 2·o + k - 1 = parity (k+1)&1 at half-resolution index o - (k == 0)
                taps.append((-1 if kh == 0 else 0, -1 if kw == 0 else 0, (kh + 1) & 1, (kw + 1) & 1,
                             t * cout, src))
    return taps


def dgrad_taps(ksize: int, stride: int, cout: int, cls: tuple[int, int] = (0, 0),
               src: int = 0) -> list[tuple[int, ...]]:
    """dY-side taps of the data gradient; for stride 2 only the taps that reach output parity
    class ``cls`` (row parity, column parity) of dX."""
    if ksize == 1:
        return [(0, 0, 0, 0, 0, src)] if (stride == 1 or cls == (0, 0)) else []
    taps = []
    for kh in range(3):
        for kw in range(3):
            t = kh * 3 + kw
            if stride == 1:
                taps.append((1 - kh, 1 - kw, 0, 0, t * cout, src))
            else:
                # dX row 2·hh + p receives dY row ho with 2·ho + kh - 1 = 2·hh + p
                if (kh + 1) % 2 != cls[0] or (kw + 1) % 2 != cls[1]:
                    continue
                dh = 1 if kh == 0 else 0
                dw = 1 if kw == 0 else 0
                taps.append((dh, dw, 0, 0, t * cout, src))
    return taps


def _tap_array(taps: list[tuple[int, ...]]):
    """7 ints per tap for the C side: the 6 documented above + the output parity class (-1 = all)."""
    flat = [v for t in taps for v in (tuple(t) + (-1,))[:7]]
    return (C.c_int * len(flat))(*flat)


def dgrad_taps_all_classes(ksize: int, cout: int, sc_cout: int | None = None) -> list[tuple[int, ...]]:
    """Taps of a stride-2 data gradient for ALL four output parity classes (7th entry = class
    2·row_parity + col_parity), plus the 1x1 projection shortcut's tap (operand pair 1) in class 0."""
    taps: list[tuple[int, ...]] = []
    for cls in ((0, 0), (0, 1), (1, 0), (1, 1)):
        idx = 2 * cls[0] + cls[1]
        taps += [t + (idx,) for t in dgrad_taps(ksize, 2, cout, cls)]
        if sc_cout is not None:
            taps += [t + (idx,) for t in dgrad_taps(1, 2, sc_cout, cls, src=1)]
    return taps


@dataclass
class WeightRef:
    """One convolution's tap-major weights inside the per-client rows ``wt[clients, t_stride]``."""

    buf: torch.Tensor      # bf16 [clients, t_stride]
    t_off: int
    rows: int              # taps · Cout
    inner: int             # Cin

    @property
    def ptr(self) -> int:
        return self.buf.data_ptr() + 2 * self.t_off

    @property
    def dims(self) -> tuple[int, int, int]:
        return (self.inner, self.rows, self.buf.stride(0))


def pick_tile(positions_per_client: int, n_out: int, clients: int, sms: int = 148) -> tuple[int, int]:
    """(bn_tile, halves) of the persistent conv kernel.

    Measured on B200 (``profiles/rn_tile_sweep_persistent.log``): 128 output channels x 256 positions
    is the best or within 5 % of the best tile for every ResNet-18 layer - it keeps two accumulator
    sets in TMEM (the epilogue of a tile overlaps the MMAs of the next) and three 48 KB stages in
    flight; the 256 x 256 tile, best for the one-tile-per-CTA kernel of the first version, has a
    single accumulator set and two stages here and is 20-50 % slower."""
    halves = 2 if positions_per_client % 256 == 0 else 1
    bn = 128 if n_out % 128 == 0 else 64
    return (bn, halves)


def conv_gemm(
    dgrad: bool,
    a: torch.Tensor, w: WeightRef, out: torch.Tensor, *,
    clients: int, grid_hw: tuple[int, int], taps: list[tuple[int, ...]],
    a_parity: bool = False, out_parity: tuple[int, int] | None = None,
    a1: torch.Tensor | None = None, w1: WeightRef | None = None,
    stats: torch.Tensor | None = None, residual: torch.Tensor | None = None,
    res_mask: torch.Tensor | None = None, out_mask: torch.Tensor | None = None, n_out: int,
    tile: tuple[int, int] | None = None, all_classes: bool = False,
    bn_x: torch.Tensor | None = None, gn_stats: torch.Tensor | None = None,
) -> None:
    """``out[pos, n] = Σ_taps A[pos + tap, :] · W`` on tensor cores (see ``csrc/rn_conv.cu``)."""
    lib = _lib()
    imgs = a.shape[0] // clients
    h, wd = grid_hw
    bn, halves = tile or pick_tile(imgs * h * wd, n_out, clients)
    a0 = _act(a, a_parity)
    a1s = _act(a1, False) if a1 is not None else None
    o = _act(out, out_parity is not None)
    wd0 = list(w.dims) + (list(w1.dims) if w1 is not None else [0, 0, 0])
    dims = (C.c_int * 6)(*wd0)
    oph, opw = out_parity if out_parity is not None else (0, 0)
    assert not all_classes or out_parity is not None, "all_classes needs a parity-view output"
    if stats is not None:
        assert stats.dtype == torch.float32 and stats.shape == (clients, 2, n_out)
    st = status_word(a.device)
    code = lib.nfb_rn_conv_gemm(
        1 if dgrad else 0, C.byref(a0), C.byref(a1s) if a1s is not None else None, w.ptr,
        w1.ptr if w1 is not None else None, dims, clients, imgs, h, wd, n_out, len(taps),
        _tap_array(taps), C.byref(o), oph, opw, _ptr(stats), _ptr(residual), _ptr(res_mask),
        _ptr(out_mask), _ptr(bn_x), _ptr(gn_stats), bn, halves, 4 if all_classes else 1,
        st.data_ptr(), _stream(a))
    _check(code, "rn_conv_gemm")


def conv_wgrad(x: torch.Tensor, dy: torch.Tensor, dwt: torch.Tensor, t_off: int, *, clients: int,
               taps: list[tuple[int, ...]], x_parity: bool = False, k_split: int = 1,
               tile: tuple[int, int] | None = None,
               sgd: tuple[torch.Tensor, torch.Tensor, float, float] | None = None) -> None:
    """``dwt[client][t_off + (tap·Cout + co)·Cin + ci] (+)= Σ_pos x[pos + tap, ci] · dy[pos, co]``.

    ``sgd=(wm, wt, lr, wd)`` (needs ``k_split == 1``): the gradient is applied to the tap-major
    fp32 master ``wm`` / bf16 copy ``wt`` inside the epilogue instead of being written."""
    lib = _lib()
    imgs = dy.shape[0] // clients
    cout = dy.shape[3]
    bn, posk = tile or (min(cout, 256), 128)
    st = status_word(x.device)
    xa, dya = _act(x, x_parity), _act(dy)
    wm_ptr = wt_ptr = None
    lr = wd = 0.0
    if sgd is not None:
        wm, wt, lr, wd = sgd
        assert wm.stride(0) == dwt.stride(0) == wt.stride(0)
        wm_ptr, wt_ptr = wm.data_ptr() + 4 * t_off, wt.data_ptr() + 2 * t_off
    code = lib.nfb_rn_conv_wgrad(C.byref(xa), C.byref(dya), clients, imgs, len(taps),
                                 _tap_array(taps), dwt.data_ptr() + 4 * t_off, dwt.stride(0),
                                 k_split, bn, posk, wm_ptr, wt_ptr, lr, wd, st.data_ptr(),
                                 _stream(x))
    _check(code, "rn_conv_wgrad")


def im2col3(x: torch.Tensor, out: torch.Tensor) -> None:
    """x fp32 ``[clients, batch, 3, 32, 32]`` (client stride free) → bf16 ``[clients·batch, 32, 32, 64]``."""
    assert x.dtype == torch.float32 and x.shape[2:] == (3, 32, 32) and x[0].is_contiguous()
    _check(_lib().nfb_rn_im2col3(x.data_ptr(), x.stride(0), x.shape[0], x.shape[1], out.data_ptr(),
                                 _stream(x)), "rn_im2col3")


def _ll(vals) -> "C.Array":
    return (C.c_longlong * len(vals))(*[int(v) for v in vals])


def bn_apply(x, y, slots, stats, bn_offs, *, clients: int, relu: bool, res=None, res_stats=None,
             res_offs=None, eps: float = 1e-5, momentum: float = 0.1) -> None:
    c = x.shape[-1]
    p = x.numel() // c // clients
    _check(_lib().nfb_rn_bn_apply(
        x.data_ptr(), _ptr(res), y.data_ptr(), slots.data_ptr(), slots.stride(0), stats.data_ptr(),
        _ll(bn_offs), _ptr(res_stats), _ll(res_offs) if res_offs is not None else None, clients, p,
        c, 1 if relu else 0, eps, momentum, _stream(x)), "rn_bn_apply")


def bn_bwd(apply: bool, dout, out, x, slots, sgrad, grad_offs, stats, bn_offs, *, clients: int,
           dx=None, x2=None, dx2=None, stats2=None, bn2_offs=None, eps: float = 1e-5,
           raw_sums: torch.Tensor | None = None) -> None:
    c = x.shape[-1]
    p = x.numel() // c // clients
    _check(_lib().nfb_rn_bn_bwd(
        1 if apply else 0, dout.data_ptr(), _ptr(out), x.data_ptr(), _ptr(x2), _ptr(dx), _ptr(dx2),
        slots.data_ptr(), slots.stride(0), sgrad.data_ptr(), sgrad.stride(0), _ll(grad_offs),
        stats.data_ptr(), _ll(bn_offs), _ptr(stats2),
        _ll(bn2_offs) if bn2_offs is not None else None, clients, p, c, eps, _ptr(raw_sums),
        _stream(x)), "rn_bn_bwd")


def gn_stats(x: torch.Tensor, stats: torch.Tensor) -> None:
    """GroupNorm(8) sums of ``x`` bf16 ``[imgs, h, w, c]`` into ``stats`` fp32 ``[imgs, 8, 2]``
    (accumulated: the buffer must be zero)."""
    imgs, h, w, c = x.shape
    assert stats.shape == (imgs, 8, 2) and stats.dtype == torch.float32
    _check(_lib().nfb_rn_gn_stats(x.data_ptr(), imgs, h * w, c, stats.data_ptr(), _stream(x)),
           "rn_gn_stats")


def gn_apply(x, y, slots, stats, gn_offs, *, clients: int, relu: bool, res=None, res_stats=None,
             res_offs=None, eps: float = 1e-5) -> None:
    imgs, h, w, c = x.shape
    _check(_lib().nfb_rn_gn_apply(
        x.data_ptr(), _ptr(res), y.data_ptr(), slots.data_ptr(), slots.stride(0), stats.data_ptr(),
        _ll(gn_offs[:2]), _ptr(res_stats), _ll(res_offs[:2]) if res_offs is not None else None,
        imgs, imgs // clients, h * w, c, 1 if relu else 0, eps, _stream(x)), "rn_gn_apply")


def gn_bwd(apply: bool, dout, x, slots, sgrad, grad_offs, stats, gn_offs, bsums, *, clients: int,
           dx=None, x2=None, dx2=None, stats2=None, gn2_offs=None, bsums2=None,
           eps: float = 1e-5) -> None:
    """``apply=False``: group sums into ``bsums`` (zero before) and dgamma / dbeta into ``sgrad``
    (atomics); ``apply=True``: dx (dx2) from those sums."""
    imgs, h, w, c = x.shape
    _check(_lib().nfb_rn_gn_bwd(
        1 if apply else 0, dout.data_ptr(), x.data_ptr(), _ptr(x2), _ptr(dx), _ptr(dx2),
        slots.data_ptr(), slots.stride(0), sgrad.data_ptr(), sgrad.stride(0), _ll(grad_offs),
        stats.data_ptr(), _ll(gn_offs[:2]), _ptr(stats2),
        _ll(gn2_offs[:2]) if gn2_offs is not None else None, bsums.data_ptr(), _ptr(bsums2),
        imgs, imgs // clients, h * w, c, eps, _stream(x)), "rn_gn_bwd")


def head(feat, labels, slots, w_off, b_off, sgrad, dw_off, db_off, dfeat, stats, *, clients: int) -> None:
    """feat bf16 ``[clients·batch, h, w, 512]``; labels int64 ``[clients, batch]`` (client stride free)."""
    batch = feat.shape[0] // clients
    _check(_lib().nfb_rn_head(
        feat.data_ptr(), clients, batch, feat.shape[1] * feat.shape[2], labels.data_ptr(),
        labels.stride(0), slots.data_ptr(), slots.stride(0), w_off, b_off, sgrad.data_ptr(),
        sgrad.stride(0), dw_off, db_off, dfeat.data_ptr(), stats.data_ptr(), _stream(feat)),
        "rn_head")


def wm_convert(table: torch.Tensor, n_layers: int, total_cols: int, slots, wm, wt, *,
               to_slots: bool) -> None:
    """Conv weights: arena rows (torch layout) → tap-major ``wm`` / ``wt``, or ``wm`` → arena rows."""
    _check(_lib().nfb_rn_wm_convert(
        table.data_ptr(), n_layers, total_cols, slots.data_ptr(), slots.stride(0), wm.data_ptr(),
        wt.data_ptr(), wt.stride(0), slots.shape[0], 1 if to_slots else 0, _stream(slots)),
        "rn_wm_convert")


def _dp(dp: dict | None, idx_base: int) -> "C.pointer | None":
    """``dp``: sumsq [clients] fp32, clip_norm, noise_std, seed, step_base (int64 device counter),
    step, client0, clients_total, row_groups - client-side DP-SGD inside the update kernels."""
    if dp is None:
        return None
    return C.byref(_Dp(dp["sumsq"].data_ptr(), float(dp["clip_norm"]), float(dp["noise_std"]),
                       int(dp["seed"]), dp["step_base"].data_ptr(), int(dp["step"]),
                       int(dp["client0"]), int(dp["clients_total"]), int(dp["row_groups"]),
                       int(idx_base)))


def tap_sgd(ranges: torch.Tensor, max_len: int, wm, dwt, momentum, wt, *, lr: float, mu: float,
            wd: float, dp: dict | None = None) -> None:
    _check(_lib().nfb_rn_tap_sgd(
        ranges.data_ptr(), ranges.shape[0], max_len, wm.data_ptr(), dwt.data_ptr(), _ptr(momentum),
        wt.data_ptr(), wt.stride(0), wm.shape[0], lr, mu, wd, _dp(dp, 0), _stream(wm)),
        "rn_tap_sgd")


def small_sgd(table: torch.Tensor, n_entries: int, slots, momentum, sgrad, *, lr: float, mu: float,
              wd: float, zero: torch.Tensor | None = None, dp: dict | None = None,
              dp_idx_base: int = 0) -> None:
    _check(_lib().nfb_rn_small_sgd(
        table.data_ptr(), n_entries, slots.data_ptr(), slots.stride(0), _ptr(momentum),
        sgrad.data_ptr(), sgrad.stride(0), slots.shape[0], lr, mu, wd, _ptr(zero),
        0 if zero is None else zero.numel(), _dp(dp, dp_idx_base), _stream(slots)), "rn_small_sgd")

"""Python face of the tcgen05 grouped GEMM / implicit-conv kernel (``csrc/tc_gemm.cu``).

All functions take *views into the arena* (bf16 activations/weights with the client
as leading dimension) and launch ONE kernel for all clients.  The image grid
convention for the conv ops: an activation is stored NHWC, row-flattened, as
``[C, R, ch]`` with ``R = B·H·W`` rows per client (``W`` = grid width, 26 for the
MNIST CNN); a 3×3 tap ``(kh, kw)`` is the row shift ``s = kh·W + kw``; outputs are
produced on the same grid (positions whose window leaves the image hold
don't-care values that consumers never read / are kept zero in gradients).

Replaces (reference): ``nn.Conv2d`` / ``nn.Linear`` forward and autograd backward
of ``nanofed/models/mnist.py:9-14,17-27`` as executed by ``trainer/base.py:145-150``.
"""

from __future__ import annotations

import ctypes as C
import os

import torch

from . import native

CONV_FWD, CONV_DGRAD, CONV_WGRAD, FC_FWD, FC_DGRAD, FC_WGRAD, FC_DGRAD_ROWS = range(7)
_BM = 128
_BN = {CONV_FWD: 64, CONV_DGRAD: 32, CONV_WGRAD: 32, FC_FWD: 64, FC_DGRAD: 64, FC_WGRAD: 128,
       FC_DGRAD_ROWS: 128}
_BK = {CONV_FWD: 32, CONV_DGRAD: 64, CONV_WGRAD: 64, FC_FWD: 64, FC_DGRAD: 64, FC_WGRAD: 64}


class _Operand(C.Structure):
    _fields_ = [("ptr", C.c_void_p), ("inner", C.c_longlong), ("rows", C.c_longlong),
                ("groups", C.c_longlong), ("row_stride", C.c_longlong),
                ("group_stride", C.c_longlong)]


class _Epilogue(C.Structure):
    _fields_ = [("out", C.c_void_p), ("out_bf16", C.c_int), ("atomic", C.c_int),
                ("out_gs", C.c_longlong), ("out_ms", C.c_longlong), ("out_ns", C.c_longlong),
                ("out_ts", C.c_longlong), ("bias", C.c_void_p), ("bias_gs", C.c_longlong),
                ("bias_on_m", C.c_int), ("relu", C.c_int), ("mask_src", C.c_void_p),
                ("m_valid", C.c_int), ("n_valid", C.c_int)]


_declared = False
_status: dict[int, torch.Tensor] = {}


def _lib() -> C.CDLL:
    global _declared
    lib = native.lib()
    if not _declared:
        lib.nfb_tc_gemm.restype = C.c_int
        lib.nfb_tc_gemm.argtypes = [C.c_int, C.POINTER(_Operand), C.POINTER(_Operand), C.c_int,
                                    C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                    C.POINTER(C.c_int), C.POINTER(_Epilogue), C.c_void_p,
                                    C.c_void_p]
        _declared = True
    return lib


def status_word(device: torch.device) -> torch.Tensor:
    """Per-device int32 that the kernels set to 1 if a pipeline wait times out."""
    idx = device.index or 0
    if idx not in _status:
        _status[idx] = torch.zeros(1, dtype=torch.int32, device=device)
    return _status[idx]


def check_status(device: torch.device) -> None:
    if int(status_word(device).item()) != 0:
        raise RuntimeError("tc_gemm: a TMA/MMA pipeline wait timed out (status word set)")


def _operand(t: torch.Tensor) -> _Operand:
    """``t``: bf16 ``[G, rows, inner]`` view with unit inner stride."""
    assert t.dtype == torch.bfloat16 and t.dim() == 3 and t.stride(2) == 1, (t.dtype, t.shape)
    return _Operand(t.data_ptr(), t.shape[2], t.shape[1], t.shape[0], t.stride(1), t.stride(0))


def gemm(
    variant: int,
    a: torch.Tensor,
    b: torch.Tensor,
    out: torch.Tensor,
    *,
    m: int,
    n: int,
    strides: tuple[int, int, int, int],      # element strides of out: (group, m, n, tap)
    taps: list[tuple[int, int, int, int]] | None = None,
    inner: int,
    k_split: int = 1,
    atomic: bool = False,
    bias: torch.Tensor | None = None,
    bias_gs: int = 0,
    bias_on_m: bool = False,
    relu: bool = False,
    mask_src: torch.Tensor | None = None,
) -> None:
    lib = _lib()
    groups = a.shape[0]
    oa, ob = _operand(a), _operand(b)
    taps = taps or [(0, 0, 0, 0)]
    table = (C.c_int * (4 * len(taps)))(*[v for tap in taps for v in tap])
    assert out.dtype in (torch.float32, torch.bfloat16)
    if atomic:
        assert out.dtype == torch.float32
    epi = _Epilogue(out.data_ptr(), 1 if out.dtype == torch.bfloat16 else 0, 1 if atomic else 0,
                    strides[0], strides[1], strides[2], strides[3],
                    bias.data_ptr() if bias is not None else None, bias_gs,
                    1 if bias_on_m else 0, 1 if relu else 0,
                    mask_src.data_ptr() if mask_src is not None else None, m, n)
    m_tiles = (m + _BM - 1) // _BM
    n_tiles = (n + _BN[variant] - 1) // _BN[variant]
    st = status_word(a.device)
    code = lib.nfb_tc_gemm(variant, C.byref(oa), C.byref(ob), m_tiles, n_tiles, groups,
                           len(taps), inner, k_split, table, C.byref(epi), st.data_ptr(),
                           torch.cuda.current_stream(a.device).cuda_stream)
    native._check(code, f"tc_gemm(variant={variant})")
    native._count()


def _conv_persistent(mode: int, a: torch.Tensor, w: torch.Tensor, out: torch.Tensor, width: int,
                     bias: torch.Tensor | None, mask: torch.Tensor | None) -> None:
    lib = _lib()
    if not getattr(lib, "_conv_declared", False):
        ll = C.c_longlong
        lib.nfb_tc_conv3x3.restype = C.c_int
        lib.nfb_tc_conv3x3.argtypes = [C.c_int, C.c_void_p, ll, ll, C.c_void_p, ll, C.c_int,
                                       C.c_int, C.c_void_p, ll, C.c_void_p, ll, C.c_void_p,
                                       C.c_void_p, C.c_void_p]
        lib._conv_declared = True
    for t in (a, w, out):
        assert t.dtype == torch.bfloat16 and t.dim() == 3 and t.is_contiguous()
    st = status_word(a.device)
    code = lib.nfb_tc_conv3x3(
        mode, a.data_ptr(), a.shape[1], a.stride(0), w.data_ptr(), w.stride(0), a.shape[0], width,
        out.data_ptr(), out.stride(0), bias.data_ptr() if bias is not None else None,
        bias.stride(0) if bias is not None else 0,
        mask.data_ptr() if mask is not None else None, st.data_ptr(),
        torch.cuda.current_stream(a.device).cuda_stream)
    native._check(code, f"tc_conv3x3(mode={mode})")
    native._count()


def conv1_fwd(x1: torch.Tensor, slots: torch.Tensor, w_off: int, b_off: int, a1: torch.Tensor,
              width: int) -> None:
    """conv (1→32, 3×3) + bias + ReLU as ONE tcgen05 MMA per 128 positions (``tc_conv1.cu``).

    x1 ``[C, B, ...]`` fp32 images of ``(W+2)²`` pixels (free client stride), slots ``[C, n]`` fp32
    parameter rows holding ``conv1.weight [32, 9]`` at ``w_off`` and ``conv1.bias`` at ``b_off``,
    a1 ``[C, B·W², 32]`` bf16 (NHWC, row-flattened).
    """
    lib = _lib()
    if not getattr(lib, "_conv1_declared", False):
        ll = C.c_longlong
        lib.nfb_tc_conv1_fwd.restype = C.c_int
        lib.nfb_tc_conv1_fwd.argtypes = [C.c_void_p, ll, C.c_void_p, ll, ll, ll, C.c_int, C.c_int,
                                         C.c_int, C.c_void_p, ll, C.c_void_p, C.c_void_p]
        lib._conv1_declared = True
    assert x1.dtype == torch.float32 and x1[0].is_contiguous() and slots.dtype == torch.float32
    assert a1.dtype == torch.bfloat16 and a1.is_contiguous() and a1.shape[2] == 32
    images = x1.shape[1]
    assert x1[0].numel() == images * (width + 2) ** 2 and a1.shape[1] == images * width * width
    st = status_word(a1.device)
    code = lib.nfb_tc_conv1_fwd(x1.data_ptr(), x1.stride(0), slots.data_ptr(), slots.stride(0),
                                int(w_off), int(b_off), a1.shape[0], images, width, a1.data_ptr(),
                                a1.stride(0), st.data_ptr(),
                                torch.cuda.current_stream(a1.device).cuda_stream)
    native._check(code, "tc_conv1_fwd")
    native._count()


def conv3x3_dgrad_wgrad1(dy: torch.Tensor, w: torch.Tensor, relu_mask: torch.Tensor,
                         x1: torch.Tensor, grads: torch.Tensor, w1_off: int, b1_off: int,
                         width: int) -> None:
    """Data gradient of the 32→64 conv fused with the weight/bias gradient of the 1→32 conv in
    front of it (``tc_conv.cu`` MODE 3): ``da1`` is masked, rounded to bf16 and fed straight to a
    second tcgen05 MMA against the 3×3 input windows — it is never written to memory.

    dy ``[C, B·W², 64]`` bf16, w ``[C, 64, 288]`` bf16, relu_mask ``[C, B·W², 32]`` bf16 (conv1
    output), x1 ``[C, B, ...]`` fp32 layer-1 images of ``(W+2)²`` pixels (client stride free),
    grads ``[C, n]`` fp32 rows: ``grads[c, w1_off + ch·9 + t]`` and ``grads[c, b1_off + ch]`` are
    accumulated with atomics (caller zeroes them).
    """
    lib = _lib()
    if not getattr(lib, "_dgw1_declared", False):
        ll = C.c_longlong
        lib.nfb_tc_conv3x3_dgrad_wgrad1.restype = C.c_int
        lib.nfb_tc_conv3x3_dgrad_wgrad1.argtypes = [C.c_void_p, ll, ll, C.c_void_p, ll, C.c_int,
                                                    C.c_int, C.c_void_p, ll, C.c_void_p, ll,
                                                    C.c_void_p, ll, ll, ll, C.c_void_p, C.c_void_p]
        lib._dgw1_declared = True
    for t in (dy, w, relu_mask):
        assert t.dtype == torch.bfloat16 and t.dim() == 3 and t.is_contiguous()
    assert x1.dtype == torch.float32 and grads.dtype == torch.float32 and grads.stride(1) == 1
    assert x1[0].is_contiguous() and x1[0].numel() * width * width == dy.shape[1] * (width + 2) ** 2
    st = status_word(dy.device)
    code = lib.nfb_tc_conv3x3_dgrad_wgrad1(
        dy.data_ptr(), dy.shape[1], dy.stride(0), w.data_ptr(), w.stride(0), dy.shape[0], width,
        relu_mask.data_ptr(), relu_mask.stride(0), x1.data_ptr(), x1.stride(0), grads.data_ptr(),
        grads.stride(0), int(w1_off), int(b1_off), st.data_ptr(),
        torch.cuda.current_stream(dy.device).cuda_stream)
    native._check(code, "tc_conv3x3_dgrad_wgrad1")
    native._count()


def conv3x3_fwd_persistent(x, w, bias, out, width: int) -> None:
    """Same contract as ``conv3x3_fwd`` (bias + ReLU), persistent slab kernel (``tc_conv.cu``)."""
    _conv_persistent(0, x, w, out, width, bias, None)


def conv3x3_pool_fwd(x, w, bias, pool_out, pool_idx, images: int, width: int, drop_p: float,
                     rng_state, seed: int) -> None:
    """conv3x3 (32→64) + bias + ReLU + 2×2 max-pool + dropout fused (``tc_conv.cu`` MODE 2).

    x ``[C, B·26·26, 32]`` bf16 → ``pool_out [C, B, 9216]`` bf16 (PyTorch flatten order) and
    ``pool_idx [C, B, 144, 64]`` uint8 codes; the conv output itself is never materialised.
    """
    lib = _lib()
    if not getattr(lib, "_pool_declared", False):
        ll = C.c_longlong
        lib.nfb_tc_conv3x3_pool_fwd.restype = C.c_int
        lib.nfb_tc_conv3x3_pool_fwd.argtypes = [C.c_void_p, ll, ll, C.c_void_p, ll, C.c_int, C.c_int,
                                                C.c_int, C.c_void_p, ll, C.c_void_p, C.c_void_p,
                                                C.c_float, C.c_void_p, C.c_ulonglong, C.c_void_p,
                                                C.c_void_p]
        lib._pool_declared = True
    assert x.is_contiguous() and w.is_contiguous() and pool_out.is_contiguous()
    st = status_word(x.device)
    code = lib.nfb_tc_conv3x3_pool_fwd(
        x.data_ptr(), x.shape[1], x.stride(0), w.data_ptr(), w.stride(0), x.shape[0], images, width,
        bias.data_ptr() if bias is not None else None, bias.stride(0) if bias is not None else 0,
        pool_out.data_ptr(), pool_idx.data_ptr(), float(drop_p),
        rng_state.data_ptr() if rng_state is not None else None, seed, st.data_ptr(),
        torch.cuda.current_stream(x.device).cuda_stream)
    native._check(code, "tc_conv3x3_pool_fwd")
    native._count()


def conv3x3_dgrad_persistent(dy, w, out, width: int, relu_mask=None) -> None:
    """Same contract as ``conv3x3_dgrad``, persistent slab kernel (``tc_conv.cu``)."""
    _conv_persistent(1, dy, w, out, width, None, relu_mask)


def conv3x3_wgrad_persistent(dy: torch.Tensor, x: torch.Tensor, grad: torch.Tensor, width: int,
                             grad_group_stride: int) -> None:
    """Same contract as ``conv3x3_wgrad`` (fp32 atomics into OIHW), persistent slab kernel."""
    lib = _lib()
    if not getattr(lib, "_wgrad_declared", False):
        ll = C.c_longlong
        lib.nfb_tc_conv3x3_wgrad.restype = C.c_int
        lib.nfb_tc_conv3x3_wgrad.argtypes = [C.c_void_p, C.c_void_p, ll, ll, ll, C.c_int, C.c_int,
                                             C.c_void_p, ll, C.c_void_p, C.c_void_p]
        lib._wgrad_declared = True
    assert dy.is_contiguous() and x.is_contiguous() and grad.dtype == torch.float32
    st = status_word(dy.device)
    code = lib.nfb_tc_conv3x3_wgrad(dy.data_ptr(), x.data_ptr(), dy.shape[1], dy.stride(0),
                                    x.stride(0), dy.shape[0], width, grad.data_ptr(),
                                    grad_group_stride, st.data_ptr(),
                                    torch.cuda.current_stream(dy.device).cuda_stream)
    native._check(code, "tc_conv3x3_wgrad")
    native._count()


def _shifts(width: int) -> list[int]:
    return [kh * width + kw for kh in range(3) for kw in range(3)]


# ---- 3x3 convolution on the flattened grid -------------------------------------------
def conv3x3_fwd(x: torch.Tensor, w: torch.Tensor, bias: torch.Tensor | None, out: torch.Tensor,
                width: int, relu: bool = True) -> None:
    """``out[c, q, co] = act(Σ_t Σ_ci x[c, q + s_t, ci]·w[c, co, t·Cin + ci] + bias[c, co])``.

    x ``[C, R, 32]``, w ``[C, 64, 288]`` (tap-major, channel-minor), out ``[C, R, 64]`` bf16.
    """
    cin = x.shape[2]
    assert cin == 32 and w.shape[1] == 64 and w.shape[2] == 9 * cin
    taps = [(0, s, t * cin, 0) for t, s in enumerate(_shifts(width))]
    gemm(CONV_FWD, x, w, out, m=x.shape[1], n=64,
         strides=(out.stride(0), out.stride(1), 1, 0), taps=taps, inner=1,
         bias=bias, bias_gs=bias.stride(0) if bias is not None else 0, relu=relu)


def conv3x3_dgrad(dy: torch.Tensor, w: torch.Tensor, out: torch.Tensor, width: int,
                  relu_mask: torch.Tensor | None = None) -> None:
    """``out[c, q, ci] = Σ_t Σ_co dy[c, q − s_t, co]·w[c, co, t·Cin + ci]`` (× ``relu_mask > 0``).

    dy ``[C, R, 64]`` (zero outside the valid output window), out ``[C, R, 32]`` bf16.
    """
    cin = out.shape[2]
    assert dy.shape[2] == 64 and cin == 32
    taps = [(0, -s, 0, t * cin) for t, s in enumerate(_shifts(width))]
    gemm(CONV_DGRAD, dy, w, out, m=dy.shape[1], n=cin,
         strides=(out.stride(0), out.stride(1), 1, 0), taps=taps, inner=1, mask_src=relu_mask)


def conv3x3_wgrad(dy: torch.Tensor, x: torch.Tensor, grad: torch.Tensor, width: int,
                  grad_group_stride: int, k_split: int = 8) -> None:
    """``grad[c, co, ci, t] += Σ_q dy[c, q, co]·x[c, q + s_t, ci]`` (PyTorch OIHW layout, fp32).

    ``grad`` points at the first client's conv weight gradient inside the flat gradient
    slab; ``grad_group_stride`` is the slab's row length.
    """
    rows = dy.shape[1]
    cin = x.shape[2]
    taps = [(0, 0, s, 0) for s in _shifts(width)]
    inner = (rows + 63) // 64
    gemm(CONV_WGRAD, dy, x, grad, m=64, n=cin,
         strides=(grad_group_stride, cin * 9, 9, 1), taps=taps, inner=inner,
         k_split=min(k_split, inner), atomic=True)


# ---- linear layers ---------------------------------------------------------------------
def linear_fwd(w: torch.Tensor, x: torch.Tensor, out: torch.Tensor, k_split: int = 8) -> None:
    """``out[c, b, j] += Σ_i x[c, b, i]·w[c, j, i]`` (fp32 atomics; zero ``out`` first)."""
    j, i = w.shape[1], w.shape[2]
    gemm(FC_FWD, w, x, out, m=j, n=x.shape[1], strides=(out.stride(0), 1, out.stride(1), 0),
         inner=i // 64, k_split=k_split, atomic=True)


def linear_wgrad_sgd(dy: torch.Tensor, x: torch.Tensor, w: torch.Tensor, w_lowp: torch.Tensor,
                     lr: float, weight_decay: float = 0.0, dp: dict | None = None,
                     sumsq_out: torch.Tensor | None = None) -> None:
    """``w[c, j, i] -= lr·(Σ_b dy[c, b, j]·x[c, b, i] + wd·w)`` and ``w_lowp = bf16(w)`` in ONE
    kernel: the weight gradient is consumed in the tcgen05 epilogue and never written.  ``w`` (fp32)
    and ``w_lowp`` (bf16) are ``[C, j, i]`` views with contiguous ``[j, i]`` blocks and arbitrary
    client strides (rows of the parameter arena).  Everything that reads the OLD weights in this
    step (the data gradient) must already be enqueued.  ``sumsq_out`` (fp32 ``[C]``, accumulated):
    Σ of the squares of the new weights per client (K-STATS side output)."""
    lib = _lib()
    if not getattr(lib, "_wgrad_sgd_declared", False):
        ll = C.c_longlong
        lib.nfb_tc_linear_wgrad_sgd.restype = C.c_int
        lib.nfb_tc_linear_wgrad_sgd.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                                C.c_int, C.c_void_p, ll, C.c_void_p, ll, C.c_float,
                                                C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]
        lib._wgrad_sgd_declared = True
    c, j, i = w.shape
    assert w.dtype == torch.float32 and w_lowp.dtype == torch.bfloat16 and w_lowp.shape == w.shape
    assert w.stride(2) == 1 and w.stride(1) == i and w_lowp.stride(2) == 1 and w_lowp.stride(1) == i
    assert dy.shape[0] == c and x.shape[0] == c and dy.shape[2] == j and x.shape[2] == i
    oa, ob = _operand(dy), _operand(x)
    st = status_word(dy.device)
    if dp is not None:
        # DP-SGD front end in the epilogue: dp = {row_sumsq, clip_norm, noise_std, seed, offset,
        # step_counter, row_len, w_off} (see nfb_tc_linear_wgrad_sgd_dp)
        if not getattr(lib, "_wgrad_sgd_dp_declared", False):
            ll = C.c_longlong
            lib.nfb_tc_linear_wgrad_sgd_dp.restype = C.c_int
            lib.nfb_tc_linear_wgrad_sgd_dp.argtypes = [
                C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, ll,
                C.c_void_p, ll, C.c_float, C.c_float, C.c_void_p, C.c_float, C.c_float,
                C.c_uint64, C.c_uint64, C.c_void_p, ll, ll, C.c_void_p, C.c_void_p]
            lib._wgrad_sgd_dp_declared = True
        ctr = dp.get("step_counter")
        code = lib.nfb_tc_linear_wgrad_sgd_dp(
            C.byref(oa), C.byref(ob), c, j, i, (dy.shape[1] + 63) // 64, w.data_ptr(), w.stride(0),
            w_lowp.data_ptr(), w_lowp.stride(0), float(lr), float(weight_decay),
            dp["row_sumsq"].data_ptr(), float(dp["clip_norm"]), float(dp["noise_std"]),
            int(dp["seed"]) & (2**64 - 1), int(dp.get("offset", 0)),
            ctr.data_ptr() if ctr is not None else None, int(dp["row_len"]), int(dp["w_off"]),
            st.data_ptr(), torch.cuda.current_stream(dy.device).cuda_stream)
        native._check(code, "tc_linear_wgrad_sgd_dp")
        native._count()
        return
    code = lib.nfb_tc_linear_wgrad_sgd(C.byref(oa), C.byref(ob), c, j, i, (dy.shape[1] + 63) // 64,
                                       w.data_ptr(), w.stride(0), w_lowp.data_ptr(),
                                       w_lowp.stride(0), float(lr), float(weight_decay),
                                       sumsq_out.data_ptr() if sumsq_out is not None else None,
                                       st.data_ptr(),
                                       torch.cuda.current_stream(dy.device).cuda_stream)
    native._check(code, "tc_linear_wgrad_sgd")
    native._count()


def fc_dgrad_is_persistent(j: int, batch: int, i: int) -> bool:
    """True when ``linear_dgrad`` runs the persistent kernel (which can also scale its output)."""
    return (j == 128 and batch <= 128 and i % 64 == 0
            and os.environ.get("NFB_NO_PERSISTENT_FC") is None)


def linear_dgrad(w: torch.Tensor, dy: torch.Tensor, out: torch.Tensor, scale: float = 1.0) -> None:
    """``out[c, b, i] = scale·Σ_j dy[c, b, j]·w[c, j, i]`` (bf16; ``scale`` only on the
    persistent path, see ``fc_dgrad_is_persistent``)."""
    j, i = w.shape[1], w.shape[2]
    if fc_dgrad_is_persistent(j, dy.shape[1], i) and out.stride(2) == 1:
        # persistent form (tc_fc.cu): dy resident, weight tiles streamed, double-buffered TMEM
        lib = _lib()
        if not getattr(lib, "_fc_dgrad_declared", False):
            lib.nfb_tc_fc_dgrad.restype = C.c_int
            lib.nfb_tc_fc_dgrad.argtypes = [C.POINTER(_Operand), C.POINTER(_Operand), C.c_void_p,
                                            C.c_longlong, C.c_longlong, C.c_float, C.c_void_p,
                                            C.c_void_p]
            lib._fc_dgrad_declared = True
        assert out.dtype == torch.bfloat16 and out.shape == (w.shape[0], dy.shape[1], i)
        oa, ob = _operand(dy), _operand(w)
        st = status_word(dy.device)
        code = lib.nfb_tc_fc_dgrad(C.byref(oa), C.byref(ob), out.data_ptr(), out.stride(1),
                                   out.stride(0), float(scale), st.data_ptr(),
                                   torch.cuda.current_stream(dy.device).cuda_stream)
        native._check(code, "tc_fc_dgrad")
        native._count()
        return
    assert scale == 1.0, "output scaling needs the persistent fc dgrad kernel"
    if dy.shape[1] <= 128 and out.is_contiguous():
        # M = batch rows (zero-padded to 128 by TMA), N = i: row-major output, TMA-store epilogue
        gemm(FC_DGRAD_ROWS, dy, w, out, m=dy.shape[1], n=i,
             strides=(out.stride(0), out.stride(1), 1, 0), inner=j // 64)
        return
    gemm(FC_DGRAD, w, dy, out, m=i, n=dy.shape[1], strides=(out.stride(0), 1, out.stride(1), 0),
         inner=j // 64)


def linear_wgrad(dy: torch.Tensor, x: torch.Tensor, grad: torch.Tensor,
                 grad_group_stride: int) -> None:
    """``grad[c, j, i] = Σ_b dy[c, b, j]·x[c, b, i]`` (fp32 store into the gradient slab)."""
    j, i = dy.shape[2], x.shape[2]
    # (M = j, N = i): each thread stores 16-byte vectors along the contiguous i dimension.
    # The swapped orientation (lanes along i, scalar stores) measured 1.7x slower.
    gemm(FC_WGRAD, dy, x, grad, m=j, n=i, strides=(grad_group_stride, i, 1, 0),
         inner=(dy.shape[1] + 63) // 64)


def conv_psgn(unfolded: torch.Tensor, grad_outputs: torch.Tensor, bias: bool = True) -> torch.Tensor:
    """K-PSGN for convolutions on tensor cores: ``out[i] = ‖δ_iᵀ U_i‖_F² (+ ‖Σ_p δ_i[p]‖²)``.

    ``unfolded`` ``[B, P, K]`` (im2col patches, ``K = Cin·kh·kw``), ``grad_outputs`` ``[B, P, Cout]``.
    Each sample is one group of the grouped tcgen05 GEMM (both operands MN-major); the epilogue
    sums the squares of the per-sample gradient straight out of TMEM, so the ``B × Cout × K``
    per-sample gradients never exist in memory.  The bias term rides along as a column of ones
    appended to the patches.  Operands are rounded to bf16 (fp32 accumulation): relative error of
    the norms ≈ 3e-3.  Reference span: ``nanofed/trainer/private.py:54-86``.
    """
    lib = _lib()
    if not getattr(lib, "_nfb_psgn_declared", False):
        lib.nfb_tc_psgn.restype = C.c_int
        lib.nfb_tc_psgn.argtypes = [C.POINTER(_Operand), C.POINTER(_Operand), C.c_int, C.c_int,
                                    C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        lib._nfb_psgn_declared = True
    b, p, k = unfolded.shape
    cout = grad_outputs.shape[2]
    assert grad_outputs.shape[:2] == (b, p)
    kp = -(-(k + (1 if bias else 0)) // 8) * 8          # 16-byte rows for the TMA descriptors
    cp = -(-cout // 8) * 8
    u = torch.zeros(b, p, kp, dtype=torch.bfloat16, device=unfolded.device)
    u[:, :, :k] = unfolded
    if bias:
        u[:, :, k] = 1.0
    d = grad_outputs
    if cp != cout or d.dtype != torch.bfloat16 or not d.is_contiguous():
        dd = torch.zeros(b, p, cp, dtype=torch.bfloat16, device=d.device)
        dd[:, :, :cout] = d
        d = dd
    out = torch.zeros(b, dtype=torch.float32, device=u.device)
    st = status_word(u.device)
    od, ou = _operand(d), _operand(u)
    code = lib.nfb_tc_psgn(C.byref(od), C.byref(ou), b, cp, kp, (p + 63) // 64, out.data_ptr(),
                           st.data_ptr(), torch.cuda.current_stream(u.device).cuda_stream)
    native._check(code, "tc_psgn")
    native._count()
    return out

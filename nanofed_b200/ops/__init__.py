"""Layer E: the op surface.  CUDA tensors → hand-written sm_100a kernels
(``ops.native`` → ``csrc/*.cu`` through the C ABI in ``include/nanofed_b200.h``);
CPU tensors → ``ops.reference`` (pure PyTorch, also the test oracle).

There is no silent fallback on a GPU: a CUDA tensor with the native library
missing raises ``NativeLibraryError`` telling the user to build it.
"""

from __future__ import annotations

from typing import Sequence

import torch

from . import philox, reference
from .native import NativeLibraryError, is_available as native_available, lib as _native_lib

__all__ = [
    "fedavg_reduce", "update_stats", "sumsq_multi", "per_tensor_stats", "segment_sumsq",
    "bcast_global", "fused_sgd_", "fused_adam_", "xent_fused",
    "philox_normal_", "philox_laplace_", "philox_uniform_", "mask_add_",
    "linear_per_sample_grad_sqnorms", "conv_per_sample_grad_sqnorms",
    "native_available", "NativeLibraryError", "reference", "philox",
]


def _native():
    from . import native

    native.lib()  # raises NativeLibraryError with build instructions if absent
    return native


def _on_gpu(*tensors: torch.Tensor | None) -> bool:
    return any(t is not None and t.is_cuda for t in tensors)


# -- K-RNG ------------------------------------------------------------------
def philox_normal_(out: torch.Tensor, scale: float, seed: int, offset: int = 0) -> torch.Tensor:
    if out.is_cuda:
        return _native().philox_fill_(out, 0, float(scale), seed, offset)
    return reference.philox_normal_(out, scale, seed, offset)


def philox_laplace_(out: torch.Tensor, scale: float, seed: int, offset: int = 0) -> torch.Tensor:
    if out.is_cuda:
        return _native().philox_fill_(out, 1, float(scale), seed, offset)
    return reference.philox_laplace_(out, scale, seed, offset)


def philox_uniform_(out: torch.Tensor, seed: int, offset: int = 0) -> torch.Tensor:
    if out.is_cuda:
        return _native().philox_fill_(out, 2, 1.0, seed, offset)
    return reference.philox_uniform_(out, seed, offset)


# -- K-STATS ------------------------------------------------------------------
def update_stats(x: torch.Tensor) -> tuple[torch.Tensor, torch.Tensor]:
    if x.is_cuda:
        return _native().update_stats(x)
    return reference.update_stats(x)


def sumsq_multi(tensors: Sequence[torch.Tensor]) -> torch.Tensor:
    tensors = [t for t in tensors if t.numel() > 0]
    if tensors and all(t.is_cuda for t in tensors):
        return _native().sumsq_multi(tensors)
    return reference.sumsq_multi(tensors)


def per_tensor_stats(tensors: Sequence[torch.Tensor]) -> tuple[torch.Tensor, torch.Tensor]:
    if tensors and all(t.is_cuda for t in tensors):
        return _native().per_tensor_stats(list(tensors))
    return reference.per_tensor_stats(tensors)


def segment_sumsq(x: torch.Tensor, segments: torch.Tensor) -> torch.Tensor:
    if x.is_cuda:
        return _native().segment_sumsq(x, segments)
    return reference.segment_sumsq(x, segments)


# -- K-AGG ------------------------------------------------------------------
def fedavg_reduce(x: torch.Tensor, counts: torch.Tensor, **kw) -> torch.Tensor:
    if x.is_cuda:
        return _native().fedavg_reduce(x, counts, **kw)
    return reference.fedavg_reduce(x, counts, **kw)


# -- K-BCAST ------------------------------------------------------------------
def bcast_global(master: torch.Tensor, slots=None, slots_lowp=None, momentum=None) -> None:
    if master.is_cuda:
        return _native().bcast_global(master, slots, slots_lowp, momentum)
    return reference.bcast_global(master, slots, slots_lowp, momentum)


# -- K-OPT ------------------------------------------------------------------
def fused_sgd_(p: torch.Tensor, g: torch.Tensor, lr: float, **kw) -> torch.Tensor:
    if p.is_cuda:
        return _native().fused_sgd_(p, g, lr, **kw)
    return reference.fused_sgd_(p, g, lr, **kw)


def fused_adam_(p, g, exp_avg, exp_avg_sq, step: int, lr: float, **kw) -> torch.Tensor:
    if p.is_cuda:
        return _native().fused_adam_(p, g, exp_avg, exp_avg_sq, step, lr, **kw)
    return reference.fused_adam_(p, g, exp_avg, exp_avg_sq, step, lr, **kw)


# -- K-XENT ------------------------------------------------------------------
def xent_fused(logits: torch.Tensor, targets: torch.Tensor, **kw):
    if logits.is_cuda:
        return _native().xent_fused(logits, targets, **kw)
    return reference.xent_fused(logits, targets, **kw)


# -- K-PSGN ------------------------------------------------------------------
def linear_per_sample_grad_sqnorms(activations, grad_outputs, bias: bool = True) -> torch.Tensor:
    if activations.is_cuda:
        return _native().linear_psgn(activations, grad_outputs, bias)
    return reference.linear_per_sample_grad_sqnorms(activations, grad_outputs, bias)


def conv_per_sample_grad_sqnorms(unfolded, grad_outputs, bias: bool = True) -> torch.Tensor:
    """``[B]`` squared per-sample weight-gradient norms of a convolution (or a Linear applied to
    sequences).  CUDA: grouped tcgen05 GEMM with a sum-of-squares epilogue (``ops.tc.conv_psgn``,
    one group per sample, gradients never leave TMEM); CPU: the Gram-matrix oracle."""
    if unfolded.is_cuda and unfolded.shape[0] <= 65535:
        from . import tc

        return tc.conv_psgn(unfolded, grad_outputs, bias)
    return reference.conv_per_sample_grad_sqnorms(unfolded, grad_outputs, bias)


# -- K-MASK ------------------------------------------------------------------
def mask_add_(x: torch.Tensor, cumulative: torch.Tensor, seed: int, offset: int) -> torch.Tensor:
    if x.is_cuda:
        return _native().mask_add_(x, cumulative, seed, offset)
    return reference.mask_add_(x, cumulative, seed, offset)

"""Pure-PyTorch definition of every op (= numerical oracle and CPU path).

Each function documents the math ONCE; ``csrc/*.cu`` implements the same
contract for sm_100a and ``tests/`` compare the two.  Nothing here is tuned.
"""

from __future__ import annotations

import math
from typing import Sequence

import torch

from . import philox

_CLIP_EPS = 1e-6  # reference ``mechanisms.py:93`` / torch ``clip_grad_norm_``


# ---------------------------------------------------------------------------
# K-RNG
# ---------------------------------------------------------------------------
def philox_normal_(out: torch.Tensor, scale: float, seed: int, offset: int = 0) -> torch.Tensor:
    z = philox.normal(out.numel(), seed, offset).reshape(out.shape)
    out.copy_((z * scale).to(out.dtype))
    return out


def philox_laplace_(out: torch.Tensor, scale: float, seed: int, offset: int = 0) -> torch.Tensor:
    z = philox.laplace(out.numel(), seed, offset).reshape(out.shape)
    out.copy_((z * scale).to(out.dtype))
    return out


def philox_uniform_(out: torch.Tensor, seed: int, offset: int = 0) -> torch.Tensor:
    out.copy_(philox.uniform(out.numel(), seed, offset).reshape(out.shape).to(out.dtype))
    return out


# ---------------------------------------------------------------------------
# K-STATS
# ---------------------------------------------------------------------------
def update_stats(x: torch.Tensor) -> tuple[torch.Tensor, torch.Tensor]:
    """Per-row ``Σ x²`` (fp32) and a non-finite flag (int32) of a ``[K, N]`` slab."""
    xf = x.to(torch.float32)
    bad = (~torch.isfinite(xf)).any(dim=1).to(torch.int32)
    return xf.pow(2).sum(dim=1), bad


def sumsq_multi(tensors: Sequence[torch.Tensor]) -> torch.Tensor:
    """``Σ_t Σ_i t[i]²`` over a list of tensors, 0-d fp32 on the tensors' device."""
    total = torch.zeros((), dtype=torch.float32, device=tensors[0].device)
    for t in tensors:
        total = total + t.detach().to(torch.float32).pow(2).sum()
    return total


def per_tensor_stats(tensors: Sequence[torch.Tensor]) -> tuple[torch.Tensor, torch.Tensor]:
    dev = tensors[0].device
    ss = torch.stack([t.detach().to(torch.float32).pow(2).sum() for t in tensors]).to(dev)
    bad = torch.stack([(~torch.isfinite(t.detach().to(torch.float32))).any() for t in tensors])
    return ss, bad.to(torch.int32)


def segment_sumsq(x: torch.Tensor, segments: torch.Tensor) -> torch.Tensor:
    """``[K, S]``: ``Σ x[k, off_s : off_s+len_s]²`` for each (offset, length) row of ``segments``."""
    xf = x.to(torch.float32)
    cols = [xf[:, int(o) : int(o) + int(n)].pow(2).sum(dim=1) for o, n in segments.tolist()]
    return torch.stack(cols, dim=1) if cols else xf.new_zeros(x.shape[0], 0)


# ---------------------------------------------------------------------------
# K-AGG
# ---------------------------------------------------------------------------
def agg_coefficients(
    counts: torch.Tensor,
    sumsq: torch.Tensor | None,
    mask: torch.Tensor | None,
    clip_norm: float,
    normalize: bool,
) -> tuple[torch.Tensor, torch.Tensor]:
    """``(coef[K], Σ w²)``: per-client multiplier ``w_k·c_k`` and the noise-folding factor."""
    eff = counts.to(torch.float32)
    if mask is not None:
        eff = eff * mask.to(torch.float32)
    w = eff / eff.sum() if normalize else eff
    coef = w
    if clip_norm > 0 and sumsq is not None:
        c = (clip_norm / (sumsq.to(torch.float32).sqrt() + _CLIP_EPS)).clamp(max=1.0)
        # an absent client (w == 0) may carry NaN statistics; it must not poison the sum
        coef = torch.where(w > 0, w * torch.nan_to_num(c, nan=1.0), torch.zeros_like(w))
    return coef, (w * w).sum()


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
    """``out[j] (+)= Σ_k w_k c_k x[k,j] + noise_std·sqrt(Σ w_k²)·z[j]·noise_mask[j]``.

    ``w_k = counts_k·mask_k / Σ(counts·mask)`` (or un-normalised), ``c_k`` the DP clip
    factor from ``sumsq`` when ``clip_norm > 0``, ``z`` = ``unit_noise`` if given else
    the Philox normal stream at ``(seed, offset)``.  Accumulation in fp32.
    """
    k, n = x.shape
    eff = counts.to(torch.float32) * (mask.to(torch.float32) if mask is not None else 1.0)
    if normalize and out is not None and not bool(eff.sum() > 0):
        if bcast is not None:
            bcast_global(out, *bcast)
        return out          # every client masked out: keep the previous model (as the kernels do)
    coef, w2 = agg_coefficients(counts, sumsq, mask, clip_norm, normalize)
    acc = (coef.view(k, 1).to(torch.float32) * x.to(torch.float32)).sum(dim=0)
    # Rows with coefficient 0 must not contribute even if they hold NaN/Inf.
    if bool((coef == 0).any()):
        keep = coef != 0
        acc = (coef[keep].view(-1, 1) * x[keep].to(torch.float32)).sum(dim=0)
    if noise_std > 0:
        z = unit_noise.to(torch.float32) if unit_noise is not None else philox.normal(
            n, seed, offset).to(x.device)
        if noise_mask is not None:
            z = z * noise_mask.to(torch.float32)
        acc = acc + (noise_std * torch.sqrt(w2)) * z
    if out is None:
        out = torch.zeros(n, dtype=torch.float32, device=x.device)
        accumulate = False
    if accumulate:
        out.add_(acc)
    else:
        out.copy_(acc)
    if out_lowp is not None:
        out_lowp.copy_(out.to(out_lowp.dtype))
    if bcast is not None:        # K-BCAST fused into the aggregation (native: same launch)
        bcast_global(out, *bcast)
    return out


# ---------------------------------------------------------------------------
# K-BCAST
# ---------------------------------------------------------------------------
def bcast_global(
    master: torch.Tensor,
    slots: torch.Tensor | None = None,
    slots_lowp: torch.Tensor | None = None,
    momentum: torch.Tensor | None = None,
) -> None:
    """Replicate the global vector into every client slot and reset optimiser state."""
    if slots is not None:
        slots.copy_(master.to(slots.dtype).unsqueeze(0).expand_as(slots))
    if slots_lowp is not None:
        slots_lowp.copy_(master.to(slots_lowp.dtype).unsqueeze(0).expand_as(slots_lowp))
    if momentum is not None:
        momentum.zero_()


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
    """torch.optim.SGD semantics on flat buffers, with optional DP-SGD pre-processing.

    ``sumsq_out`` ``[rows]`` (accumulated): Σ of the squares of the updated parameters per row,
    over the columns outside ``skip_range`` (K-STATS side output of the last local step).

    Cohort form: ``row_sumsq`` ``[rows]`` clips row ``r`` by ``min(1, clip_norm / (sqrt(ss_r) +
    1e-6))``, ``noise_mask`` ``[row_len]`` limits the noise to real parameter entries and
    ``step_counter`` shifts the noise stream by ``step·ceil(n/4)`` Philox groups.

    ``g' = g·grad_scale + noise_std·z`` (clip factor and Gaussian noise, both optional),
    then ``g' += wd·p``; momentum buffer ``b = g'`` on the first step else
    ``b = μ·b + (1-dampening)·g'``; ``p -= lr·(g' + μ·b if nesterov else b)``.
    ``skip_range=(lo, hi)``: columns ``[lo, hi)`` of every ``row_len`` row keep their values
    (parameters, momentum and low-precision copy): they were updated elsewhere.
    """
    keep = None
    if skip_range is not None:
        lo, hi = skip_range
        keep = [(t, t.view(-1, row_len)[:, lo:hi].clone())
                for t in (p, momentum_buf, lowp_out) if t is not None]
    gf = g.to(torch.float32)
    if grad_scale is not None:
        gf = gf * grad_scale.to(torch.float32)
    if row_sumsq is not None:
        c = torch.clamp(clip_norm / (row_sumsq.to(torch.float32).sqrt() + 1e-6), max=1.0)
        gf = (gf.view(-1, row_len) * c.view(-1, 1).to(gf.device)).reshape(g.shape)
    if noise_std > 0:
        shift = 0
        if step_counter is not None:
            shift = int(step_counter.reshape(-1)[0]) * ((g.numel() + 3) // 4) * 4
        z = philox.normal(g.numel(), seed, offset + shift).reshape(g.shape).to(g.device)
        if noise_mask is not None:
            z = (z.view(-1, row_len) * noise_mask.to(z.dtype).view(1, -1).to(z.device)).reshape(g.shape)
        gf = gf + noise_std * z
    if weight_decay != 0:
        gf = gf + weight_decay * p.to(torch.float32)
    if momentum != 0 and momentum_buf is not None:
        if first_step:
            momentum_buf.copy_(gf)
        else:
            momentum_buf.mul_(momentum).add_(gf, alpha=1 - dampening)
        gf = gf + momentum * momentum_buf if nesterov else momentum_buf.to(torch.float32)
    p.add_(gf.to(p.dtype), alpha=-lr)
    if lowp_out is not None:
        lowp_out.copy_(p.to(lowp_out.dtype))
    if zero_grad_prefix > 0:
        g.view(-1, row_len)[:, :zero_grad_prefix] = 0
    if zero_grad_from > 0:
        g.view(-1, row_len)[:, zero_grad_from:] = 0
    if keep is not None:
        for t, saved in keep:
            t.view(-1, row_len)[:, skip_range[0]:skip_range[1]] = saved
    if sumsq_out is not None:
        sq = p.to(torch.float32).view(-1, row_len) ** 2
        if skip_range is not None:
            sq[:, skip_range[0]:skip_range[1]] = 0
        sumsq_out.add_(sq.sum(1).to(sumsq_out.device))
    if perm_out is not None:
        # the parameter stored [rows][inner][taps] at columns lo.. is also written [rows][taps][inner]
        out_t, lo, rows, inner, taps = perm_out
        block = p.view(-1, row_len)[:, lo:lo + rows * inner * taps].view(-1, rows, inner, taps)
        flat = block.permute(0, 1, 3, 2).reshape(block.shape[0], -1)
        out_t.view(block.shape[0], -1)[:, :flat.shape[1]] = flat.to(out_t.dtype)
    return p


def fused_adam_(
    p: torch.Tensor,
    g: torch.Tensor,
    exp_avg: torch.Tensor,
    exp_avg_sq: torch.Tensor,
    step: int,
    lr: float,
    *,
    beta1: float = 0.9,
    beta2: float = 0.999,
    eps: float = 1e-8,
    weight_decay: float = 0.0,
    decoupled: bool = False,
    grad_scale: torch.Tensor | None = None,
    lowp_out: torch.Tensor | None = None,
) -> torch.Tensor:
    """torch.optim.Adam / AdamW (``decoupled=True``) on flat buffers; ``step`` is 1-based."""
    gf = g.to(torch.float32)
    if grad_scale is not None:
        gf = gf * grad_scale.to(torch.float32)
    if weight_decay != 0:
        if decoupled:
            p.mul_(1 - lr * weight_decay)
        else:
            gf = gf + weight_decay * p.to(torch.float32)
    exp_avg.mul_(beta1).add_(gf, alpha=1 - beta1)
    exp_avg_sq.mul_(beta2).addcmul_(gf, gf, value=1 - beta2)
    bc1 = 1 - beta1**step
    bc2 = 1 - beta2**step
    denom = (exp_avg_sq.sqrt() / math.sqrt(bc2)).add_(eps)
    p.addcdiv_(exp_avg, denom, value=-lr / bc1)
    if lowp_out is not None:
        lowp_out.copy_(p.to(lowp_out.dtype))
    return p


# ---------------------------------------------------------------------------
# K-XENT
# ---------------------------------------------------------------------------
def xent_fused(
    logits: torch.Tensor,
    targets: torch.Tensor,
    *,
    dlogits: torch.Tensor | None = None,
    stats: torch.Tensor | None = None,
    grad_scale: float | None = None,
) -> tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """Log-softmax + NLL + accuracy + gradient in one pass over ``[..., B, C]`` logits.

    Returns ``(loss_sum, correct, dlogits)`` per leading group (``[...]``);
    ``dlogits = (softmax − onehot)·grad_scale`` with ``grad_scale`` defaulting to
    ``1/B`` (mean reduction).  ``stats[..., 0:3] += (loss_sum, correct, B)`` if given.
    """
    lf = logits.to(torch.float32)
    b, c = lf.shape[-2], lf.shape[-1]
    logp = torch.log_softmax(lf, dim=-1)
    t = targets.to(torch.int64)
    picked = logp.gather(-1, t.unsqueeze(-1)).squeeze(-1)
    loss_sum = -picked.sum(dim=-1)
    correct = (lf.argmax(dim=-1) == t).sum(dim=-1).to(torch.float32)
    scale = (1.0 / b) if grad_scale is None else grad_scale
    grad = logp.exp()
    grad.scatter_add_(-1, t.unsqueeze(-1), -torch.ones_like(picked).unsqueeze(-1))
    grad = grad * scale
    if dlogits is None:
        dlogits = grad.to(logits.dtype)
    else:
        dlogits.copy_(grad.to(dlogits.dtype))
    if stats is not None:
        stats[..., 0] += loss_sum
        stats[..., 1] += correct
        stats[..., 2] += float(b)
    return loss_sum, correct, dlogits


# ---------------------------------------------------------------------------
# K-PSGN
# ---------------------------------------------------------------------------
def linear_per_sample_grad_sqnorms(
    activations: torch.Tensor, grad_outputs: torch.Tensor, bias: bool = True
) -> torch.Tensor:
    """Ghost norms: ``‖∂L_i/∂W‖² = ‖a_i‖²·‖δ_i‖²`` (+ ``‖δ_i‖²`` for the bias), ``[B]``."""
    a2 = activations.to(torch.float32).flatten(1).pow(2).sum(dim=1)
    d2 = grad_outputs.to(torch.float32).flatten(1).pow(2).sum(dim=1)
    return a2 * d2 + (d2 if bias else 0.0)


def conv_per_sample_grad_sqnorms(
    unfolded: torch.Tensor, grad_outputs: torch.Tensor, bias: bool = True
) -> torch.Tensor:
    """Per-sample conv weight-gradient norms from unfolded inputs.

    ``unfolded``: ``[B, P, Cin·kh·kw]`` (im2col patches), ``grad_outputs``: ``[B, P, Cout]``.
    ``∂L_i/∂W = δ_iᵀ U_i`` so ``‖·‖_F² = ⟨U_i U_iᵀ, δ_i δ_iᵀ⟩`` — two ``P×P`` Gram
    matrices instead of the ``Cout × Cin·kh·kw`` gradient per sample.
    """
    u = unfolded.to(torch.float32)
    d = grad_outputs.to(torch.float32)
    gu = u @ u.transpose(1, 2)
    gd = d @ d.transpose(1, 2)
    out = (gu * gd).sum(dim=(1, 2))
    if bias:
        out = out + d.sum(dim=1).pow(2).sum(dim=1)
    return out


# ---------------------------------------------------------------------------
# K-MASK
# ---------------------------------------------------------------------------
def mask_add_(x: torch.Tensor, cumulative: torch.Tensor, seed: int, offset: int) -> torch.Tensor:
    """``x += m`` and ``cumulative += m`` with ``m ~ U(0,1)`` from the Philox stream."""
    m = philox.uniform(x.numel(), seed, offset).reshape(x.shape).to(x.device)
    x.add_(m.to(x.dtype))
    cumulative.add_(m.to(cumulative.dtype))
    return x

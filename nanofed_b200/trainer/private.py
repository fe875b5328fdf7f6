"""DP-SGD trainer: batch-gradient clipping + Gaussian noise + accounting.

Parity: reference ``trainer/private.py:16-154`` — constructor signature,
``_clip_gradients`` (global-norm clip of the **batch** gradient to
``max_gradient_norm``), ``_add_noise`` (per-parameter ``N(0, (σC)²)`` + one
accountant event ``(σ, batch_size)``), ``_apply_privacy_to_gradients``,
``train_batch`` (returns ``samples_processed == B``, ``epoch = batch = 0``),
``get_privacy_spent`` / ``validate_privacy_budget``
(``test_private_trainer.py:36-191``).

Fixes: ``callbacks`` are forwarded (the reference drops them,
``private.py:49``) and ``train_epoch`` is private too — every step of the
inherited loop goes through clip+noise (the reference's ``train_epoch`` silently
trains without privacy).

B200 path: on CUDA gradients the clip needs no host round-trip — the global norm
is one fused multi-tensor sum-of-squares launch (K-STATS), the coefficient stays
a device scalar, and noise comes from the Philox kernel when the generator lives
on the same device.  ``ops.per_sample_grad_sqnorms`` (K-PSGN) offers true
per-sample norms for linear/conv layers as an extension beyond the reference.
"""

from __future__ import annotations

from typing import Sequence

import torch
import torch.nn as nn

from nanofed_b200.privacy.accountant import GaussianAccountant, PrivacySpent
from nanofed_b200.privacy.config import PrivacyConfig
from nanofed_b200.privacy.noise import GaussianNoiseGenerator

from .base import Callback, TrainingConfig, TrainingMetrics
from .torch import TorchTrainer


class PrivateTrainer(TorchTrainer):
    def __init__(
        self,
        training_config: TrainingConfig,
        privacy_config: PrivacyConfig,
        accountant: GaussianAccountant | None = None,
        noise_generator: GaussianNoiseGenerator | None = None,
        callbacks: list[Callback] | None = None,
        per_sample_clipping: bool = False,
    ) -> None:
        super().__init__(training_config, callbacks)
        self._per_sample = per_sample_clipping
        self._privacy_config = privacy_config
        self._accountant = accountant or GaussianAccountant(privacy_config)
        self._noise_gen = noise_generator or GaussianNoiseGenerator()

    # -- privacy primitives --------------------------------------------------
    def _clip_gradients(self, model: nn.Module) -> None:
        bound = self._privacy_config.max_gradient_norm
        grads = [p.grad for p in model.parameters() if p.grad is not None]
        if not grads:
            return
        if all(g.is_cuda for g in grads):
            from nanofed_b200 import ops

            norm = ops.sumsq_multi(grads).sqrt()
            coef = (bound / (norm + 1e-6)).clamp(max=1.0)
            torch._foreach_mul_(grads, coef)
        else:
            torch.nn.utils.clip_grad_norm_(model.parameters(), bound)

    def _add_noise(self, model: nn.Module, batch_size: int) -> None:
        sigma = self._privacy_config.noise_multiplier
        std = sigma * self._privacy_config.max_gradient_norm
        for p in model.parameters():
            if p.grad is None:
                continue
            draw = self._noise_gen.generate(tuple(p.grad.shape), scale=std)
            if draw.device != p.grad.device:
                draw = draw.to(p.grad.device, non_blocking=True)
            p.grad.add_(draw.to(p.grad.dtype))
        self._accountant.add_noise_event(sigma=sigma, samples=batch_size)

    def _apply_privacy_to_gradients(self, model: nn.Module, batch_size: int) -> None:
        self._clip_gradients(model)
        self._add_noise(model, batch_size)

    # -- per-sample clipping (north-star extension, K-PSGN) ---------------------------------
    def per_sample_grad_norms(
        self, model: nn.Module, data: torch.Tensor, target: torch.Tensor
    ) -> tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
        """``(norms [B], per-sample losses [B], output)`` WITHOUT materialising per-sample
        gradients ("ghost norms").  For ``nn.Linear`` ``‖∂L_i/∂W‖² = ‖a_i‖²‖δ_i‖²`` (hand-written
        kernel ``ops.linear_per_sample_grad_sqnorms``); for ``nn.Conv2d`` the Frobenius product of
        the unfolded-input and output-gradient Gram matrices.  Other parametrised layers are
        rejected (their per-sample gradients need the full Jacobian).  The autograd graph is
        kept so the caller can run the re-weighted second backward."""
        import torch.nn.functional as F

        from nanofed_b200 import ops

        acts: dict[nn.Module, torch.Tensor] = {}
        sq = torch.zeros(len(data), device=data.device, dtype=torch.float32)
        handles = []

        def fwd_hook(mod, inputs, _out):
            acts[mod] = inputs[0].detach()

        armed = {"on": True}   # hooks baked into the retained graph fire again on the 2nd backward

        def bwd_hook(mod, _gin, gout):
            nonlocal sq
            if not armed["on"] or mod not in acts:
                return
            a, d = acts.pop(mod), gout[0].detach()
            if isinstance(mod, nn.Linear):
                if a.dim() != 2:
                    a2, d2 = a.flatten(1, -2), d.flatten(1, -2)       # [B, T, in] sequences
                    sq = sq + ops.conv_per_sample_grad_sqnorms(a2, d2, bias=mod.bias is not None)
                else:
                    sq = sq + ops.linear_per_sample_grad_sqnorms(a, d, bias=mod.bias is not None)
            else:  # Conv2d
                if mod.groups != 1:
                    raise NotImplementedError("per-sample clipping: grouped convolutions")
                unf = F.unfold(a, mod.kernel_size, mod.dilation, mod.padding, mod.stride)
                sq = sq + ops.conv_per_sample_grad_sqnorms(
                    unf.transpose(1, 2), d.flatten(2).transpose(1, 2), bias=mod.bias is not None)

        for mod in model.modules():
            own = [p for p in mod.parameters(recurse=False) if p.requires_grad]
            if not own:
                continue
            if not isinstance(mod, (nn.Linear, nn.Conv2d)):
                raise NotImplementedError(
                    f"per-sample clipping supports Linear and Conv2d layers, found {type(mod).__name__}")
            handles.append(mod.register_forward_hook(fwd_hook))
            handles.append(mod.register_full_backward_hook(bwd_hook))
        try:
            output = model(data)
            losses = F.cross_entropy(output, target, reduction="none")
            losses.sum().backward(retain_graph=True)
        finally:
            armed["on"] = False
            for h in handles:
                h.remove()
        return sq.clamp_min(0).sqrt(), losses, output

    def _per_sample_step(self, model, data, target, optimizer):
        cfg = self._privacy_config
        optimizer.zero_grad(set_to_none=True)
        norms, losses, output = self.per_sample_grad_norms(model, data, target)
        coef = (cfg.max_gradient_norm / (norms + 1e-6)).clamp(max=1.0)
        optimizer.zero_grad(set_to_none=True)
        (losses * coef.detach()).sum().backward()            # Σ_i c_i ∇L_i, one more backward
        batch = len(data)
        self._add_noise(model, batch)                          # N(0, (σC)²) on the clipped SUM
        for p in model.parameters():
            if p.grad is not None:
                p.grad.div_(batch)
        optimizer.step()
        return losses.detach().mean(), output

    # -- training ------------------------------------------------------------
    def _step(self, model, data, target, optimizer):  # type: ignore[override]
        if self._per_sample:
            return self._per_sample_step(model, data, target, optimizer)
        optimizer.zero_grad(set_to_none=True)
        output = model(data)
        loss = self.compute_loss(output, target)
        loss.backward()
        self._apply_privacy_to_gradients(model, len(data))
        optimizer.step()
        return loss, output

    def train_batch(
        self,
        model: nn.Module,
        batch: Sequence[torch.Tensor],
        optimizer: torch.optim.Optimizer,
    ) -> TrainingMetrics:
        inputs, targets = batch
        loss, outputs = self._step(model, inputs, targets, optimizer)
        return TrainingMetrics(
            loss=float(loss.item()),
            accuracy=float(self.compute_accuracy(outputs, targets)),
            epoch=0,
            batch=0,
            samples_processed=len(inputs),
        )

    def get_privacy_spent(self) -> PrivacySpent:
        return self._accountant.get_privacy_spent()

    def validate_privacy_budget(self) -> bool:
        return self._accountant.validate_budget()

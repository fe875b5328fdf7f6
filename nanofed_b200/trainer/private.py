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
    ) -> None:
        super().__init__(training_config, callbacks)
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

    # -- training ------------------------------------------------------------
    def _step(self, model, data, target, optimizer):  # type: ignore[override]
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

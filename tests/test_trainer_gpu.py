"""Trainer paths on a CUDA device: the fused cross-entropy/metrics kernel, DP-SGD noise from the
Philox kernel and per-sample clipping through K-PSGN."""

import pytest
import torch
import torch.nn as nn
from torch.utils.data import DataLoader, TensorDataset

from nanofed_b200.privacy import PrivacyConfig
from nanofed_b200.trainer import PrivateTrainer, TorchTrainer, TrainingConfig

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _small_cnn():
    return nn.Sequential(nn.Conv2d(1, 4, 3), nn.ReLU(), nn.Flatten(), nn.Linear(4 * 6 * 6, 5))


def test_per_sample_norms_on_cuda_use_the_native_kernel():
    from nanofed_b200.ops import native

    torch.manual_seed(0)
    model = _small_cnn().to(DEV)
    x, y = torch.randn(6, 1, 8, 8, device=DEV), torch.randint(0, 5, (6,), device=DEV)
    t = PrivateTrainer(TrainingConfig(1, 6, 0.1, device=DEV),
                       PrivacyConfig(max_gradient_norm=0.7, noise_multiplier=1e-9),
                       per_sample_clipping=True)
    before = native.LAUNCHES
    norms, _, _ = t.per_sample_grad_norms(model, x, y)
    assert native.LAUNCHES > before, "K-PSGN was not launched for the Linear layer"
    explicit = []
    for i in range(6):
        model.zero_grad()
        nn.functional.cross_entropy(model(x[i:i + 1]), y[i:i + 1]).backward()
        explicit.append(torch.sqrt(sum(p.grad.pow(2).sum() for p in model.parameters())))
    # conv norms come from bf16 tensor-core operands (fp32 accumulation): 1e-2
    torch.testing.assert_close(norms, torch.stack(explicit), rtol=1e-2, atol=1e-4)


def test_conv_psgn_kernel_matches_per_sample_gradients():
    """K-PSGN for convolutions (grouped tcgen05 GEMM, sum-of-squares epilogue) at the MNIST CNN's
    conv2 shape against explicit per-sample weight + bias gradients in fp32."""
    import torch.nn.functional as F

    from nanofed_b200 import ops
    from nanofed_b200.ops import native, tc

    torch.manual_seed(2)
    b, cin, cout, hw = 16, 32, 64, 26
    x = torch.randn(b, cin, hw, hw, device=DEV)
    dy = torch.randn(b, cout, hw - 2, hw - 2, device=DEV)
    unf = F.unfold(x, 3).transpose(1, 2)
    before = native.LAUNCHES
    got = ops.conv_per_sample_grad_sqnorms(unf, dy.flatten(2).transpose(1, 2), bias=True)
    torch.cuda.synchronize()
    tc.check_status(x.device)
    assert native.LAUNCHES > before
    w = torch.zeros(cout, cin, 3, 3, device=DEV, requires_grad=True)
    bias = torch.zeros(cout, device=DEV, requires_grad=True)
    want = []
    for i in range(b):
        gw, gb = torch.autograd.grad((F.conv2d(x[i:i + 1], w, bias) * dy[i:i + 1]).sum(), (w, bias))
        want.append(gw.pow(2).sum() + gb.pow(2).sum())
    torch.testing.assert_close(got, torch.stack(want), rtol=1e-2, atol=0)
    # odd sizes: K = 9 (one input channel), Cout = 4, P = 36: padding to 16-byte rows is internal
    x1, d1 = torch.randn(5, 1, 8, 8, device=DEV), torch.randn(5, 4, 6, 6, device=DEV)
    got1 = ops.conv_per_sample_grad_sqnorms(F.unfold(x1, 3).transpose(1, 2),
                                            d1.flatten(2).transpose(1, 2), bias=False)
    from nanofed_b200.ops import reference

    want1 = reference.conv_per_sample_grad_sqnorms(F.unfold(x1, 3).transpose(1, 2),
                                                   d1.flatten(2).transpose(1, 2), bias=False)
    torch.testing.assert_close(got1, want1, rtol=1e-2, atol=0)


def test_per_sample_step_applies_clipped_mean():
    torch.manual_seed(1)
    model = _small_cnn().to(DEV)
    x, y = torch.randn(6, 1, 8, 8, device=DEV), torch.randint(0, 5, (6,), device=DEV)
    clip = 0.5
    want = [torch.zeros_like(p) for p in model.parameters()]
    for i in range(6):
        model.zero_grad()
        nn.functional.cross_entropy(model(x[i:i + 1]), y[i:i + 1]).backward()
        g = [p.grad.clone() for p in model.parameters()]
        n = float(torch.sqrt(sum(v.pow(2).sum() for v in g)))
        for acc, v in zip(want, g):
            acc += min(clip / (n + 1e-6), 1.0) * v
    t = PrivateTrainer(TrainingConfig(1, 6, 0.1, device=DEV),
                       PrivacyConfig(max_gradient_norm=clip, noise_multiplier=1e-9),
                       per_sample_clipping=True)
    before = [p.detach().clone() for p in model.parameters()]
    opt = torch.optim.SGD(model.parameters(), lr=1.0)
    model.zero_grad()
    t.train_batch(model, (x, y), opt)
    for p, b0, g in zip(model.parameters(), before, want):
        torch.testing.assert_close(b0 - p.detach(), g / 6, rtol=1.5e-2, atol=1e-4)


def test_torch_trainer_epoch_on_cuda_learns_and_reports_device_metrics():
    torch.manual_seed(2)
    x = torch.randn(256, 1, 8, 8)
    y = (x.flatten(1).sum(1) > 0).long()
    loader = DataLoader(TensorDataset(x, y), batch_size=32)
    model = _small_cnn().to(DEV)
    opt = torch.optim.SGD(model.parameters(), lr=0.1)
    t = TorchTrainer(TrainingConfig(epochs=1, batch_size=32, learning_rate=0.1, device=DEV))
    first = None
    for epoch in range(6):
        t.train_epoch(model, loader, opt, epoch)
        if first is None:
            first = t.last_epoch_metrics.loss
    assert t.last_epoch_metrics.loss < first
    assert t.last_epoch_metrics.samples_processed == 256
    assert 0.0 <= t.last_epoch_metrics.accuracy <= 1.0


def test_private_trainer_on_cuda_draws_noise_from_the_philox_kernel():
    from nanofed_b200.ops import native

    torch.manual_seed(3)
    x, y = torch.randn(64, 1, 8, 8), torch.randint(0, 5, (64,))
    loader = DataLoader(TensorDataset(x, y), batch_size=32)
    model = _small_cnn().to(DEV)
    opt = torch.optim.SGD(model.parameters(), lr=0.05)
    t = PrivateTrainer(TrainingConfig(epochs=1, batch_size=32, learning_rate=0.05, device=DEV),
                       PrivacyConfig(max_gradient_norm=1.0, noise_multiplier=1.1))
    before = native.LAUNCHES
    t.train_epoch(model, loader, opt, 0)
    assert native.LAUNCHES > before
    spent = t.get_privacy_spent()
    assert spent.epsilon_spent > 0 and t.validate_privacy_budget() in (True, False)
    assert all(torch.isfinite(p).all() for p in model.parameters())

"""CPU oracle self-checks (closed forms / torch cross-checks)."""

import math

import torch

from nanofed_b200 import ops
from nanofed_b200.ops import philox, reference as ref


def test_philox_known_answer_and_offsets():
    import numpy as np

    w = philox.philox4x32(np.array([0], dtype=np.uint64), np.array([0], dtype=np.uint64), 0)
    assert [hex(v) for v in w[0]] == ["0x6627e8d5", "0xe169c58d", "0xbc57ac4c", "0x9b00dbd8"]
    full = philox.normal(64, 7, 0)
    assert torch.equal(philox.normal(16, 7, 8), full[8:24])


def test_fedavg_reduce_reference_matches_definition():
    k, n = 5, 256
    x = torch.randn(k, n)
    counts = torch.tensor([1.0, 2, 3, 4, 5])
    out = ops.fedavg_reduce(x, counts)
    want = (counts / counts.sum()).view(-1, 1).mul(x).sum(0)
    torch.testing.assert_close(out, want)
    ss, _ = ops.update_stats(x)
    out = ops.fedavg_reduce(x, counts, sumsq=ss, clip_norm=0.5)
    c = (0.5 / (ss.sqrt() + 1e-6)).clamp(max=1)
    torch.testing.assert_close(out, ((counts / counts.sum()) * c).view(-1, 1).mul(x).sum(0))


def test_folded_noise_has_right_std():
    k, n = 4, 200_000
    x = torch.zeros(k, n)
    counts = torch.tensor([1.0, 1, 1, 1])
    out = ops.fedavg_reduce(x, counts, noise_std=2.0, seed=3)
    assert abs(float(out.std()) - 2.0 * math.sqrt(4 * 0.25**2)) < 0.01


def test_sgd_and_adam_reference_match_torch():
    p0 = torch.randn(300)
    pt = torch.nn.Parameter(p0.clone())
    opt = torch.optim.SGD([pt], lr=0.05, momentum=0.9, weight_decay=1e-2)
    p, buf = p0.clone(), torch.zeros(300)
    for s in range(4):
        g = torch.randn(300)
        pt.grad = g.clone()
        opt.step()
        ops.fused_sgd_(p, g, 0.05, momentum_buf=buf, momentum=0.9, weight_decay=1e-2,
                       first_step=(s == 0))
    torch.testing.assert_close(p, pt.data)
    pt = torch.nn.Parameter(p0.clone())
    opt = torch.optim.Adam([pt], lr=1e-2)
    p, m, v = p0.clone(), torch.zeros(300), torch.zeros(300)
    for s in range(1, 4):
        g = torch.randn(300)
        pt.grad = g.clone()
        opt.step()
        ops.fused_adam_(p, g, m, v, s, 1e-2)
    torch.testing.assert_close(p, pt.data, rtol=1e-5, atol=1e-6)


def test_xent_reference_matches_autograd():
    logits = torch.randn(3, 8, 10, requires_grad=True)
    t = torch.randint(0, 10, (3, 8))
    loss, hit, dl = ref.xent_fused(logits.detach(), t)
    want = sum(torch.nn.functional.cross_entropy(logits[i], t[i]) for i in range(3))
    want.backward()
    torch.testing.assert_close(dl, logits.grad, rtol=1e-5, atol=1e-6)
    assert hit.tolist() == [(logits[i].argmax(1) == t[i]).sum().item() for i in range(3)]


def test_conv_psgn_reference_matches_vmap():
    b, cin, cout, hw = 3, 2, 4, 6
    x = torch.randn(b, cin, hw, hw)
    w = torch.randn(cout, cin, 3, 3, requires_grad=True)
    per = []
    for i in range(b):
        y = torch.nn.functional.conv2d(x[i : i + 1], w)
        (g,) = torch.autograd.grad(y.pow(2).sum() / 2, w)
        per.append(g.pow(2).sum())
    y = torch.nn.functional.conv2d(x, w).detach()
    unf = torch.nn.functional.unfold(x, 3).transpose(1, 2)
    dy = y.flatten(2).transpose(1, 2)
    got = ops.conv_per_sample_grad_sqnorms(unf, dy, bias=False)
    torch.testing.assert_close(got, torch.stack(per), rtol=1e-4, atol=1e-4)


def test_fused_sgd_skip_range_leaves_columns_alone():
    from nanofed_b200.ops import reference as ref

    rows, row_len = 3, 32
    p = torch.arange(rows * row_len, dtype=torch.float32)
    g = torch.ones(rows * row_len)
    lowp = torch.zeros(rows * row_len, dtype=torch.bfloat16)
    p0 = p.clone()
    ref.fused_sgd_(p, g, 0.5, lowp_out=lowp, row_len=row_len, zero_grad_prefix=4, skip_range=(8, 24))
    pv, p0v = p.view(rows, row_len), p0.view(rows, row_len)
    assert torch.equal(pv[:, 8:24], p0v[:, 8:24])
    assert torch.equal(pv[:, :8], p0v[:, :8] - 0.5) and torch.equal(pv[:, 24:], p0v[:, 24:] - 0.5)
    assert float(lowp.view(rows, row_len)[:, 8:24].abs().sum()) == 0.0
    assert float(g.view(rows, row_len)[:, :4].abs().sum()) == 0.0


def test_fused_sgd_permuted_side_output():
    from nanofed_b200.ops import reference as ref

    rows, row_len, lo, r, inner, taps = 2, 64, 8, 2, 3, 4      # parameter [2][3][4] at column 8
    p = torch.arange(rows * row_len, dtype=torch.float32)
    g = torch.zeros(rows * row_len)
    out = torch.zeros(rows, r * inner * taps, dtype=torch.bfloat16)
    ref.fused_sgd_(p, g, 0.1, row_len=row_len, perm_out=(out, lo, r, inner, taps))
    src = p.view(rows, row_len)[:, lo:lo + r * inner * taps].view(rows, r, inner, taps)
    want = src.permute(0, 1, 3, 2).reshape(rows, -1)
    assert torch.equal(out.float(), want)

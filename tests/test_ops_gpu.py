"""sm_100a kernels vs the pure-PyTorch fp32 oracle (``ops.reference``)."""

import math

import pytest
import torch

from nanofed_b200 import ops
from nanofed_b200.ops import reference as ref

pytestmark = pytest.mark.gpu
DEV = "cuda"


def test_native_library_is_loaded():
    from nanofed_b200.ops import native

    assert native.lib().nfb_version() >= 100


@pytest.mark.parametrize("kind", ["normal", "laplace", "uniform"])
def test_philox_matches_host_replica(kind):
    n, seed, off = 100_003, 1234567, 64
    out = torch.empty(n, device=DEV)
    if kind == "normal":
        ops.philox_normal_(out, 2.0, seed, off)
        want = ops.philox.normal(n, seed, off) * 2.0
    elif kind == "laplace":
        ops.philox_laplace_(out, 0.5, seed, off)
        want = ops.philox.laplace(n, seed, off) * 0.5
    else:
        ops.philox_uniform_(out, seed, off)
        want = ops.philox.uniform(n, seed, off)
    torch.testing.assert_close(out.cpu(), want, rtol=2e-4, atol=2e-5)


def test_philox_statistics():
    z = torch.empty(1 << 20, device=DEV)
    ops.philox_normal_(z, 3.0, 99, 0)
    assert abs(float(z.mean())) < 0.02 and abs(float(z.std()) - 3.0) < 0.02
    ops.philox_laplace_(z, 2.0, 99, 0)
    assert abs(float(z.abs().mean()) - 2.0) < 0.02


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("k,n", [(1, 64), (3, 1024), (32, 1_199_936), (40, 8_000)])
def test_update_stats(dtype, k, n):
    x = (torch.randn(k, n, device=DEV) * 0.1).to(dtype)
    ss, bad = ops.update_stats(x)
    want, _ = ref.update_stats(x.float().cpu())
    torch.testing.assert_close(ss.cpu(), want, rtol=1e-4, atol=1e-6)
    assert int(bad.sum()) == 0
    # deterministic: two launches, identical bits
    ss2, _ = ops.update_stats(x)
    assert torch.equal(ss, ss2)
    x[k - 1, n // 2] = float("nan")
    _, bad = ops.update_stats(x)
    assert bad.tolist() == [0] * (k - 1) + [1]


def test_sumsq_multi_and_per_tensor():
    ts = [torch.randn(s, device=DEV) for s in [(5,), (128, 9216), (10,), (3, 3, 7), ()]]
    total = ops.sumsq_multi(ts)
    want = sum(float(t.double().pow(2).sum()) for t in ts)
    assert math.isclose(float(total), want, rel_tol=1e-5)
    ss, bad = ops.per_tensor_stats([t for t in ts if t.numel()])
    assert ss.shape[0] == 5 and int(bad.sum()) == 0


def test_segment_sumsq():
    x = torch.randn(4, 4096, device=DEV)
    segs = torch.tensor([[0, 100], [128, 1000], [2048, 2048], [5, 1]])
    got = ops.segment_sumsq(x, segs)
    want = ref.segment_sumsq(x.cpu(), segs)
    torch.testing.assert_close(got.cpu(), want, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("k,n", [(1, 64), (3, 640), (7, 4096), (32, 1_199_936), (19, 50_048)])
def test_fedavg_reduce_plain(dtype, k, n):
    x = torch.randn(k, n, device=DEV).to(dtype)
    counts = torch.randint(1, 1000, (k,), device=DEV).float()
    out = ops.fedavg_reduce(x, counts)
    want = ref.fedavg_reduce(x.float().cpu(), counts.cpu())
    torch.testing.assert_close(out.cpu(), want, rtol=1e-5, atol=1e-5)


def test_fedavg_reduce_full_dp_path():
    k, n = 9, 65_536
    x = torch.randn(k, n, device=DEV) * 0.05
    counts = torch.arange(1, k + 1, device=DEV).float()
    mask = torch.ones(k, device=DEV)
    mask[3] = 0
    x[3] = float("nan")  # masked-out client may hold garbage
    ss, _ = ops.update_stats(x)
    noise_mask = (torch.rand(n, device=DEV) > 0.1).to(torch.uint8)
    lowp = torch.empty(n, device=DEV, dtype=torch.bfloat16)
    kw = dict(sumsq=ss, mask=mask, clip_norm=1.0, noise_std=0.3, noise_mask=noise_mask,
              seed=42, offset=128)
    out = ops.fedavg_reduce(x, counts, out_lowp=lowp, **kw)
    xc = x.cpu().clone()
    want = ref.fedavg_reduce(xc, counts.cpu(), sumsq=ss.cpu(), mask=mask.cpu(), clip_norm=1.0,
                             noise_std=0.3, noise_mask=noise_mask.cpu(), seed=42, offset=128)
    torch.testing.assert_close(out.cpu(), want, rtol=1e-4, atol=2e-5)
    torch.testing.assert_close(lowp.float().cpu(), want.bfloat16().float(), rtol=1e-2, atol=1e-2)
    # injected noise: exact algebra
    z = torch.randn(n, device=DEV)
    out2 = ops.fedavg_reduce(x, counts, unit_noise=z, **kw)
    want2 = ref.fedavg_reduce(xc, counts.cpu(), sumsq=ss.cpu(), mask=mask.cpu(), clip_norm=1.0,
                              noise_std=0.3, noise_mask=noise_mask.cpu(), unit_noise=z.cpu())
    torch.testing.assert_close(out2.cpu(), want2, rtol=1e-5, atol=1e-5)


def test_fedavg_reduce_accumulate_waves():
    k, n = 8, 4096
    x = torch.randn(k, n, device=DEV)
    c = torch.rand(k, device=DEV) + 0.5
    acc = torch.zeros(n, device=DEV)
    ops.fedavg_reduce(x[:4], c[:4], out=acc, normalize=False, accumulate=True)
    ops.fedavg_reduce(x[4:], c[4:], out=acc, normalize=False, accumulate=True)
    want = (c.view(-1, 1) * x).sum(0)
    torch.testing.assert_close(acc, want, rtol=1e-5, atol=1e-5)


def test_bcast_global():
    n, k = 8192, 5
    g = torch.randn(n, device=DEV)
    slots = torch.empty(k, n, device=DEV)
    lowp = torch.empty(k, n, device=DEV, dtype=torch.bfloat16)
    mom = torch.ones(k, n, device=DEV)
    ops.bcast_global(g, slots, lowp, mom)
    assert torch.equal(slots, g.expand(k, n))
    assert torch.equal(lowp, g.bfloat16().expand(k, n))
    assert float(mom.abs().sum()) == 0.0


@pytest.mark.parametrize("momentum,nesterov,wd", [(0.0, False, 0.0), (0.9, False, 1e-3),
                                                  (0.9, True, 0.0)])
def test_fused_sgd_matches_torch(momentum, nesterov, wd):
    n = 10_007
    p0 = torch.randn(n, device=DEV)
    pt = torch.nn.Parameter(p0.clone())
    opt = torch.optim.SGD([pt], lr=0.1, momentum=momentum, nesterov=nesterov, weight_decay=wd)
    p = torch.empty(n + 1, device=DEV)[:n]  # keep 16B alignment of base
    p = p0.clone()
    buf = torch.zeros(n, device=DEV)
    lowp = torch.empty(n, device=DEV, dtype=torch.bfloat16)
    for step in range(3):
        g = torch.randn(n, device=DEV)
        pt.grad = g.clone()
        opt.step()
        ops.fused_sgd_(p, g, 0.1, momentum_buf=buf, momentum=momentum, nesterov=nesterov,
                       weight_decay=wd, first_step=(step == 0), lowp_out=lowp)
    torch.testing.assert_close(p, pt.data, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(lowp.float(), p.bfloat16().float())


def test_fused_sgd_dp_front_end():
    n = 4096
    p = torch.zeros(n, device=DEV)
    g = torch.ones(n, device=DEV)
    scale = torch.tensor(0.5, device=DEV)
    ops.fused_sgd_(p, g, 1.0, grad_scale=scale, noise_std=2.0, seed=5, offset=0)
    want = -(0.5 + 2.0 * ops.philox.normal(n, 5, 0))
    torch.testing.assert_close(p.cpu(), want, rtol=1e-4, atol=1e-4)


def test_fused_sgd_skips_a_column_range_and_zeroes_the_atomic_prefix():
    rows, row_len, lo, hi, prefix = 5, 1024, 128, 896, 64
    p = torch.randn(rows * row_len, device=DEV)
    g = torch.randn(rows * row_len, device=DEV)
    lowp = torch.full((rows * row_len,), 7.0, device=DEV, dtype=torch.bfloat16)
    pr, gr, lr_ = p.cpu().clone(), g.cpu().clone(), lowp.cpu().clone()
    kw = dict(weight_decay=0.01, row_len=row_len, zero_grad_prefix=prefix, skip_range=(lo, hi),
              zero_grad_from=960, noise_std=0.3, seed=11, offset=8)
    ops.fused_sgd_(p, g, 0.1, lowp_out=lowp, **kw)
    ref.fused_sgd_(pr, gr, 0.1, lowp_out=lr_, **kw)
    torch.testing.assert_close(p.cpu(), pr, rtol=1e-5, atol=1e-6)
    assert torch.equal(g.cpu(), gr)
    torch.testing.assert_close(lowp.cpu().float(), lr_.float(), rtol=0, atol=1e-2)
    assert float((lowp.view(rows, row_len)[:, lo:hi] - 7.0).abs().max()) == 0.0   # untouched


@pytest.mark.parametrize("rows", [5, 300])
def test_fused_sgd_sum_of_squares_side_output(rows):
    """K-STATS folded into K-OPT: per-row sum of the squared UPDATED parameters outside the skipped
    range, accumulated into the output (shared-memory path up to 256 rows, direct atomics above)."""
    row_len, lo, hi = 512, 128, 384
    p = torch.randn(rows * row_len, device=DEV)
    g = torch.randn(rows * row_len, device=DEV)
    out = torch.full((rows,), 2.0, device=DEV)
    pr, gr, outr = p.cpu().clone(), g.cpu().clone(), out.cpu().clone()
    kw = dict(weight_decay=0.01, row_len=row_len, skip_range=(lo, hi))
    ops.fused_sgd_(p, g, 0.1, sumsq_out=out, **kw)
    ref.fused_sgd_(pr, gr, 0.1, sumsq_out=outr, **kw)
    torch.testing.assert_close(p.cpu(), pr, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(out.cpu(), outr, rtol=1e-5, atol=1e-5)
    want = 2.0 + (p.view(rows, row_len).double() ** 2).sum(1) - (p.view(rows, row_len)[:, lo:hi].double() ** 2).sum(1)
    torch.testing.assert_close(out.double(), want, rtol=1e-5, atol=1e-5)


def test_fused_sgd_permuted_side_output_matches_oracle():
    rows, row_len = 3, 64 * 32 * 9 + 256
    lo = 128
    p = torch.randn(rows * row_len, device=DEV)
    g = torch.randn(rows * row_len, device=DEV)
    out = torch.zeros(rows, 64 * 288, device=DEV, dtype=torch.bfloat16)
    pr, gr, outr = p.cpu().clone(), g.cpu().clone(), out.cpu().clone()
    kw = dict(row_len=row_len, zero_grad_prefix=lo + 64 * 288, skip_range=(lo + 64 * 288, row_len - 64))
    ops.fused_sgd_(p, g, 0.1, perm_out=(out, lo, 64, 32, 9), **kw)
    ref.fused_sgd_(pr, gr, 0.1, perm_out=(outr, lo, 64, 32, 9), **kw)
    torch.testing.assert_close(p.cpu(), pr, rtol=1e-5, atol=1e-6)
    assert torch.equal(out.cpu(), outr)
    # and it is exactly what the stand-alone permutation kernel produces from the new weights
    w = p.view(rows, row_len)[:, lo:lo + 64 * 288].view(rows, 64, 32, 9)
    assert torch.equal(out.view(rows, 64, 9, 32), w.permute(0, 1, 3, 2).bfloat16())


def test_fused_adam_matches_torch():
    n = 5000
    p0 = torch.randn(n, device=DEV)
    pt = torch.nn.Parameter(p0.clone())
    opt = torch.optim.AdamW([pt], lr=1e-2, weight_decay=0.01)
    p, m, v = p0.clone(), torch.zeros(n, device=DEV), torch.zeros(n, device=DEV)
    for step in range(1, 4):
        g = torch.randn(n, device=DEV)
        pt.grad = g.clone()
        opt.step()
        ops.fused_adam_(p, g, m, v, step, 1e-2, weight_decay=0.01, decoupled=True)
    torch.testing.assert_close(p, pt.data, rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_xent_fused(dtype):
    g, b, c = 6, 64, 10
    logits = (torch.randn(g, b, c, device=DEV) * 3).to(dtype)
    t = torch.randint(0, c, (g, b), device=DEV)
    stats = torch.zeros(g, 3, device=DEV)
    loss, hit, dl = ops.xent_fused(logits, t, stats=stats)
    rl, rh, rd = ref.xent_fused(logits.float().cpu(), t.cpu())
    torch.testing.assert_close(loss.cpu(), rl, rtol=1e-4, atol=1e-3)
    torch.testing.assert_close(hit.cpu(), rh)
    tol = 1e-5 if dtype == torch.float32 else 1e-2
    torch.testing.assert_close(dl.float().cpu(), rd, rtol=tol, atol=tol)
    torch.testing.assert_close(stats[:, 2].cpu(), torch.full((g,), float(b)))
    # torch cross-check
    want = torch.nn.functional.cross_entropy(logits[0].float(), t[0], reduction="sum")
    assert math.isclose(float(loss[0]), float(want), rel_tol=1e-4)


def test_linear_psgn_matches_autograd():
    b, fin, fout = 16, 96, 24
    a = torch.randn(b, fin, device=DEV)
    d = torch.randn(b, fout, device=DEV)
    got = ops.linear_per_sample_grad_sqnorms(a, d, bias=True)
    per = torch.einsum("bi,bo->boi", a, d).flatten(1).pow(2).sum(1) + d.pow(2).sum(1)
    torch.testing.assert_close(got, per, rtol=1e-4, atol=1e-4)


def test_mask_add():
    x = torch.zeros(1001, device=DEV)
    cum = torch.zeros(1001, device=DEV)
    ops.mask_add_(x, cum, 77, 0)
    torch.testing.assert_close(x.cpu(), ops.philox.uniform(1001, 77, 0), rtol=1e-6, atol=1e-7)
    assert torch.equal(x, cum)


def test_cuda_tensor_without_library_is_loud(monkeypatch):
    from nanofed_b200.ops import native

    monkeypatch.setattr(native, "_lib", None)
    monkeypatch.setenv("NANOFED_B200_LIB", "/nonexistent/lib.so")
    with pytest.raises(native.NativeLibraryError):
        ops.update_stats(torch.zeros(1, 64, device=DEV))


def test_fedavg_reduce_with_fused_broadcast_equals_two_kernels():
    """K-AGG + K-BCAST in one launch: the client rows it reads are overwritten in place with the
    new global model (fp32 + bf16), momentum rows are zeroed; DP noise and masks included; with
    every client masked out the OLD model is what gets broadcast."""
    from nanofed_b200 import ops

    torch.manual_seed(4)
    k, n = 6, 70_016
    slots = torch.randn(k, n, device="cuda")
    counts = torch.randint(1, 90, (k,), device="cuda").float()
    mask = torch.ones(k, device="cuda")
    mask[2] = 0
    slots[2] = float("nan")                      # masked rows are never read
    kw = dict(mask=mask, noise_std=0.1, seed=9, offset=128)
    want = ops.fedavg_reduce(slots.clone(), counts, **kw)
    lowp = torch.zeros(k, n, dtype=torch.bfloat16, device="cuda")
    mom = torch.ones(k, n, device="cuda")
    master = torch.zeros(n, device="cuda")
    mlow = torch.zeros(n, dtype=torch.bfloat16, device="cuda")
    ops.fedavg_reduce(slots, counts, out=master, out_lowp=mlow, bcast=(slots, lowp, mom), **kw)
    torch.cuda.synchronize()
    assert torch.equal(master, want)
    assert all(torch.equal(slots[i], want) for i in range(k))
    assert all(torch.equal(lowp[i], want.bfloat16()) for i in range(k))
    assert float(mom.abs().sum()) == 0.0
    # all masked: the previous model is kept AND still reaches the rows
    slots.normal_()
    ops.fedavg_reduce(slots, counts, out=master, out_lowp=mlow, mask=torch.zeros(k, device="cuda"),
                      bcast=(slots, lowp, None))
    assert torch.equal(master, want) and torch.equal(slots[k - 1], want)


def test_simulation_round_has_no_broadcast_launch_after_the_first():
    from nanofed_b200.ops import native
    from nanofed_b200.parallel.cohort import CohortConfig, FedSimulation

    def run(fuse):
        sim = FedSimulation(CohortConfig(model="mnist_cnn", clients_per_gpu=4, samples_per_client=128,
                                         batch_size=64, device="cuda", fuse_bcast=fuse))
        if hasattr(sim.engine, "p1"):
            sim.engine.p1 = sim.engine.p2 = 0.0
        sim.run_round()
        native.reset_launch_count()
        sim.run_round()
        eager = native.LAUNCHES
        torch.cuda.synchronize()
        return eager, sim.arena.master.clone()

    fused_launches, fused = run(True)
    plain_launches, plain = run(False)
    assert fused_launches == plain_launches - 1          # K-BCAST is gone from the round
    assert float((fused - plain).norm() / plain.norm()) < 1e-3

"""Fused multi-GPU K-AGG kernel, G ranks emulated in one cooperative launch on one GPU."""

import pytest
import torch

from nanofed_b200 import ops
from nanofed_b200.ops import reference as ref
from nanofed_b200.parallel.peer import emulate_allreduce

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("g,k,n", [(2, 3, 4096), (4, 8, 65_536), (8, 4, 1_199_936)])
def test_emulated_allreduce_matches_flat_fedavg(g, k, n):
    torch.manual_seed(0)
    slots = torch.randn(g, k, n, device="cuda")
    counts = torch.randint(1, 500, (g, k), device="cuda").float()
    master, lowp = emulate_allreduce(slots, counts, calls=2)  # 2 calls: barrier reuse works
    want = ref.fedavg_reduce(slots.view(g * k, n).cpu(), counts.view(-1).cpu())
    for r in range(g):
        torch.testing.assert_close(master[r].cpu(), want, rtol=1e-5, atol=1e-5)
        assert torch.equal(master[r], master[0])
    torch.testing.assert_close(lowp[0].float().cpu(), want.bfloat16().float(), rtol=1e-2, atol=1e-2)


def test_emulated_allreduce_chunk_pipeline():
    """Several pipeline chunks (incl. a ragged last one): producers and consumers overlap."""
    g, k, n = 4, 3, 70_016
    slots = torch.randn(g, k, n, device="cuda")
    counts = torch.randint(1, 50, (g, k), device="cuda").float()
    master, _ = emulate_allreduce(slots, counts, calls=3, chunk_elems=8192)
    want = ref.fedavg_reduce(slots.view(g * k, n).cpu(), counts.view(-1).cpu())
    torch.testing.assert_close(master[0].cpu(), want, rtol=1e-5, atol=1e-5)
    assert all(torch.equal(master[r], master[0]) for r in range(g))


def test_emulated_allreduce_dp_matches_single_gpu_kernel():
    g, k, n = 4, 5, 32_768
    slots = torch.randn(g, k, n, device="cuda") * 0.02
    counts = torch.randint(1, 100, (g, k), device="cuda").float()
    mask = torch.ones(g, k, device="cuda")
    mask[1, 2] = 0
    ss, _ = ops.update_stats(slots.view(g * k, n))
    nm = torch.ones(n, dtype=torch.uint8, device="cuda")
    nm[::7] = 0
    kw = dict(clip_norm=1.0, noise_std=0.25, seed=11, offset=64)
    master, _ = emulate_allreduce(slots, counts, mask=mask, sumsq=ss.view(g, k), noise_mask=nm, **kw)
    single = ops.fedavg_reduce(slots.view(g * k, n), counts.view(-1), mask=mask.view(-1), sumsq=ss,
                               noise_mask=nm, **kw)
    torch.testing.assert_close(master[0], single, rtol=1e-4, atol=1e-5)
    assert torch.equal(master[0], master[g - 1])


def test_emulated_allreduce_streaming_waves_and_gathered_stats():
    """K clients fed in 3 waves (the first two local-only, accumulating the unnormalised partial):
    same result as one call over all clients, with DP clipping; every rank ends up holding every
    rank's metric rows."""
    g, k, n = 4, 6, 40_960
    torch.manual_seed(3)
    slots = torch.randn(g, k, n, device="cuda") * 0.05
    counts = torch.randint(1, 100, (g, k), device="cuda").float()
    ss, _ = ops.update_stats(slots.view(g * k, n))
    stats = torch.arange(g * 9, device="cuda", dtype=torch.float32).view(g, 9) + 0.5
    kw = dict(sumsq=ss.view(g, k), clip_norm=2.0)
    one, _ = emulate_allreduce(slots, counts, **kw)
    waves, _, got = emulate_allreduce(slots, counts, waves=3, calls=2, chunk_elems=8192, stats=stats,
                                      return_stats=True, **kw)
    torch.testing.assert_close(waves[0], one[0], rtol=1e-5, atol=1e-6)
    assert all(torch.equal(waves[r], waves[0]) for r in range(g))
    for r in range(g):
        assert torch.equal(got[r], stats)


def test_emulated_allreduce_keeps_the_model_when_every_client_is_masked():
    g, k, n = 2, 3, 8192
    slots = torch.full((g, k, n), float("nan"), device="cuda")
    counts = torch.ones(g, k, device="cuda")
    keep = torch.randn(n, device="cuda")
    master, _ = emulate_allreduce(slots, counts, mask=torch.zeros(g, k, device="cuda"),
                                  master_init=keep)
    assert torch.equal(master[0], keep) and torch.equal(master[1], keep)
    # single-GPU kernel: same rule
    out = keep.clone()
    ops.fedavg_reduce(slots.view(g * k, n), counts.view(-1), mask=torch.zeros(g * k, device="cuda"),
                      out=out)
    assert torch.equal(out, keep)


def test_barrier_timeout_is_reported_not_hung():
    """A wait that cannot be satisfied in time sets the status word; the wrapper raises."""
    g, k, n = 4, 2, 1 << 20
    slots = torch.randn(g, k, n, device="cuda")
    counts = torch.ones(g, k, device="cuda")
    with pytest.raises(RuntimeError, match="timed out"):
        emulate_allreduce(slots, counts, spin_ns=1)    # 1 ns: the first unsatisfied probe expires


def test_emulated_allreduce_with_broadcast_tail():
    """K-BCAST as the tail of the fused kernel: after the completion barrier every CTA of a rank
    replicates the new model into that rank's client rows (fp32 + bf16) and zeroes momentum."""
    g, k, n, rows = 4, 3, 70_016, 5
    torch.manual_seed(5)
    slots = torch.randn(g, k, n, device="cuda")
    counts = torch.randint(1, 50, (g, k), device="cuda").float()
    b_slots = torch.full((g, rows, n), -7.0, device="cuda")
    b_lowp = torch.zeros(g, rows, n, dtype=torch.bfloat16, device="cuda")
    b_mom = torch.ones(g, rows, n, device="cuda")
    master, _ = emulate_allreduce(slots, counts, calls=2, chunk_elems=8192,
                                  bcast=(b_slots, b_lowp, b_mom))
    want = ref.fedavg_reduce(slots.view(g * k, n).cpu(), counts.view(-1).cpu()).cuda()
    torch.testing.assert_close(master[0], want, rtol=1e-5, atol=1e-5)
    for r in range(g):
        for j in range(rows):
            assert torch.equal(b_slots[r, j], master[r])
            assert torch.equal(b_lowp[r, j], master[r].bfloat16())
    assert float(b_mom.abs().sum()) == 0.0

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

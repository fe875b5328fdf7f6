"""tcgen05 GEMM / conv kernels vs plain PyTorch fp32 (the checks of tools/tc_check.py)."""

import importlib.util
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu

_spec = importlib.util.spec_from_file_location(
    "tc_check", Path(__file__).resolve().parent.parent / "tools" / "tc_check.py")
tc_check = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(tc_check)


@pytest.mark.parametrize("name", sorted(tc_check.CHECKS))
def test_variant_matches_fp32_reference(name):
    from nanofed_b200.ops import native, tc

    before = native.LAUNCHES
    torch.manual_seed(0)
    got, ref = tc_check.CHECKS[name]()
    torch.cuda.synchronize()
    tc.check_status(got.device)
    tc_check._check_guards()          # nothing written outside the guarded outputs
    assert torch.isfinite(got).all()
    scale = ref.abs().max().item()
    # fp32 oracle (TF32 off).  fp32 outputs: 1e-3 of the output range, bf16 outputs: 1e-2
    tol = tc_check.tolerance(name)
    bad = int(((got - ref).abs() > tol * scale).sum())
    assert bad == 0, f"{name}: {bad}/{got.numel()} elements off (range {scale:.3g})"
    assert native.LAUNCHES > before, "no native kernel was launched"

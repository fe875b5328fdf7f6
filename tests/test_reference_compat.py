"""Drop-in check: the REFERENCE's own test-suite, executed against this package through the
``nanofed`` → ``nanofed_b200`` import alias (nothing is copied; the tests run from
/root/reference).  Skipped where the reference checkout is not mounted (e.g. the GPU box)."""

import subprocess
import sys
from pathlib import Path

import pytest

import os

REF_TESTS = Path(os.environ.get("NANOFED_REFERENCE_SRC", "/root/reference")) / "tests"
ROOT = Path(__file__).resolve().parent.parent


@pytest.mark.skipif(not REF_TESTS.exists(), reason="reference checkout not available")
def test_reference_suite_passes_against_this_package(tmp_path):
    cmd = [sys.executable, "-m", "pytest", str(REF_TESTS), "-p", "tools.ref_compat_plugin",
           "--rootdir", str(tmp_path), "-c", "/dev/null", "-q", "-p", "no:cacheprovider",
           # upstream flakes (they fail on the reference itself too): the random z-score test
           # ~1 in 6, and test_gradient_clipping compares a clipped norm to its bound with `<=`
           # in float32 (100.00001 vs 100.0)
           "-k", "not test_statistical_validation and not test_gradient_clipping"]
    proc = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=900)
    tail = proc.stdout.strip().splitlines()[-1] if proc.stdout.strip() else proc.stderr[-500:]
    assert proc.returncode == 0, proc.stdout[-3000:]
    assert "passed" in tail and "failed" not in tail


def test_alias_maps_modules_to_the_same_objects():
    code = ("import nanofed_b200.compat as c; c.install_alias(force=True); "
            "import nanofed, nanofed_b200; from nanofed.server.aggregator.fedavg import FedAvgAggregator as A; "
            "from nanofed_b200.server.aggregator.fedavg import FedAvgAggregator as B; "
            "assert A is B and nanofed.Coordinator is nanofed_b200.Coordinator; print('ok')")
    out = subprocess.run([sys.executable, "-c", code], cwd=ROOT, capture_output=True, text=True,
                         timeout=120)
    assert out.returncode == 0 and "ok" in out.stdout, out.stderr[-1000:]


def test_public_api_surface_matches_the_reference():
    """Every public name / parameter of every ``nanofed.*`` module exists in ``nanofed_b200.*``."""
    import subprocess
    import sys
    from pathlib import Path

    root = Path(__file__).resolve().parent.parent
    if not (root / "baseline" / "_ref" / "nanofed").exists():
        import pytest

        pytest.skip("reference not installed under baseline/_ref")
    p = subprocess.run([sys.executable, str(root / "tools" / "api_diff.py")], capture_output=True,
                       text=True, timeout=300)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]

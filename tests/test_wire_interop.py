"""Wire and checkpoint compatibility with the UNMODIFIED reference: its
HTTPClient completes a FedAvg round against our server, our HTTPClient works against its server,
and model versions written by either ModelManager are loaded by the other."""

import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
pytestmark = pytest.mark.skipif(not (ROOT / "baseline" / "_ref" / "nanofed").exists(),
                                reason="reference not installed under baseline/_ref")


@pytest.mark.parametrize("mode", ["ref_client", "ref_server", "checkpoint"])
def test_reference_and_this_package_interoperate_over_http(mode):
    p = subprocess.run([sys.executable, str(ROOT / "tests" / "_interop_worker.py"), mode],
                       capture_output=True, text=True, timeout=600)
    assert p.returncode == 0 and "INTEROP_OK" in p.stdout, p.stdout[-1500:] + p.stderr[-3000:]

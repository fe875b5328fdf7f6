"""Shared pytest configuration.

``gpu`` marks tests that need a real B200 (run with ``-m gpu`` through ``gpurun``);
everything else must pass on the CPU-only build box.
"""

import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

os.environ.setdefault("MASTER_ADDR", "127.0.0.1")

# The interoperability tests need the unmodified reference under baseline/_ref.  Where its source
# tree is mounted (the build box) install it before collection so their skipif sees it; on a GPU
# box it arrives with the snapshot.
try:
    from baseline.install_ref import install as _install_reference

    _install_reference()
except Exception:
    pass


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200, sm_100a)")
    # every oracle in this suite is fp32: PyTorch's default lets cuDNN convolutions use TF32
    try:
        import torch

        torch.backends.cudnn.allow_tf32 = False
        torch.backends.cuda.matmul.allow_tf32 = False
    except Exception:
        pass


def pytest_collection_modifyitems(config, items):
    try:
        import torch

        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(autouse=True)
def _quiet_logger():
    """Keep the framework logger from flooding test output."""
    import logging

    logging.getLogger("nanofed").setLevel(logging.WARNING)
    yield

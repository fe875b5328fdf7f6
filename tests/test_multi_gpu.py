"""Multi-GPU checks on real peer memory (skipped on boxes with fewer than two GPUs).

The host-side logic of the same code runs on CPU with gloo in test_dist_cpu.py; the kernel
itself is also tested on one GPU in emulation mode (test_peer_gpu.py).
"""

import os
import subprocess
import sys
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2,
                    reason="needs >= 2 GPUs")
def test_fused_allreduce_and_round_on_two_gpus():
    worker = Path(__file__).with_name("_mgpu_worker.py")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", "29733", str(worker)]
    p = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, (p.stdout[-2000:] + "\n" + p.stderr[-3000:])
    assert "MULTI_GPU_OK" in p.stdout

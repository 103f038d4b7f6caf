"""Iterated SpMV over two GPUs (NCCL) against one GPU; needs a box with >= 2 GPUs, skipped elsewhere."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_power_iteration_two_gpus_matches_one():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29731", os.path.join(ROOT, "tools", "power_parity.py")],
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and "POWER PARITY OK" in r.stdout, r.stdout[-4000:] + r.stderr[-4000:]

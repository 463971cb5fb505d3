"""N > 1 on real GPUs: views sharded over 2 processes / 2 GPUs, gathered to rank 0 with sr_gather (NCCL through the C
ABI), compared with the oracle.  Skipped on boxes with fewer than two devices."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_two_gpu_sharded_views_gathered_over_nccl_match_the_oracle():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "tests", "nccl_composite_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, cwd=ROOT, timeout=600)
    assert r.returncode == 0 and "bit-exact vs oracle" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]

"""NCCL multi-GPU checks (need >= 2 GPUs on the box; skipped otherwise): tools/mgpu_check.py under
torch.distributed.run compares the expert-parallel sample all-to-all and the ray-sharded render with the same
work done on one GPU.  The host-side logic of both is covered on CPU (gloo) by tests/test_parallel_cpu.py."""
import os
import subprocess
import sys

import pytest
import torch

from conftest import ROOT

pytestmark = pytest.mark.gpu


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_two_rank_nccl_a2a_and_sharded_render():
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "tools", "mgpu_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("-> ok") == 7, r.stdout

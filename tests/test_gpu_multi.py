"""The multi-GPU single-frame path across processes (needs two GPUs; skipped on a one-GPU box, where bench.py's config-3 leg
is the check): tile-row bands composed by the NCCL gather and by the fused peer-store exchange, bit-identical to one GPU."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu


def test_two_ranks_compose_the_single_gpu_frame():
    import torch
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29541", os.path.join(root, "tests", "multi_worker.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=root)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-4000:]
    assert "MULTI_OK world=2" in res.stdout

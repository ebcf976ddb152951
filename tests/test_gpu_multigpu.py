"""Data-parallel path on real GPUs (skipped on a single-GPU box): launches tests/multigpu_worker.py under torchrun with
one rank per GPU (NCCL).  The host-side logic is also covered on CPU by tests/test_parallel_gloo.py."""
import socket
import subprocess
import sys
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parents[1]


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs: NOT verified on this box — the all-reduced gradient bucket over NCCL, identical weights on every rank after Adam, sharded inference == whole-batch inference, sync-BN step == single-device step (verified on 2 x B200 by the builder: profiles/r02_multigpu_2gpu.txt; host logic on CPU: tests/test_parallel_gloo.py)")
def test_data_parallel_step_two_ranks(lib):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), str(ROOT / "tests" / "multigpu_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=str(ROOT))
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "multigpu ok world=2" in r.stdout

"""GPU, 2 ranks over NCCL: the sharded sort end to end -- fused K7 + peer-store
exchange (symmetric memory) and the all-to-all fallback -- against numpy.
Skipped on boxes with fewer than two GPUs (the host logic is covered on CPU by
tests/test_mgpu_host.py, the fused kernel on one GPU by tests/test_parity_gpu.py)."""
import os
import socket
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_sharded_sort_two_gpus():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tests", "_mgpu_worker.py")]
    p = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0 and "MGPU_OK" in p.stdout, (p.stdout[-2000:], p.stderr[-4000:])

"""Multi-GPU results on hardware: spawns one NCCL rank per GPU (2, when the box has them) and
checks every sharded path against the single-GPU result and the oracle -- the reference's
"distributed == local for the same seed" design (tests/unit/DenseSketchApplyElementalTest.cpp:28-106).
Skipped on a single-GPU box; bench.py --gpus N asserts the same on its own legs."""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def test_sharded_paths_match_single_gpu():
    import torch
    ngpu = torch.cuda.device_count()
    if ngpu < 2:
        pytest.skip("needs at least 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    world = 2
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(root, "tests", "nccl_worker.py")]
    r = subprocess.run(cmd, cwd=root, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "NCCL_WORKER_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-2000:]

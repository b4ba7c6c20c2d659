"""Fused exchange over real peer mappings (needs >= 2 GPUs; skipped on a 1-GPU box — the same kernels
are covered there by tests/test_gpu_comm_emulated.py)."""
import os
import subprocess
import sys

import pytest
import torch

from conftest import ROOT, free_port

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


def _ngpu():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
@pytest.mark.parametrize("world", [2, 4, 8])
def test_fused_exchanger_over_ipc(world):
    if _ngpu() < world:
        pytest.skip(f"needs {world} GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(free_port()), os.path.join(ROOT, "tests", "mgpu_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and "MGPU_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-6000:]

import os
import socket
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "multigpu: needs >= 2 CUDA devices with peer access")


def free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world_size, port, fn, args):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world_size), LOCAL_RANK=str(rank),
                      MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import torch
    torch.set_num_threads(1)
    fn(rank, world_size, *args)


def run_distributed(fn, world_size, *args):
    """Spawn ``world_size`` processes with the env a launcher would set; fn(rank, ws, *args)."""
    import torch.multiprocessing as mp
    mp.spawn(_worker, args=(world_size, free_port(), fn, args), nprocs=world_size, join=True)


@pytest.fixture
def dist_runner():
    return run_distributed

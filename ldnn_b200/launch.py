"""Process / device bring-up: one process per GPU.

Replaces the reference's import-time rendezvous constants and ``cuda:0`` pinning
(``Balanced All-Reduce/main.py:14-15,19,25``): rank ``r`` binds
``cuda:LOCAL_RANK`` and a launcher-provided ``MASTER_ADDR/PORT`` is respected
(the reference's ``localhost:29500`` is only the fallback).  The MPI variants'
``MPI.COMM_WORLD`` (``Balanced Ring/main.py:15-17``) maps onto the same
``torch.distributed`` group — NCCL/gloo here are bootstrap + control plane only;
the gradient exchange itself runs in our P2P kernels (``parallel/fused.py``).
"""
from __future__ import annotations

import datetime
import os
import random
from dataclasses import dataclass

import numpy as np
import torch
import torch.distributed as dist


@dataclass
class Runtime:
    rank: int
    world_size: int
    local_rank: int
    device: torch.device
    backend: str
    owns_group: bool

    @property
    def is_cuda(self) -> bool:
        return self.device.type == "cuda"

    def single_node(self) -> bool:
        """True when every rank of the job runs on this host (one NVSwitch / CUDA-IPC domain): the fused P2P
        exchange needs that.  Launchers export LOCAL_WORLD_SIZE; otherwise the hostnames are compared once
        (collective)."""
        if self.world_size == 1:
            return True
        if not hasattr(self, "_single_node"):
            lws = os.environ.get("LOCAL_WORLD_SIZE")
            if lws is not None:
                self._single_node = int(lws) == self.world_size
            else:
                import socket
                names = [None] * self.world_size
                dist.all_gather_object(names, socket.gethostname())
                self._single_node = len(set(names)) == 1
        return self._single_node


def launcher_env(env=None):
    """(rank, world size, local rank) from whichever launcher started this process.

    torchrun / torch.distributed.launch set RANK / WORLD_SIZE / LOCAL_RANK.  The reference's four ring variants are
    started as ``mpirun -n N python main.py`` (Balanced Ring/main.py:11-18 reads MPI.COMM_WORLD), so the variables of
    the common MPI launchers are understood too — Open MPI (OMPI_COMM_WORLD_*), MPICH / Intel MPI / Hydra (PMI_*),
    Slurm's srun (SLURM_PROCID, SLURM_NTASKS, SLURM_LOCALID) — and an existing `mpirun` line keeps working: the ranks
    rendezvous over torch.distributed at MASTER_ADDR:MASTER_PORT (127.0.0.1:29500 unless set; one host by default)."""
    env = os.environ if env is None else env
    for rk, wk, lk in (("RANK", "WORLD_SIZE", "LOCAL_RANK"),
                       ("OMPI_COMM_WORLD_RANK", "OMPI_COMM_WORLD_SIZE", "OMPI_COMM_WORLD_LOCAL_RANK"),
                       ("PMI_RANK", "PMI_SIZE", "MPI_LOCALRANKID"),
                       ("PMIX_RANK", "PMIX_SIZE", "PMIX_LOCAL_RANK"),
                       ("SLURM_PROCID", "SLURM_NTASKS", "SLURM_LOCALID")):
        if rk in env and wk in env:
            rank, world = int(env[rk]), int(env[wk])
            return rank, world, int(env.get(lk, str(rank)))
    return 0, 1, 0


def init_runtime(backend: str = "gloo", seed=None, force_cpu: bool = False) -> Runtime:
    rank, world, local_rank = launcher_env()
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29500")

    use_cuda = torch.cuda.is_available() and not force_cpu
    if use_cuda:
        ndev = torch.cuda.device_count()
        device = torch.device("cuda", local_rank % ndev)
        torch.cuda.set_device(device)
    else:
        device = torch.device("cpu")
        if backend == "nccl":
            backend = "gloo"

    owns = False
    if not dist.is_initialized():
        kw = {}
        if backend == "nccl":
            kw["device_id"] = device
        dist.init_process_group(backend, rank=rank, world_size=world,
                                timeout=datetime.timedelta(minutes=10), **kw)
        owns = True
    rt = Runtime(dist.get_rank(), dist.get_world_size(), local_rank, device, backend, owns)
    seed_everything(seed, rt.rank)
    return rt


def seed_everything(seed, rank: int = 0) -> None:
    """The reference never seeds.  ``seed=None`` keeps that; a seed makes
    model init identical on all ranks and data draws rank-decorrelated."""
    if seed is None:
        return
    torch.manual_seed(seed)
    np.random.seed((seed * 9973 + rank) % (2 ** 31 - 1))
    random.seed(seed * 7919 + rank)


def shutdown(rt: Runtime) -> None:
    if rt.owns_group and dist.is_initialized():
        dist.destroy_process_group()

"""RESTATED reference arm for the Ring / Double-Ring variants.

The reference's four ring directories cannot be "switched to the NCCL backend": they hard-wire
``mpi4py`` (``Balanced Ring/main.py:11-18``; not installed here and torch has no MPI backend), they
stage every tensor through host numpy (``Balanced Ring/communication.py:15-28``), and on CUDA tensors
three of their four exchange functions rebind the local name and change nothing.
What can be measured is the reference's OWN training code with its exchange re-expressed 1:1 on
``torch.distributed`` P2P over NCCL:

* model and ``train_local_epoch`` are imported unmodified from ``baseline/_ref/<variant dir>``
  (an empty ``mpi4py`` stand-in module satisfies the import line; nothing of it is called);
* the exchange keeps the reference's peers, its per-parameter granularity (one message per parameter
  tensor, driven by the trainer's loop, ``Balanced Ring/trainer.py:147-168``,
  ``Balanced Double-Ring/trainer.py:150-171``) and its separate elementwise combine
  (``tensor.add_(recv).div_(2)``, ``communication.py:30``; ``add_.add_.div_(3)``,
  ``Balanced Double-Ring/communication.py:39-40``); only Isend/Irecv/Waitall on host copies become
  ``dist.batch_isend_irecv`` on the device tensors.

Everything this file reports is labelled ``restated``: it is the reference's intended math on NCCL, not
a run of the reference's shipped (CPU-only, MPI) exchange.
"""
from __future__ import annotations

import os
import sys
import types

import torch
import torch.distributed as dist

VARIANT_DIR = {
    ("ring", False): "Balanced Ring", ("ring", True): "Disbalanced Ring",
    ("double_ring", False): "Balanced Double-Ring", ("double_ring", True): "Disalanced Double-Ring",
    ("allreduce", False): "Balanced All-Reduce", ("allreduce", True): "Disbalanced All-Reduce",
}


def install_mpi_stub() -> None:
    """``from mpi4py import MPI`` at the top of the ring variants' modules must import; nothing else
    of mpi4py is touched on the path timed here (the exchange below replaces every ``comm.*`` call)."""
    if "mpi4py" in sys.modules:
        return
    m = types.ModuleType("mpi4py")
    m.MPI = types.SimpleNamespace(COMM_WORLD=None, SUM="SUM", MAX="MAX", FLOAT="FLOAT", DOUBLE="DOUBLE",
                                  Request=types.SimpleNamespace(Waitall=lambda reqs: None))
    m.__doc__ = "stand-in so the reference's ring modules import without mpi4py (see baseline/ring_nccl.py)"
    sys.modules["mpi4py"] = m


def import_reference(ref_root: str, topology: str, disbalanced: bool = False):
    """(EnhancedCNNModel, train_local_epoch, variant directory name) of the matching reference variant."""
    name = VARIANT_DIR[(topology, disbalanced)]
    vdir = os.path.join(ref_root, name)
    if not os.path.isdir(vdir):
        raise FileNotFoundError(vdir)
    install_mpi_stub()
    for mod in ("model", "trainer", "communication", "dataloader", "validator"):
        sys.modules.pop(mod, None)           # the six directories reuse the same module names
    sys.path.insert(0, vdir)
    try:
        from model import EnhancedCNNModel    # noqa: E402  (reference module, unmodified)
        from trainer import train_local_epoch  # noqa: E402
    finally:
        sys.path.remove(vdir)
    return EnhancedCNNModel, train_local_epoch, name


def _exchange(tensor: torch.Tensor, rank: int, world_size: int, hops):
    """Send `tensor` to rank+h and receive from rank-h for every hop h — one batched NCCL group,
    like the reference's Irecv/Isend/Waitall block.  A peer that is this rank itself (world sizes 1, 2)
    is a local copy."""
    recvs, ops = [], []
    for h in hops:
        dst, src = (rank + h) % world_size, (rank - h + world_size) % world_size
        r = torch.zeros_like(tensor)                      # communication.py:12 (zeros_like per call)
        recvs.append(r)
        if src == rank:
            r.copy_(tensor)
        else:
            ops.append(dist.P2POp(dist.irecv, r, src))
            ops.append(dist.P2POp(dist.isend, tensor, dst))
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()
    return recvs


def ring_all_reduce_equal(tensor, rank, world_size):
    """Balanced Ring/communication.py:5-30 on device tensors."""
    (recv,) = _exchange(tensor, rank, world_size, (1,))
    tensor.add_(recv).div_(2)


def ring_all_reduce_weighted(tensor, rank, world_size, local_weight):
    """Balanced Ring/communication.py:33-62."""
    (recv,) = _exchange(tensor, rank, world_size, (1,))
    tensor.mul_(local_weight).add_(recv, alpha=1 - local_weight)


def double_ring_all_reduce(tensor, rank, world_size):
    """Balanced Double-Ring/communication.py:5-40."""
    r1, r2 = _exchange(tensor, rank, world_size, (1, 2))
    tensor.add_(r1).add_(r2)
    tensor.div_(3)


def double_ring_all_reduce_weighted(tensor, rank, world_size, local_weight):
    """Balanced Double-Ring/communication.py:43-77."""
    r1, r2 = _exchange(tensor, rank, world_size, (1, 2))
    w = (1 - local_weight) / 2
    tensor.mul_(local_weight).add_(r1, alpha=w).add_(r2, alpha=w)


def exchange_model(model, rank, world_size, topology="ring", aggregation_by="gradients",
                   aggregation_type="equal", local_weight=0.5):
    """The trainer's per-parameter driver loop (Balanced Ring/trainer.py:147-168,
    Balanced Double-Ring/trainer.py:150-171): one exchange call per parameter tensor."""
    if topology == "ring":
        fn = ring_all_reduce_weighted if aggregation_type == "weighted" else ring_all_reduce_equal
    else:
        fn = double_ring_all_reduce_weighted if aggregation_type == "weighted" else double_ring_all_reduce
    extra = (local_weight,) if aggregation_type == "weighted" else ()
    for param in model.parameters():
        if aggregation_by == "gradients":
            if param.grad is not None:
                fn(param.grad.data, rank, world_size, *extra)
        elif param.requires_grad:
            fn(param.data, rank, world_size, *extra)

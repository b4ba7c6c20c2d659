"""Checkpoint save / resume (S-CKPT).

The reference never saves anything (no ``torch.save``/``load`` in the tree; SURVEY
§0.1 row 5).  The layout kept compatible is ``EnhancedCNNModel().state_dict()``:
128 keys, fp32 OIHW conv weights, fp32 running stats, int64 ``num_batches_tracked``
(``Balanced All-Reduce/model.py:74-100``) — whatever the engine's internal layout
or compute dtype is, ``engine.state_dict()`` exports exactly that, so a checkpoint
written here loads into the upstream ``nn.Module`` with ``load_state_dict`` and
vice versa.

Replicas.  Under the reference's schedule (local SGD, one exchange per global epoch, gradient aggregation that
never reaches the weights — SURVEY B1) and under gossip ring / double-ring, the ranks' models legitimately differ,
and BN running statistics and optimizer moments are per-rank in every schedule (B6).  The main file is rank 0's
state in the upstream layout (what upstream's rank-0 test evaluation uses, ``main.py:61-62``); every other rank
writes its own model/optimizer beside it as ``<path>.rank<r>`` and restores from it on resume, so a resume
continues every replica's own trajectory instead of collapsing all ranks onto rank 0.  If the side file is
missing or belongs to another run (epoch / world size differ) rank 0's state is used and a warning says so.
"""
from __future__ import annotations

import os
from typing import Optional

import numpy as np
import torch


def _shard_blob(shards):
    return {"train": np.asarray(shards.train_indices), "val": np.asarray(shards.val_indices),
            "ratios": list(shards.ratios), "fixed_classes": shards.fixed_classes}


def gather_shards(shards, rt):
    """Collective: every rank's shard state, as a list indexed by rank (each rank trains on its own shard, so a
    checkpoint written by rank 0 has to carry all of them).  Returns None for a single process."""
    if shards is None or rt is None or rt.world_size == 1:
        return None
    import torch.distributed as dist
    out = [None] * rt.world_size
    dist.all_gather_object(out, _shard_blob(shards))
    return out


def save_checkpoint(path: str, engine, history=None, shards=None, epoch: int = 0, cfg=None, all_shards=None) -> None:
    """``shards`` is the writer's own shard state; ``all_shards`` (from :func:`gather_shards`) every rank's."""
    blob = {
        "model": {k: v.detach().cpu() for k, v in engine.state_dict().items()},
        "optimizer": engine.optimizer_state_dict(),
        "epoch": int(epoch),
        "history": None if history is None else dict(vars(history)),
        "indices": None if shards is None else _shard_blob(shards),
        "indices_per_rank": all_shards,
        "config": None if cfg is None else cfg.to_dict(),
    }
    tmp = path + ".tmp"
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    torch.save(blob, tmp)
    os.replace(tmp, path)


def rank_path(path: str, rank: int) -> str:
    return path if rank == 0 else f"{path}.rank{rank}"


def save_rank_state(path: str, engine, rank: int, world_size: int, epoch: int) -> None:
    """Rank r > 0: its own replica (model incl. BN buffers + optimizer) beside the main checkpoint."""
    blob = {"model": {k: v.detach().cpu() for k, v in engine.state_dict().items()},
            "optimizer": engine.optimizer_state_dict(), "epoch": int(epoch), "rank": int(rank),
            "world_size": int(world_size)}
    out = rank_path(path, rank)
    os.makedirs(os.path.dirname(os.path.abspath(out)), exist_ok=True)
    torch.save(blob, out + ".tmp")
    os.replace(out + ".tmp", out)


def load_checkpoint(path: str, engine, history=None, shards=None, rank: int = 0, world_size: int = 1) -> int:
    """Restores engine (+history, +shards in place when given).  Returns the next
    global epoch.  Also accepts a bare ``state_dict`` file written by upstream-style code.

    Shards: rank ``r`` takes its own saved shard when the checkpoint carries one per rank for this world size;
    a single-process checkpoint restores into a single process; otherwise (world size changed) the freshly
    computed shards are kept — restoring rank 0's shard everywhere would make every rank train on the same data."""
    blob = torch.load(path, map_location="cpu", weights_only=False)
    if "model" not in blob:                      # bare state_dict
        engine.load_state_dict(blob)
        return 0
    mine = blob
    if rank > 0:                                 # this rank's own replica, when the same run wrote one
        rp = rank_path(path, rank)
        side = torch.load(rp, map_location="cpu", weights_only=False) if os.path.exists(rp) else None
        if side is not None and side.get("epoch") == blob.get("epoch") and side.get("world_size") == world_size:
            mine = side
        elif world_size > 1:
            import warnings
            warnings.warn(f"checkpoint {path}: no matching replica file for rank {rank} "
                          f"({'missing' if side is None else 'different epoch / world size'}); restoring rank 0's "
                          "model and optimizer state on this rank — replicas that had diverged are collapsed")
    engine.load_state_dict(mine["model"])
    if mine.get("optimizer") is not None:
        engine.load_optimizer_state_dict(mine["optimizer"])
    if history is not None and blob.get("history"):
        for k, v in blob["history"].items():
            setattr(history, k, v)
    if shards is not None:
        per_rank = blob.get("indices_per_rank")
        ix = None
        if per_rank is not None and len(per_rank) == world_size:
            ix = per_rank[rank]
        elif per_rank is None and world_size == 1 and blob.get("indices"):
            ix = blob["indices"]
        if ix is not None:
            shards.train_indices, shards.val_indices = np.asarray(ix["train"]), np.asarray(ix["val"])
            shards.ratios, shards.fixed_classes = list(ix["ratios"]), ix["fixed_classes"]
    return int(blob.get("epoch", 0))

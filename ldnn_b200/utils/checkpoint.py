"""Checkpoint save / resume (S-CKPT).

The reference never saves anything (no ``torch.save``/``load`` in the tree; SURVEY
§0.1 row 5).  The layout kept compatible is ``EnhancedCNNModel().state_dict()``:
128 keys, fp32 OIHW conv weights, fp32 running stats, int64 ``num_batches_tracked``
(``Balanced All-Reduce/model.py:74-100``) — whatever the engine's internal layout
or compute dtype is, ``engine.state_dict()`` exports exactly that, so a checkpoint
written here loads into the upstream ``nn.Module`` with ``load_state_dict`` and
vice versa.
"""
from __future__ import annotations

import os
from typing import Optional

import numpy as np
import torch


def save_checkpoint(path: str, engine, history=None, shards=None, epoch: int = 0, cfg=None) -> None:
    blob = {
        "model": {k: v.detach().cpu() for k, v in engine.state_dict().items()},
        "optimizer": engine.optimizer_state_dict(),
        "epoch": int(epoch),
        "history": None if history is None else dict(vars(history)),
        "indices": None if shards is None else {
            "train": np.asarray(shards.train_indices), "val": np.asarray(shards.val_indices),
            "ratios": list(shards.ratios), "fixed_classes": shards.fixed_classes},
        "config": None if cfg is None else cfg.to_dict(),
    }
    tmp = path + ".tmp"
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    torch.save(blob, tmp)
    os.replace(tmp, path)


def load_checkpoint(path: str, engine, history=None, shards=None) -> int:
    """Restores engine (+history, +shards in place when given).  Returns the next
    global epoch.  Also accepts a bare ``state_dict`` file written by upstream-style code."""
    blob = torch.load(path, map_location="cpu", weights_only=False)
    if "model" not in blob:                      # bare state_dict
        engine.load_state_dict(blob)
        return 0
    engine.load_state_dict(blob["model"])
    if blob.get("optimizer") is not None:
        engine.load_optimizer_state_dict(blob["optimizer"])
    if history is not None and blob.get("history"):
        for k, v in blob["history"].items():
            setattr(history, k, v)
    if shards is not None and blob.get("indices"):
        ix = blob["indices"]
        shards.train_indices, shards.val_indices = np.asarray(ix["train"]), np.asarray(ix["val"])
        shards.ratios, shards.fixed_classes = list(ix["ratios"]), ix["fixed_classes"]
    return int(blob.get("epoch", 0))

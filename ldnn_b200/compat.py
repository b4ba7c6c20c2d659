"""Reference-signature API, so scripts written against the upstream modules keep working.

Each variant directory's ``trainer.py / communication.py / dataloader.py / model.py /
validator.py / evaluator.py / vizualizator.py`` re-exports from here with its topology
and balance pinned.  Signatures follow upstream:

* ``get_loaders(batch_size, world_size, rank, model, device[, fixed_ratio])``
  (``Balanced All-Reduce/dataloader.py:9``, ``Disbalanced Ring/dataloader.py:9``)
* ``train_global(model, trainloader, val_loader, trainset, valset, indices_train,
  indices_val, criterion, optimizer, scheduler, device, rank, world_size, …)`` → 12 lists
  (``Balanced All-Reduce/trainer.py:11,192``; MPI variants thread a ``comm`` argument,
  ``Balanced Ring/trainer.py:11`` — accepted and ignored: transport is torch.distributed
  plus our P2P kernels)
* ``train_local_epoch(model, trainloader, criterion, optimizer, device, scheduler)``
* ``validate(model, val_loader, criterion, device)``, ``evaluate(model, loader, criterion, device, rank)``
* ``average_{gradients,weights}_{equal,weighted}``, ``ring_all_reduce_{equal,weighted}``,
  ``double_ring_all_reduce[_weighted]`` — in place, and correct on CUDA tensors
  (upstream's ring functions are silent no-ops on GPU).
"""
from __future__ import annotations

import os
from types import SimpleNamespace
from typing import Optional

import numpy as np
import torch
import torch.distributed as dist
import torch.nn as nn

from . import data as D
from .config import Config, resolve
from .engine import Engine, FlatState, OptimConfig, build_layout
from .parallel.exchange import DistExchanger, ExchangeSpec
from . import trainer as T


# ---------------------------------------------------------------------------
# runtime discovery for code that passes (device, rank, world_size) around
# ---------------------------------------------------------------------------
def _runtime(device=None, rank=None, world_size=None):
    init = dist.is_available() and dist.is_initialized()
    r = rank if rank is not None else (dist.get_rank() if init else 0)
    ws = world_size if world_size is not None else (dist.get_world_size() if init else 1)
    dev = torch.device(device) if device is not None else torch.device("cpu")
    backend = dist.get_backend() if init else "gloo"
    return SimpleNamespace(rank=r, world_size=ws, local_rank=r, device=dev, backend=backend,
                           owns_group=False, is_cuda=dev.type == "cuda")


class ModuleEngine(Engine):
    """Adapts an arbitrary ``nn.Module`` + torch optimizer/scheduler/criterion to the
    engine interface; parameters and grads are re-seated into flat slabs so the
    exchange still runs on one contiguous buffer."""

    def __init__(self, model: nn.Module, criterion=None, optimizer=None, scheduler=None, device=None):
        self.model, self.criterion = model, criterion or nn.CrossEntropyLoss()
        self.optimizer, self.scheduler = optimizer, scheduler
        self.device = torch.device(device) if device is not None else next(model.parameters()).device
        self.spec = getattr(model, "spec", None)
        named = [(k, p) for k, p in model.named_parameters() if p.requires_grad]
        slots, total = build_layout([(k, tuple(p.shape)) for k, p in named])
        lr = optimizer.param_groups[0]["lr"] if optimizer is not None else 1e-3
        self.flat = FlatState(slots, total, self.device, OptimConfig(lr=lr))
        with torch.no_grad():
            for k, p in named:
                v = self.flat.view(self.flat.params, k)
                v.copy_(p.data)
                p.data = v
                p.grad = self.flat.view(self.flat.grads, k)

    def train_step(self, x, y, update=True):
        self.model.train()
        self.flat.grads.zero_()
        out = self.model(x)
        loss = self.criterion(out, y)
        loss.backward()
        if update:
            self.optimizer.step() if self.optimizer is not None else self.flat.optimizer_step()
        return loss.detach(), (out.detach().argmax(1) == y).sum()

    @torch.no_grad()
    def eval_step(self, x, y):
        self.model.eval()
        out = self.model(x)
        pred = out.argmax(1)
        return self.criterion(out, y), (pred == y).sum(), pred

    def scheduler_step(self):
        if self.scheduler is not None:
            self.scheduler.step()
        self.flat.scheduler_step()

    def state_dict(self):
        return {k: v.detach().clone() for k, v in self.model.state_dict().items()}

    def load_state_dict(self, sd):
        with torch.no_grad():
            own = self.model.state_dict()
            for k, v in sd.items():
                own[k].copy_(v.to(own[k].device))


def _as_engine(model, criterion=None, optimizer=None, scheduler=None, device=None) -> Engine:
    return model if isinstance(model, Engine) else ModuleEngine(model, criterion, optimizer, scheduler, device)


# ---------------------------------------------------------------------------
# dataloader.py surface
# ---------------------------------------------------------------------------
_BUNDLE = {}


def _bundle(device=None, synthetic=None):
    key = "b"
    if key not in _BUNDLE:
        cfg = Config(synthetic=bool(int(os.environ.get("LDNN_SYNTHETIC", "0"))) if synthetic is None else synthetic,
                     quiet=True)
        _BUNDLE[key] = D.get_data(cfg)
    return _BUNDLE[key]


def get_loaders(batch_size, world_size, rank, model, device, fixed_ratio=None, *, balance="balanced",
                probe_batches=10, bundle=None):
    b = bundle or _bundle(device)
    rt = _runtime(device, rank, world_size)
    cfg = resolve(Config(batch_size=batch_size, balance=balance, probe_batches=probe_batches,
                         fixed_ratio=0.5 if fixed_ratio is None else fixed_ratio))
    st = D.initial_shards(b, cfg, rt, _as_engine(model, device=device))
    out = (D.Loader(b.trainset, st.train_indices, batch_size), D.Loader(b.valset, st.val_indices, batch_size),
           D.Loader(b.testset, None, batch_size), b.trainset, b.valset, st.train_indices, st.val_indices)
    return out + (st.fixed_classes,) if balance == "disbalanced" else out


def get_dataset_partition(dataset, rank, efficiency_ratios, fixed_ratio=None):
    if fixed_ratio is None:
        idx = D.partition_balanced(len(dataset), rank, efficiency_ratios)
        return D.SubsetView(dataset, idx), list(idx)
    idx, fixed = D.partition_class_skew(dataset.labels, rank, efficiency_ratios, fixed_ratio)
    return D.SubsetView(dataset, idx), list(idx), fixed


def get_next_dataset_partition(dataset, prev_indices, prev_fraction, next_fraction, epoch_duration,
                               epoch_duration_tensor, fixed_classes=None, fixed_ratio=0.5, replace=True):
    tot = float(epoch_duration_tensor.item()) if torch.is_tensor(epoch_duration_tensor) else float(epoch_duration_tensor)
    idx = D.next_partition(len(dataset), prev_indices, prev_fraction, next_fraction, epoch_duration / max(tot, 1e-12),
                           replace, labels=dataset.labels, fixed_classes=fixed_classes, fixed_ratio=fixed_ratio)
    return D.SubsetView(dataset, idx), list(idx)


def get_subset_loaders(trainset, valset, train_indices, val_indices, batch_size, prev_fraction, next_fraction,
                       epoch_duration, epoch_duration_tensor, fixed_classes=None, fixed_ratio=0.5, replace=True):
    _, tr = get_next_dataset_partition(trainset, train_indices, prev_fraction, next_fraction, epoch_duration,
                                       epoch_duration_tensor, fixed_classes, fixed_ratio, replace)
    _, va = get_next_dataset_partition(valset, val_indices, prev_fraction, next_fraction, epoch_duration,
                                       epoch_duration_tensor, fixed_classes, fixed_ratio, replace)
    return D.Loader(trainset, tr, batch_size), D.Loader(valset, va, batch_size), tr, va


def estimate_epoch_duration(trainloader, world_size, model, device, num_batches=10):
    rt = _runtime(device, None, world_size)
    return D.estimate_epoch_duration(_as_engine(model, device=device), trainloader, rt, num_batches)


# ---------------------------------------------------------------------------
# trainer.py / validator.py / evaluator.py surface
# ---------------------------------------------------------------------------
def train_local_epoch(model, trainloader, criterion, optimizer, device, scheduler, *unused):
    eng = _as_engine(model, criterion, optimizer, scheduler, device)
    loss, acc, losses, _ = T.train_local_epoch(eng, trainloader, _runtime(device))
    return loss, acc, losses.tolist()


def validate(model, val_loader, criterion, device, *unused):
    return T.validate(_as_engine(model, criterion, device=device), val_loader, _runtime(device))


def evaluate(model, loader, criterion, device, rank):
    rt = _runtime(device, rank)
    ncls = int(loader.dataset.labels.max()) + 1 if len(loader.dataset) else 10
    return T.evaluate(_as_engine(model, criterion, device=device), loader, rt, max(ncls, 10))


def make_train_global(topology: str, balance: str):
    """Reference-signature ``train_global`` for one variant directory."""
    mpi_style = topology != "allreduce"

    def train_global(model, trainloader, val_loader, trainset, valset, indices_train, indices_val, criterion,
                     optimizer, scheduler, device, rank, world_size, num_local_epochs, num_global_epochs,
                     timelimit, batch_size, prev_fraction, next_fraction, *rest):
        rest = list(rest)
        if mpi_style and rest and not isinstance(rest[0], (int, float, str, list, tuple)):
            rest.pop(0)                                  # the MPI communicator: not needed
        fixed_classes, fixed_ratio = None, 0.5
        if balance == "disbalanced":
            # DisR passes (local_weight, fixed_classes, …); DisDR passes (fixed_classes, local_weight, …)
            if isinstance(rest[0], (list, tuple)):
                fixed_classes, local_weight = list(rest[0]), rest[1]
            else:
                local_weight, fixed_classes = rest[0], list(rest[1])
            aggregation_type, aggregation_by = rest[2], rest[3]
            if len(rest) > 4:
                fixed_ratio = rest[4]
        else:
            local_weight, aggregation_type, aggregation_by = rest[0], rest[1], rest[2]
        cfg = resolve(Config(topology=topology, balance=balance, epochs_local=num_local_epochs,
                             epochs_global=num_global_epochs, time_limit=timelimit, batch_size=batch_size,
                             prev_fraction=prev_fraction, next_fraction=next_fraction, local_weight=local_weight,
                             aggregation_type=aggregation_type, aggregation_by=aggregation_by,
                             fixed_ratio=fixed_ratio))
        rt = _runtime(device, rank, world_size)
        eng = _as_engine(model, criterion, optimizer, scheduler, device)
        bundle = D.DataBundle(trainset, valset, valset, int(trainset.labels.max()) + 1 if len(trainset) else 10, "user")
        shards = D.ShardState(np.asarray(indices_train), np.asarray(indices_val),
                              [1.0 / world_size] * world_size, fixed_classes)
        ex = CompatExchanger(ExchangeSpec.from_config(cfg, world_size), rank)     # fused P2P kernels on a GPU job
        return T.train_global(eng, bundle, shards, cfg, rt, ex).as_tuple()

    return train_global


# ---------------------------------------------------------------------------
# communication.py surface
# ---------------------------------------------------------------------------
class CompatExchanger(DistExchanger):
    """The exchanger behind the reference-signature entry points.  On a single-node GPU job every exchange of an fp32
    CUDA tensor runs on the hand-written P2P kernels (staged through a symmetric slab, `StagedFusedExchange`) — the
    variant directories' ``communication.py`` / ``trainer.py`` are product paths, not NCCL calls.  CPU tensors, other
    dtypes and multi-node jobs take the torch.distributed realisation it inherits (`LDNN_COMPAT_DIST=1` forces it)."""

    @torch.no_grad()
    def exchange_(self, t: torch.Tensor) -> torch.Tensor:
        if self.spec.world_size > 1 and self.group is None:
            from .parallel.fused import StagedFusedExchange
            dev = t.device
            if StagedFusedExchange.usable(None, t):
                rt = _runtime(dev, self.rank, self.spec.world_size)
                return StagedFusedExchange.get(rt).exchange_(t, self.spec)
        return super().exchange_(t)


def shutdown():
    """Release what the exchange functions of this module allocated on the GPU (symmetric staging slab, signal pads).
    Collective; optional — for long-lived hosts that keep the process after training."""
    from .parallel.fused import StagedFusedExchange
    StagedFusedExchange.close_all()


def _ex(topology, kind, world_size, local_weight=0.5, rank=None):
    r = rank if rank is not None else (dist.get_rank() if dist.is_initialized() else 0)
    sem = "allreduce" if topology == "allreduce" else "gossip"
    return CompatExchanger(ExchangeSpec(topology, sem, kind, local_weight, None, world_size), r)


def _each(model, what):
    if isinstance(model, Engine):
        yield model.flat.grads if what == "grad" else model.flat.params
        return
    for p in model.parameters():
        if what == "grad":
            if p.grad is not None:
                yield p.grad.data
        else:
            yield p.data


def average_gradients_equal(model, world_size):
    ex = _ex("allreduce", "equal", world_size)
    for t in _each(model, "grad"):
        ex.exchange_(t)


def average_weights_equal(model, world_size):
    ex = _ex("allreduce", "equal", world_size)
    for t in _each(model, "param"):
        ex.exchange_(t)


def average_gradients_weighted(model, world_size, local_weight):
    ex = _ex("allreduce", "weighted", world_size, local_weight)
    for t in _each(model, "grad"):
        ex.exchange_(t)


def average_weights_weighted(model, world_size, local_weight):
    ex = _ex("allreduce", "weighted", world_size, local_weight)
    for t in _each(model, "param"):
        ex.exchange_(t)


def ring_all_reduce_equal(tensor, rank, world_size):
    return _ex("ring", "equal", world_size, rank=rank).exchange_(tensor)


def ring_all_reduce_weighted(tensor, rank, world_size, local_weight):
    return _ex("ring", "weighted", world_size, local_weight, rank).exchange_(tensor)


def double_ring_all_reduce(tensor, rank, world_size):
    return _ex("double_ring", "equal", world_size, rank=rank).exchange_(tensor)


def double_ring_all_reduce_weighted(tensor, rank, world_size, local_weight):
    return _ex("double_ring", "weighted", world_size, local_weight, rank).exchange_(tensor)

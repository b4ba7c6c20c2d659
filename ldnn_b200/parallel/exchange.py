"""Exchange semantics (what is mixed with what) and the torch.distributed realisation.

One *mixing matrix* ``W`` (``result_r = Σ_p W[r,p]·x_p``) describes every variant of
the reference's ``communication.py`` files — and is the fp64 oracle in the tests:

=============  =========================================  ==========================
topology       reference math ("gossip" semantics)        source
=============  =========================================  ==========================
allreduce      global mean / self-weighted mean           BalAR/communication.py:4-31
ring           ½·self + ½·x[r-1]   (w·self+(1-w)·x[r-1])   BalR/communication.py:5-62
double_ring    ⅓(self + x[r-1] + x[r-2])  (w, (1-w)/2 ×2)  BalDR/communication.py:5-77
=============  =========================================  ==========================

``semantics="allreduce"`` is the north-star reading of ring / double-ring: the *result*
is the global (optionally shard-weighted) mean, and the topology only selects the
schedule (RS+AG around one ring, or around two counter-rotating rings on the two
halves of the bucket).  ``aggregation_type="shard"`` uses ``w_p = n_p/Σn``.

Defined behaviour at the sizes the reference breaks on: ``ws==1`` is
the identity everywhere (upstream divides by ``ws-1``); double-ring at ``ws==2``
gives ``(2·self + other)/3`` exactly as upstream's aliasing peers do.

`DistExchanger` runs this math over ``torch.distributed`` (gloo on CPU; NCCL as the
*measured baseline* on GPU).  The product path is `parallel/fused.py`.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional, Sequence

import numpy as np
import torch
import torch.distributed as dist


@dataclass
class ExchangeSpec:
    topology: str = "allreduce"          # allreduce | ring | double_ring
    semantics: str = "allreduce"         # allreduce | gossip
    aggregation_type: str = "equal"      # equal | weighted | shard
    local_weight: float = 0.5
    shard_weights: Optional[Sequence[float]] = None   # n_p / Σn, len == world_size
    world_size: int = 1

    @staticmethod
    def from_config(cfg, world_size: int, shard_weights=None) -> "ExchangeSpec":
        return ExchangeSpec(cfg.topology, cfg.semantics, cfg.aggregation_type, cfg.local_weight,
                            shard_weights, world_size)

    def neighbour_offsets(self) -> List[int]:
        """Which ``r-d`` peers a gossip step pulls from."""
        if self.topology == "ring":
            return [1]
        if self.topology == "double_ring":
            return [1, 2]
        return []

    def is_gossip(self) -> bool:
        return self.semantics == "gossip" and self.topology != "allreduce"


def mixing_matrix(spec: ExchangeSpec) -> np.ndarray:
    ws = spec.world_size
    W = np.zeros((ws, ws), dtype=np.float64)
    if ws == 1:
        W[0, 0] = 1.0
        return W
    sw = None
    if spec.aggregation_type == "shard":
        sw = np.asarray(spec.shard_weights if spec.shard_weights is not None else [1.0 / ws] * ws, np.float64)
        sw = sw / sw.sum()
    if not spec.is_gossip():
        for r in range(ws):
            if spec.aggregation_type == "equal":
                W[r, :] = 1.0 / ws
            elif spec.aggregation_type == "shard":
                W[r, :] = sw
            else:
                W[r, :] = (1.0 - spec.local_weight) / (ws - 1)
                W[r, r] = spec.local_weight
        return W
    offs = spec.neighbour_offsets()
    for r in range(ws):
        if spec.aggregation_type == "equal":
            c_self, c_nb = 1.0 / (len(offs) + 1), [1.0 / (len(offs) + 1)] * len(offs)
        elif spec.aggregation_type == "weighted":
            c_self, c_nb = spec.local_weight, [(1.0 - spec.local_weight) / len(offs)] * len(offs)
        else:
            members = [r] + [(r - d) % ws for d in offs]
            tot = sum(sw[m] for m in members)
            c_self, c_nb = sw[r] / tot, [sw[(r - d) % ws] / tot for d in offs]
        W[r, r] += c_self
        for d, c in zip(offs, c_nb):
            W[r, (r - d) % ws] += c
    return W


def gossip_coefficients(spec: ExchangeSpec, rank: int):
    """(c_self, [(offset d, c_d)…]) for one rank — what the pull kernels take."""
    W = mixing_matrix(spec)
    ws = spec.world_size
    offs = spec.neighbour_offsets()
    if ws == 1:
        return 1.0, []
    # peers may alias (ws==2 double ring): split the merged entry back per offset
    if spec.aggregation_type == "equal":
        c = 1.0 / (len(offs) + 1)
        return c, [(d, c) for d in offs]
    if spec.aggregation_type == "weighted":
        return spec.local_weight, [(d, (1.0 - spec.local_weight) / len(offs)) for d in offs]
    sw = np.asarray(spec.shard_weights, np.float64)
    sw = sw / sw.sum()
    members = [rank] + [(rank - d) % ws for d in offs]
    tot = sum(sw[m] for m in members)
    return float(sw[rank] / tot), [(d, float(sw[(rank - d) % ws] / tot)) for d in offs]


def reduce_form(spec: ExchangeSpec, rank: int):
    """Non-gossip results all have the form ``a_r·x_r + Σ_p b_p·x_p``.
    Returns ``(a_r, b)``: equal → (0, 1/ws); shard → (0, w); weighted →
    (w − (1−w)/(ws−1), (1−w)/(ws−1))."""
    ws = spec.world_size
    if ws == 1:
        return 0.0, [1.0]
    if spec.aggregation_type == "equal":
        return 0.0, [1.0 / ws] * ws
    if spec.aggregation_type == "shard":
        sw = np.asarray(spec.shard_weights if spec.shard_weights is not None else [1.0 / ws] * ws, np.float64)
        sw = sw / sw.sum()
        return 0.0, [float(x) for x in sw]
    o = (1.0 - spec.local_weight) / (ws - 1)
    return spec.local_weight - o, [o] * ws


class DistExchanger:
    """The exchange over ``torch.distributed`` point-to-point / collectives.

    CPU/gloo realisation of every variant, and the "only calls NCCL" baseline on GPU.
    Works on any tensor (flat slab or a single parameter)."""

    def __init__(self, spec: ExchangeSpec, rank: int, group=None):
        self.spec, self.rank, self.group = spec, rank, group
        self.launches = 0

    @torch.no_grad()
    def exchange_(self, t: torch.Tensor) -> torch.Tensor:
        sp, ws, r = self.spec, self.spec.world_size, self.rank
        if ws == 1:
            return t
        if sp.is_gossip():
            c_self, nbrs = gossip_coefficients(sp, r)
            recvs, ops = [], []
            for d, _ in nbrs:
                if (r - d) % ws == r:                 # peer aliases self (ws==2 double ring)
                    recvs.append(t.clone())
                    continue
                buf = torch.empty_like(t)
                recvs.append(buf)
                ops.append(dist.P2POp(dist.irecv, buf, (r - d) % ws, group=self.group))
                ops.append(dist.P2POp(dist.isend, t, (r + d) % ws, group=self.group))
            if ops:
                for w in dist.batch_isend_irecv(ops):
                    w.wait()
            t.mul_(c_self)
            for buf, (_, c) in zip(recvs, nbrs):
                t.add_(buf, alpha=c)
            return t
        a, b = reduce_form(sp, r)
        if a == 0.0:
            t.mul_(b[r])
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        else:
            own = t.clone()
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
            t.mul_(b[r]).add_(own, alpha=a)
        return t

    # -- engine-level entry points (same names as the fused exchanger) ---------
    def exchange_grads(self, engine, apply_update: bool = False) -> None:
        self.exchange_(engine.flat.grads)
        if apply_update:
            engine.flat.optimizer_step()
            engine.params_changed()

    def exchange_weights(self, engine) -> None:
        self.exchange_(engine.flat.params)
        engine.params_changed()

    def broadcast_state(self, engine, src: int = 0) -> None:
        """Initial rank-0 broadcast of the whole ``state_dict`` incl. BN buffers
        (``main.py:40-42`` — C1; buffers are synced only here)."""
        if self.spec.world_size == 1:
            return
        sd = engine.state_dict()
        for k in sd:
            dist.broadcast(sd[k], src=src, group=self.group)
        engine.load_state_dict(sd)

    def set_shard_weights(self, w) -> None:
        self.spec.shard_weights = list(w)

    def gather_moments(self, engine) -> bool:
        """Optimizer moments are replicated on this path (every rank runs the whole update): nothing to do.
        (The fused exchanger shards them — ``FusedExchanger.gather_moments``.)"""
        return False

    def close(self) -> None:
        pass

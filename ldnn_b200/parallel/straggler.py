"""Straggler time limit (``--time_limit``) — the feature the reference intends but cannot reach.

Upstream: a rank that has finished its local epochs raises ``epoch_finished_signal``; ranks still training
notice it, start a clock, and ``break`` once ``time_limit`` seconds have passed ("first finisher waits
``time_limit`` s, then everyone moves on", ``Balanced All-Reduce/main.py:90``, ``trainer.py:42-44,112-119,
134-139``).  There the signal is only ever raised after the loop and every local epoch contains blocking
collectives, so it never fires.

Here it is real, for the schedule where it is meaningful: ``--sync global_epoch`` with uneven shards, where a
local epoch is a run of purely local steps and small-shard ranks finish early.

* A rank that completes a local epoch **announces** it: one word per rank, carrying a strictly increasing
  ticket ``(global epoch, local epoch)``, so nothing is ever reset and stale announcements cannot fire.
* The others **poll** between steps without a host-device sync.  Once any peer's ticket has reached the
  current one they start a clock; ``time_limit`` seconds later they stop stepping and go to the epoch's
  (collective) metric gather with the steps they have — nobody can deadlock, because the only collective
  sits after the loop and every rank reaches it.

Two transports behind one interface:

``SymmetricFlag``  GPU + fused exchange: the word lives in every rank's signal pad (the K-BAR storage, peer-mapped
                   symmetric memory); announcing = a peer store into each pad; polling = an async copy of the
                   local pad's words into pinned host memory, checked one poll later (no synchronisation).
``StoreFlag``      CPU / gloo (and any job without the fused exchanger): a key in the process group's store.
"""
from __future__ import annotations

import time
from typing import Optional

import torch
import torch.distributed as dist


def ticket(global_epoch: int, local_epoch: int) -> int:
    return global_epoch * 4096 + local_epoch + 1


class _Base:
    def __init__(self, rt, time_limit: float):
        self.rt, self.time_limit = rt, float(time_limit)
        self.cur = 0
        self.t_first: Optional[float] = None
        self.cut_short = False

    def begin(self, tk: int) -> None:
        """Start of a local epoch with ticket `tk`."""
        self.cur, self.t_first, self.cut_short = tk, None, False

    def _someone_finished(self) -> bool:
        raise NotImplementedError

    def should_stop(self) -> bool:
        """Called between steps.  True once a peer finished this local epoch more than `time_limit` s ago."""
        if self.rt.world_size == 1:
            return False
        if self.t_first is None:
            if self._someone_finished():
                self.t_first = time.time()
            else:
                return False
        if time.time() - self.t_first >= self.time_limit:
            self.cut_short = True
            return True
        return False

    def announce(self) -> None:
        raise NotImplementedError


class StoreFlag(_Base):
    """Announcements as keys of the default process group's store (TCPStore): works on CPU/gloo."""

    def __init__(self, rt, time_limit: float):
        super().__init__(rt, time_limit)
        self.store = None
        if rt.world_size > 1 and dist.is_initialized():
            try:
                self.store = dist.distributed_c10d._get_default_store()
            except Exception:
                self.store = None

    def announce(self) -> None:
        if self.store is not None:
            self.store.set(f"ldnn/finished/{self.cur}", str(self.rt.rank))

    def _someone_finished(self) -> bool:
        return self.store is not None and bool(self.store.check([f"ldnn/finished/{self.cur}"]))


class SymmetricFlag(_Base):
    """Announcements as words of the peer-mapped signal pad (see FusedExchanger.flag_views)."""

    def __init__(self, rt, time_limit: float, exchanger):
        super().__init__(rt, time_limit)
        self.peer_words, self.mine = exchanger.flag_views()     # [per peer: its slot for MY rank], my own slots
        self.host = torch.zeros(self.mine.numel(), dtype=torch.int32).pin_memory()
        self.ev: Optional[torch.cuda.Event] = None
        self.stream = torch.cuda.Stream(rt.device)

    def announce(self) -> None:
        cur = torch.cuda.current_stream(self.rt.device)
        self.stream.wait_stream(cur)                 # "finished" means the epoch's last step has executed
        with torch.cuda.stream(self.stream):
            for w in self.peer_words:
                w.fill_(self.cur)

    def _someone_finished(self) -> bool:
        hit = False
        if self.ev is not None and self.ev.query():      # the copy issued at the previous poll has landed
            hit = int(self.host.max()) >= self.cur
            self.ev = None
        if self.ev is None and not hit:
            with torch.cuda.stream(self.stream):
                self.host.copy_(self.mine, non_blocking=True)
                self.ev = torch.cuda.Event()
                self.ev.record(self.stream)
        return hit


def make_straggler(cfg, rt, exchanger):
    """None when the limit cannot apply (one rank; per-step schedule, where every step is a collective)."""
    if rt.world_size == 1 or cfg.sync == "step" or cfg.time_limit is None or cfg.time_limit < 0:
        return None
    if hasattr(exchanger, "flag_views") and rt.is_cuda:
        return SymmetricFlag(rt, cfg.time_limit, exchanger)
    return StoreFlag(rt, cfg.time_limit)

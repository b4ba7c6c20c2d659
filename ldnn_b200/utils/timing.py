"""Device-side timing helpers: CUDA events on the launching stream, max over ranks.

The reference times with host ``time.time()`` and no device sync
(``trainer.py:39,179-180``; ``dataloader.py:120,134-136``).  Everything reported by
this framework is event-timed after a synchronize and reduced with MAX over ranks.
"""
from __future__ import annotations

import subprocess
import threading
import time
from contextlib import contextmanager
from typing import List, Optional

import torch
import torch.distributed as dist


class PhaseTrace:
    """Host-side phase timeline (train / validate / exchange per epoch) in Chrome-trace format — open the file in
    chrome://tracing or Perfetto.  `nsys` is not part of this image; the phases that matter end in a device
    synchronisation (metrics are read back), so host timestamps at this granularity are device-accurate.
    Enabled by ``--trace PATH`` (one ``PATH.rank<r>.json`` per rank) or ``LDNN_TRACE=PATH``."""

    def __init__(self):
        self.enabled = False
        self.events: List[dict] = []
        self._t0 = time.perf_counter()

    def enable(self) -> None:
        self.enabled, self.events, self._t0 = True, [], time.perf_counter()

    def add(self, name: str, t0: float, t1: float, cat: str = "phase", **args) -> None:
        if self.enabled:
            self.events.append({"name": name, "cat": cat, "ph": "X", "ts": 1e6 * (t0 - self._t0),
                                "dur": 1e6 * (t1 - t0), "tid": 0, "args": args})

    def dump(self, path: str, rank: int = 0) -> Optional[str]:
        if not self.enabled:
            return None
        import json
        import os
        out = f"{path}.rank{rank}.json"
        os.makedirs(os.path.dirname(os.path.abspath(out)), exist_ok=True)
        with open(out, "w") as f:
            json.dump({"traceEvents": [dict(e, pid=rank) for e in self.events], "displayTimeUnit": "ms"}, f)
        return out


TRACE = PhaseTrace()


@contextmanager
def nvtx_range(name: str):
    """NVTX range around a phase (shows up in nsys/ncu timelines) + an entry in the phase trace when enabled."""
    on = torch.cuda.is_available()
    if on:
        torch.cuda.nvtx.range_push(name)
    t0 = time.perf_counter()
    try:
        yield
    finally:
        if on:
            torch.cuda.nvtx.range_pop()
        TRACE.add(name, t0, time.perf_counter())


class CudaTimer:
    def __init__(self, device=None):
        self.device = device
        self.start_ev = torch.cuda.Event(enable_timing=True)
        self.stop_ev = torch.cuda.Event(enable_timing=True)

    def start(self):
        torch.cuda.synchronize(self.device)
        self.start_ev.record()

    def stop(self) -> float:
        self.stop_ev.record()
        torch.cuda.synchronize(self.device)
        return self.start_ev.elapsed_time(self.stop_ev)  # ms


def max_over_ranks(value: float, device) -> float:
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def barrier_sync(device=None):
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.barrier()
    if torch.cuda.is_available():
        torch.cuda.synchronize(device)


def flush_l2(device, nbytes: int = 256 << 20, _cache={}):
    """Write a buffer larger than the 126 MB L2 (timing hygiene between iterations)."""
    key = (str(device), nbytes)
    if key not in _cache:
        _cache[key] = torch.empty(nbytes, dtype=torch.uint8, device=device)
    _cache[key].zero_()


class ClockSampler:
    """Samples ``nvidia-smi`` SM clocks + throttle reasons while a timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int = 0, period_s: float = 0.2):
        self.idx, self.period = gpu_index, period_s
        self.samples: List[List[str]] = []
        self._stop = threading.Event()
        self._th: Optional[threading.Thread] = None

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.idx}", f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                row = [c.strip() for c in out.strip().split("\n")[0].split(",")]
                if len(row) >= 7:
                    self.samples.append(row)
            except Exception:
                pass
            self._stop.wait(self.period)

    def __enter__(self):
        self._th = threading.Thread(target=self._run, daemon=True)
        self._th.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._th is not None:
            self._th.join(timeout=6)

    def summary(self) -> dict:
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.samples:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except ValueError:
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}

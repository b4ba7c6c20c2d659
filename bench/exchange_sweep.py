#!/usr/bin/env python
"""Gradient-exchange bandwidth sweep 1 KB – 1 GB (BASELINE config 5): fused P2P all-reduce / ring /
double-ring vs NCCL all_reduce on the same buffers.  Launch with torchrun (one process per GPU).
Device-timed (CUDA events), max over ranks; algorithmic bytes = payload size; bus bandwidth =
2*(ws-1)/ws * bytes / time (the all-reduce convention)."""
import argparse
import ctypes as C
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


class _Flat:
    pass


class _Eng:
    """Minimal engine facade: flat fp32 slabs in symmetric memory."""

    def __init__(self, n, dev):
        from ldnn_b200.ops import native_ops as K
        self._symm, self.flat, self.shadow = {}, _Flat(), None
        for name in ("params", "grads"):
            t, h, p = K.symm_alloc(n * 4, dev)
            self._symm[name] = (t, h, p)
            setattr(self.flat, name, t.view(torch.float32))
        self.flat.total, self.flat.m, self.flat.v = n, None, None

    def symm_handle(self, name):
        return self._symm[name][1], self._symm[name][2], self.flat.total * 4

    def params_changed(self):
        pass


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max_mb", type=int, default=1024)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--blocks", type=int, default=64)   # the sweep has the GPU to itself: 64 CTAs for every kernel
    ap.add_argument("--out", default="gpurun_out/exchange_sweep.json")
    a = ap.parse_args()
    from ldnn_b200.launch import init_runtime, shutdown
    from ldnn_b200.parallel.exchange import ExchangeSpec
    from ldnn_b200.parallel.fused import FusedExchanger
    from ldnn_b200.utils.timing import max_over_ranks
    rt = init_runtime("nccl", seed=0)
    ws, dev = rt.world_size, rt.device
    nmax = a.max_mb * (1 << 20) // 4
    eng = _Eng(nmax, dev)
    eng.flat.grads.normal_()
    rows = []
    sizes = [1 << k for k in range(10, 31) if (1 << k) <= a.max_mb << 20]
    for topo in ["allreduce", "ring", "double_ring", "nccl"]:
        ex = None
        if topo != "nccl":
            ex = FusedExchanger(ExchangeSpec(topo, "allreduce", "equal", 0.5, None, ws), rt, eng, comm_blocks=a.blocks)
        for nbytes in sizes:
            n = max(nbytes // 4, 4 * ws)
            if ex is not None:
                ex.n = n
            buf = eng.flat.grads[:n]

            def one():
                if ex is None:
                    dist.all_reduce(buf)
                    buf.mul_(1.0 / ws)            # the reference's separate scale kernel
                else:
                    ex._launch_reduce(ex._exchange_struct("grads", 0))
            iters = a.iters if nbytes <= (64 << 20) else max(3, a.iters // 4)
            for _ in range(3):
                one()
            dist.barrier(); torch.cuda.synchronize()
            t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0.record()
            for _ in range(iters):
                one()
            t1.record(); torch.cuda.synchronize()
            ms = max_over_ranks(t0.elapsed_time(t1) / iters, dev)
            rows.append({"impl": topo, "bytes": n * 4, "us": ms * 1e3, "algbw_GBs": n * 4 / ms / 1e6,
                         "busbw_GBs": 2 * (ws - 1) / ws * n * 4 / ms / 1e6, "world": ws})
            buf.normal_()
        if ex is not None:
            ex.check(); ex.close()
    if rt.rank == 0:
        os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
        with open(a.out, "w") as f:
            json.dump(rows, f, indent=1)
        print(f"{'bytes':>12} " + " ".join(f"{t:>14}" for t in ["allreduce", "ring", "double_ring", "nccl"]) + "   (busbw GB/s | us)")
        for nb in sorted({r["bytes"] for r in rows}):
            cells = []
            for t in ["allreduce", "ring", "double_ring", "nccl"]:
                r = [x for x in rows if x["impl"] == t and x["bytes"] == nb][0]
                cells.append(f"{r['busbw_GBs']:7.1f}|{r['us']:6.0f}")
            print(f"{nb:>12} " + " ".join(f"{c:>14}" for c in cells))
    shutdown(rt)


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Gradient-exchange bandwidth sweep 1 KB – 1 GB (BASELINE config 5): fused P2P all-reduce / ring /
double-ring (and the in-switch NVLS all-reduce when LDNN_NVLS=1 gives a multicast mapping) vs NCCL all_reduce on
the same buffers.  Launch with torchrun (one process per GPU).

Two modes per size:
  reduce  the bare exchange: ours = one kernel (reduce + 1/W scale); NCCL arm = dist.all_reduce + the reference's
          separate scale kernel (Balanced All-Reduce/communication.py:24-25)
  update  the fused product form: ours = ONE kernel (reduce + scale + Adam on the owned slice + push of the updated
          parameters); NCCL arm = all_reduce + scale + torch's fused Adam over the whole slab.

Device-timed (CUDA events), max over ranks; algorithmic bytes = payload size; bus bandwidth =
2*(ws-1)/ws * bytes / time (the all-reduce convention); `link_frac` = busbw / 770 GB/s (the measured peer-copy
rate per direction per GPU, B200_PROFILING.md) = achieved fraction of the NVLink roofline; for `update` the roofline
is the slower of that link time and the HBM time of the tail.  SM clocks and throttle reasons are sampled during
the whole sweep and written into the same output files as the table."""
import argparse
import ctypes as C
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

LINK_GBS = 770.0     # measured peer copy, per direction per GPU
HBM_GBS = 6554.6     # MEASURED_PEAKS.json


class _Flat:
    pass


class _Eng:
    """Minimal engine facade: flat fp32 slabs in symmetric memory (multicast-capable when LDNN_NVLS=1)."""

    def __init__(self, n, dev, with_moments):
        from ldnn_b200.ops import native_ops as K
        self._symm, self.flat, self.shadow = {}, _Flat(), None
        for name in ("params", "grads"):
            if K.nvls_requested():
                t, h, p = K.symm_alloc_multicast(n * 4, dev)
            else:
                t, h, p = K.symm_alloc(n * 4, dev)
            self._symm[name] = (t, h, p)
            setattr(self.flat, name, t.view(torch.float32))
        self.flat.total = n
        self.flat.m = torch.zeros(n, device=dev) if with_moments else None
        self.flat.v = torch.zeros(n, device=dev) if with_moments else None

    def symm_handle(self, name):
        return self._symm[name][1], self._symm[name][2], self.flat.total * 4

    def symm_tensor(self, name):
        return self._symm[name][0]

    def params_changed(self):
        pass


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max_mb", type=int, default=1024)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--blocks", type=int, default=64)   # the sweep has the GPU to itself: 64 CTAs for every kernel
    ap.add_argument("--modes", default="reduce,update")
    ap.add_argument("--out", default="gpurun_out/exchange_sweep.json")
    ap.add_argument("--impls", default="allreduce,allreduce_nvls,ring,double_ring,nccl")
    ap.add_argument("--min_kb", type=int, default=1)
    ap.add_argument("--nvlink_at_mb", type=int, default=0,
                    help="at this payload also read the driver's NVLink byte counters around 10 calls (0 = off; on the "
                         "B200 pool used here `nvidia-smi nvlink -gt d` reads 0 KiB on every link, so it is off)")
    a = ap.parse_args()
    from ldnn_b200.engine import OptimConfig
    from ldnn_b200.launch import init_runtime, shutdown
    from ldnn_b200.ops import _lib as L
    from ldnn_b200.parallel.exchange import ExchangeSpec
    from ldnn_b200.parallel.fused import FusedExchanger
    from ldnn_b200.utils.timing import ClockSampler, max_over_ranks
    rt = init_runtime("nccl", seed=0)
    ws, dev = rt.world_size, rt.device
    nmax = a.max_mb * (1 << 20) // 4
    modes = [m for m in a.modes.split(",") if m]
    eng = _Eng(nmax, dev, "update" in modes)
    eng.flat.grads.normal_()
    eng.flat.params.normal_()
    opt = OptimConfig(kind="adam", lr=1e-3)
    rows = []
    sizes = [1 << k for k in range(10, 31) if (a.min_kb << 10) <= (1 << k) <= a.max_mb << 20]
    # "allreduce" = the two-shot P2P kernel; "allreduce_nvls" = the in-switch kernel on the same (multicast-capable) slabs
    impls = [t for t in a.impls.split(",") if t]
    clk = ClockSampler(dev.index or 0, period_s=0.1)
    clk.__enter__()
    nvls_seen = False
    for mode in modes:
        for topo in impls:
            ex = None
            if topo != "nccl":
                ex = FusedExchanger(ExchangeSpec(topo.replace("_nvls", ""), "allreduce", "equal", 0.5, None, ws), rt, eng,
                                    comm_blocks=a.blocks)
                if topo == "allreduce":
                    ex.nvls = False                   # two-shot kernel (peer loads / stores), whatever the slabs allow
                elif topo == "allreduce_nvls":
                    ex.nvls = bool(getattr(ex, "nvls_capable", ex.nvls))   # measure it whatever the size policy says
                    if not ex.nvls:                   # no multicast mapping on this host: no such column
                        ex.close()
                        continue
                    nvls_seen = True
            for nbytes in sizes:
                n = max(nbytes // 4, 4 * ws)
                if ex is not None:
                    ex.n = n
                buf = eng.flat.grads[:n]
                step = [0]
                topt = None
                if mode == "update" and ex is None:
                    pp = torch.nn.Parameter(eng.flat.params[:n])
                    pp.grad = buf
                    topt = torch.optim.Adam([pp], lr=opt.lr, fused=True)     # torch's fused (multi-tensor) Adam

                def one():
                    step[0] += 1
                    if ex is None:
                        dist.all_reduce(buf)
                        buf.mul_(1.0 / ws)            # the reference's separate scale kernel
                        if topt is not None:
                            topt.step()
                    elif mode == "update":
                        ex._launch_reduce(ex._exchange_struct("grads", L.X_UPDATE, L.make_optim(opt, step[0], opt.lr)))
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
                nb = n * 4
                bus = 2 * (ws - 1) / ws * nb / ms / 1e6
                t_link = 2 * (ws - 1) / ws * nb / (LINK_GBS * 1e9)
                t_hbm = ((3 + 7 / ws) * nb / (HBM_GBS * 1e9)) if mode == "update" else (3 * nb / (HBM_GBS * 1e9))
                rows.append({"mode": mode, "impl": topo, "bytes": nb, "us": ms * 1e3, "algbw_GBs": nb / ms / 1e6,
                             "busbw_GBs": bus, "link_frac": bus / LINK_GBS if ws > 1 else None,
                             "roofline_frac": max(t_link, t_hbm) / (ms / 1e3), "world": ws,
                             "nvls": topo == "allreduce_nvls",
                             "nvlink": nvlink_probe(one, nb, dev, rt) if (a.nvlink_at_mb and nb == a.nvlink_at_mb << 20) else None})
                buf.normal_()
            if ex is not None:
                ex.check(); ex.close()
    clk.__exit__(None, None, None)
    clocks = clk.summary()
    if rt.rank == 0:
        os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
        with open(a.out, "w") as f:
            json.dump({"world": ws, "clocks": clocks, "nvls": nvls_seen, "link_GBs": LINK_GBS, "rows": rows}, f, indent=1)
        print(f"# exchange sweep, {ws} x B200, blocks={a.blocks}, NVLS all-reduce: {nvls_seen}")
        print(f"# clocks during the sweep: {json.dumps(clocks)}")
        for mode in modes:
            print(f"## mode = {mode}   cells: busbw GB/s | us | fraction of roofline")
            print(f"{'bytes':>12} " + " ".join(f"{t:>22}" for t in impls))
            for nb in sorted({r["bytes"] for r in rows}):
                cells = []
                for t in impls:
                    rr = [x for x in rows if x["impl"] == t and x["bytes"] == nb and x["mode"] == mode]
                    if not rr:
                        cells.append("-")
                        continue
                    r = rr[0]
                    cells.append(f"{r['busbw_GBs']:7.1f}|{r['us']:6.0f}|{r['roofline_frac']:5.2f}")
                print(f"{nb:>12} " + " ".join(f"{c:>22}" for c in cells))
        probes = [r for r in rows if r.get("nvlink")]
        if probes:
            print("## NVLink traffic of rank 0's GPU (nvidia-smi nvlink -gt d, all links), per call, vs the algorithmic "
                  "2(W-1)/W x payload each way")
            for r in probes:
                v = r["nvlink"]
                if "error" in v:
                    print(f"{r['mode']:>7} {r['impl']:>15}: counters unavailable ({v['error']})")
                    continue
                print(f"{r['mode']:>7} {r['impl']:>15} {r['bytes']:>11} B: tx {v['tx_MB_per_call']:8.1f} MB  rx "
                      f"{v['rx_MB_per_call']:8.1f} MB  algorithmic {v['algorithmic_MB_each_way']:8.1f} MB")
    shutdown(rt)


def _nvlink_kib(idx):
    """(tx KiB, rx KiB) summed over the links of GPU `idx`, from the driver's NVLink data counters."""
    import re
    import subprocess
    out = subprocess.run(["nvidia-smi", "nvlink", "-gt", "d", "-i", str(idx)], capture_output=True, text=True,
                         timeout=20).stdout
    tx = sum(int(v) for v in re.findall(r"Data Tx:\s*(\d+)\s*KiB", out))
    rx = sum(int(v) for v in re.findall(r"Data Rx:\s*(\d+)\s*KiB", out))
    return tx, rx


def nvlink_probe(one, nbytes, dev, rt, calls=10):
    """Bytes that crossed rank 0's NVLink ports during `calls` exchanges (collective: every rank runs them)."""
    try:
        dist.barrier(); torch.cuda.synchronize()
        before = _nvlink_kib(dev.index or 0) if rt.rank == 0 else (0, 0)
        dist.barrier()
        for _ in range(calls):
            one()
        torch.cuda.synchronize(); dist.barrier()
        after = _nvlink_kib(dev.index or 0) if rt.rank == 0 else (0, 0)
        ws = rt.world_size
        return {"tx_MB_per_call": (after[0] - before[0]) * 1024 / calls / 1e6,
                "rx_MB_per_call": (after[1] - before[1]) * 1024 / calls / 1e6,
                "algorithmic_MB_each_way": 2 * (ws - 1) / ws * nbytes / 1e6, "calls": calls}
    except Exception as e:     # counters not exposed on this host
        return {"error": f"{type(e).__name__}: {e}"[:200]}


if __name__ == "__main__":
    main()

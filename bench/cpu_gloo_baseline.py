#!/usr/bin/env python
"""BASELINE config 1: Balanced All-Reduce on CPU / gloo, world_size 4, synthetic data — the plumbing
baseline.  Runs the unmodified reference (`baseline/_ref`) and this framework's CPU path for a few steps
and prints samples/s for both (wall clock, max over ranks).  Launch:
    python -m torch.distributed.run --nproc-per-node 4 --master-addr 127.0.0.1 bench/cpu_gloo_baseline.py --steps 3
"""
import argparse
import json
import os
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--batch", type=int, default=64)
    a = ap.parse_args()
    dist.init_process_group("gloo")
    rank, ws = dist.get_rank(), dist.get_world_size()
    torch.set_num_threads(max(1, (os.cpu_count() or 4) // ws))
    g = torch.Generator().manual_seed(1 + rank)
    x = torch.randn(a.steps * a.batch, 3, 32, 32, generator=g)
    y = torch.randint(0, 10, (a.steps * a.batch,), generator=g)
    out = {}

    # ---- reference (unmodified modules from baseline/_ref) ----
    vdir = os.path.join(ROOT, "baseline", "_ref", "Balanced All-Reduce")
    if os.path.isdir(vdir):
        sys.path.insert(0, vdir)
        import torch.nn as nn
        from torch.utils.data import DataLoader, TensorDataset
        from model import EnhancedCNNModel as RefModel
        from trainer import train_local_epoch as ref_epoch
        from communication import average_gradients_equal as ref_avg
        torch.manual_seed(0)
        m = RefModel()
        opt = torch.optim.Adam(m.parameters(), lr=1e-3)
        sched = torch.optim.lr_scheduler.StepLR(opt, step_size=25)
        ld = DataLoader(TensorDataset(x, y), batch_size=a.batch, shuffle=False)
        dist.barrier()
        t0 = time.perf_counter()
        ref_epoch(m, ld, nn.CrossEntropyLoss(), opt, torch.device("cpu"), sched)
        ref_avg(m, ws)
        dist.barrier()
        out["reference_s"] = time.perf_counter() - t0
        sys.path.pop(0)
        for k in ("model", "trainer", "communication", "dataloader", "validator"):
            sys.modules.pop(k, None)
        del m, opt

    # ---- ours (TorchEngine on flat slabs + DistExchanger) ----
    from ldnn_b200.engine import OptimConfig, TorchEngine
    from ldnn_b200.models import enhanced_cnn_spec
    from ldnn_b200.parallel.exchange import DistExchanger, ExchangeSpec
    torch.manual_seed(0)
    eng = TorchEngine(enhanced_cnn_spec(), "cpu", OptimConfig())
    ex = DistExchanger(ExchangeSpec("allreduce", "allreduce", "equal", 0.5, None, ws), rank)
    dist.barrier()
    t0 = time.perf_counter()
    for i in range(a.steps):
        eng.train_step(x[i * a.batch:(i + 1) * a.batch], y[i * a.batch:(i + 1) * a.batch])
    ex.exchange_grads(eng)
    dist.barrier()
    out["ours_s"] = time.perf_counter() - t0
    t = torch.tensor([out.get("reference_s", 0.0), out["ours_s"]], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        n = a.steps * a.batch * ws
        print(json.dumps({"config": "Balanced All-Reduce, CPU/gloo", "world_size": ws, "steps": a.steps,
                          "threads_per_rank": torch.get_num_threads(),
                          "reference_samples_per_s": (n / float(t[0])) if float(t[0]) > 0 else None,
                          "ours_samples_per_s": n / float(t[1])}))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Headline benchmark: train samples/s of the repo model (EnhancedCNN, batch 64 per GPU,
Adam) under a variant's schedule, on N GPUs of one node.

    python bench.py --gpus N --steps K --warmup W [--topology allreduce|ring|double_ring]
                    [--sync global_epoch|step] [--dtype tf32|bf16] [--impl ours|reference]

Timed region = exactly K training steps (forward + backward + optimizer update on a fresh
batch each step) followed by the variant's exchange — the reference's schedule, in which a
"global epoch" is a run of local steps and ONE exchange (Balanced All-Reduce/trainer.py:46-47,
141-150).  ``--sync step`` measures the product schedule instead (fused exchange+update every
step).  `value` is device-timed (CUDA events, barrier + synchronize on both sides, max over
ranks) on device-resident synthetic data; `e2e` times the same K steps through the public
engine API with a pinned-host → device copy of every step's batch and a device → host read of
every step's loss.  Weak scaling: per-GPU batch is fixed at 64.

Precision: the default ``--dtype tf32`` is the reference's own — fp32 tensors with TF32 tensor-core
convolutions (torch's default ``cudnn.allow_tf32``; no autocast anywhere upstream) — run on our
``tcgen05 kind::tf32`` kernels; both arms print ``"dtype": "tf32"``.  The faster bf16 engine is reported
as a second line under ``extra.bf16`` and never as the headline.

After the headline the same process measures, for a few steps each and under ``extra``: the other
topologies on the reference's schedule, the per-step fused exchange+update schedule for all three
topologies, and an on-device check of every exchange kernel against the fp64 mixing-matrix oracle.

`--impl reference` runs the UNMODIFIED reference (baseline/_ref, see baseline/install_ref.py)
through its own `train_local_epoch` + `average_gradients_equal` on the NCCL backend; for
``--topology ring|double_ring`` (and under its ``extra``) the reference's model and training loop with
its neighbour exchange restated on NCCL P2P (baseline/ring_nccl.py), labelled ``restated``.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BATCH = 64
METRIC = "train samples/sec (EnhancedCNN, batch 64/GPU, Adam; K steps + 1 exchange)"
DTYPE_HEADLINE = "tf32"          # both arms: fp32 storage, TF32 tensor-core convolutions
L2_NOTE = ("per-step working set (params+moments+grads ~0.9 GB) and a 16-batch input pool exceed the 126 MB L2; "
           "a fresh batch every step")


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=100)
    p.add_argument("--warmup", type=int, default=10)
    p.add_argument("--impl", choices=["ours", "reference"], default="ours")
    p.add_argument("--topology", choices=["allreduce", "ring", "double_ring"], default="allreduce")
    p.add_argument("--sync", choices=["global_epoch", "step"], default="global_epoch")
    p.add_argument("--model", choices=["enhanced_cnn", "resnet50"], default="enhanced_cnn")
    p.add_argument("--batch", type=int, default=BATCH)
    p.add_argument("--exchange_impl", choices=["fused", "dist"], default="fused")
    p.add_argument("--engine", choices=["native", "torch"], default="native")
    p.add_argument("--dtype", choices=["tf32", "bf16"], default="tf32",
                   help="native engine precision: tf32 = fp32 storage + TF32 tensor-core convolutions (what the "
                        "reference runs: fp32 tensors, torch's default cudnn.allow_tf32); bf16 = bf16 storage/MMA")
    p.add_argument("--no_overlap", action="store_true", help="--sync step: exchange after backward instead of per bucket")
    p.add_argument("--comm_blocks", type=int, default=0)
    p.add_argument("--shard_ratios", type=float, nargs="+", default=None,
                   help="Disbalanced variants: per-rank shard sizes (e.g. 4 2 1 1).  --sync global_epoch: rank r runs "
                        "K*ratio_r/max(ratio) steps before the one exchange; --sync step: every rank steps K times "
                        "with per-rank batch = batch*ratio_r/max(ratio)")
    p.add_argument("--aggregation_type", choices=["equal", "weighted", "shard"], default="equal")
    p.add_argument("--no_extra", action="store_true", help="headline only (skip the extra topologies / schedules)")
    p.add_argument("--extra_steps", type=int, default=30, help="timed steps of each `extra` measurement")
    return p.parse_args()


def env_world():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


def emit(d):
    print(json.dumps(d), flush=True)


def ratios_for(a, world):
    """Per-rank shard ratios (Disbalanced variants) — all ones for the Balanced ones."""
    if not a.shard_ratios:
        return [1.0] * world
    if len(a.shard_ratios) != world:
        raise SystemExit(f"--shard_ratios needs {world} values")
    return [float(v) for v in a.shard_ratios]


def config_block(a, world, global_batch, topology, sync):
    """The `config` object — the SAME key set from both arms (arm-specific facts go under `extra`)."""
    return {"model": a.model, "global_batch": int(global_batch), "per_gpu_batch": a.batch, "optimizer": "adam",
            "parallelism": f"dp{world}", "topology": topology, "sync": sync,
            "aggregation_type": a.aggregation_type, "shard_ratios": a.shard_ratios, "l2": L2_NOTE}


# =========================================================================== ours
class _Data:
    """Pool of distinct synthetic batches of dataloader.py's shape (device-resident + pinned host twins)."""

    def __init__(self, dev, rank, hw, ncls, pool=16):
        self.dev, self.rank, self.hw, self.ncls, self.pool = dev, rank, hw, ncls, pool
        self._by_batch = {}

    def get(self, batch):
        import torch
        if batch not in self._by_batch:
            g = torch.Generator(device="cpu").manual_seed(1234 + self.rank)
            xs = [torch.randn(batch, 3, self.hw, self.hw, generator=g).to(self.dev) for _ in range(self.pool)]
            ys = [torch.randint(0, self.ncls, (batch,), generator=g).to(self.dev) for _ in range(self.pool)]
            self._by_batch[batch] = (xs, ys, [x.cpu().pin_memory() for x in xs], [y.cpu().pin_memory() for y in ys])
        return self._by_batch[batch]


def measure(a, rt, engine, ex, data, sync, steps, warmup, with_e2e=True, ratios=None):
    """K steps of one (engine, exchanger, schedule) combination: device-timed value (+ end-to-end value).
    Returns a dict with samples/s (whole job), ms/step (max over ranks), launches, per-rank batch."""
    import torch
    import torch.distributed as dist
    from ldnn_b200.ops import _lib as L
    from ldnn_b200.utils.timing import barrier_sync, max_over_ranks
    dev, rank, world = rt.device, rt.rank, rt.world_size
    ratios = ratios or [1.0] * world
    share = ratios[rank] / max(ratios)
    sync_step = sync == "step"
    # Disbalanced shards: each rank's share of the work follows its ratio (the reference's uneven
    # random_split, Disbalanced */dataloader.py) and the exchange is weighted by n_r / sum(n).
    my_steps = steps if sync_step else max(1, round(steps * share))
    batch = max(1, round(a.batch * share)) if (sync_step and share != 1.0) else a.batch
    xs, ys, xs_pin, ys_pin = data.get(batch)
    pool = data.pool
    overlap = sync_step and not a.no_overlap and hasattr(engine, "train_step_sync") and hasattr(ex, "exchange_bucket")

    def step(x, y):
        if sync_step:
            if overlap:
                loss, _ = engine.train_step_sync(x, y, ex)          # bucketed exchange overlapped with backward
            else:
                loss, _ = engine.train_step(x, y, update=False)
                ex.exchange_grads(engine, apply_update=True)
        else:
            loss, _ = engine.train_step(x, y, update=True)
        return loss

    def epoch(k):
        for i in range(k):
            step(xs[i % pool], ys[i % pool])
        if not sync_step:
            ex.exchange_grads(engine, apply_update=False)

    def launches_now():
        return (L.launch_count() if L.available() else 0) + getattr(engine, "graph_kernel_launches", 0)

    epoch(max(warmup, 3))
    barrier_sync(dev)
    l0 = launches_now()
    barrier_sync(dev)
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    epoch(my_steps)
    t1.record()
    barrier_sync(dev)
    launches = launches_now() - l0        # eager launches + kernels replayed from the step's CUDA graphs
    ms = max_over_ranks(t0.elapsed_time(t1), dev)
    samples = float(batch * my_steps)                              # this rank's; summed over ranks below
    if world > 1:
        t = torch.tensor([samples], device=dev, dtype=torch.float64)
        dist.all_reduce(t)
        samples = float(t[0])
    out = {"value": samples / (ms / 1e3), "ms_per_step": ms / steps, "samples": samples, "launches": int(launches),
           "overlap": bool(overlap), "batch": batch, "steps": steps}
    if not with_e2e:
        return out

    # ---- end to end through the public API: pinned H2D of every batch, D2H of every loss ----
    copy_stream = torch.cuda.Stream(dev)
    host_loss = torch.zeros(2).pin_memory()

    def e2e_epoch(k):
        """Every step: H2D of its batch from pinned memory (prefetched one step ahead on a copy stream) and a
        D2H of its loss into pinned memory.  The host reads the loss of step i-1 after it has queued step i
        (how a logging trainer runs), so the launch of the next step is not held back by the read; all k
        losses are on the host before the region ends."""
        tot = 0.0
        cur = torch.cuda.current_stream()
        done = [None, None]
        with torch.cuda.stream(copy_stream):
            nxt = (xs_pin[0].to(dev, non_blocking=True), ys_pin[0].to(dev, non_blocking=True))
            ev = torch.cuda.Event(); ev.record(copy_stream)
        for i in range(k):
            cur.wait_event(ev)
            x, y = nxt
            x.record_stream(cur); y.record_stream(cur)      # allocated on the copy stream, consumed here
            if i + 1 < k:
                with torch.cuda.stream(copy_stream):
                    j = (i + 1) % pool
                    nxt = (xs_pin[j].to(dev, non_blocking=True), ys_pin[j].to(dev, non_blocking=True))
                    ev = torch.cuda.Event(); ev.record(copy_stream)
            loss = step(x, y)
            host_loss[i & 1:(i & 1) + 1].copy_(loss.detach().view(1), non_blocking=True)
            done[i & 1] = torch.cuda.Event(); done[i & 1].record(cur)
            if i > 0:
                done[(i - 1) & 1].synchronize()            # step i-1's loss is on the host
                tot += float(host_loss[(i - 1) & 1])
        done[(k - 1) & 1].synchronize()
        tot += float(host_loss[(k - 1) & 1])
        if not sync_step:
            ex.exchange_grads(engine, apply_update=False)
        return tot

    e2e_epoch(3)
    barrier_sync(dev)
    w0 = time.perf_counter()
    e2e_epoch(my_steps)
    barrier_sync(dev)
    e2e_s = max_over_ranks(time.perf_counter() - w0, dev)
    out["e2e"] = {"value": samples / e2e_s, "unit": "samples/s",
                  "h2d_bytes_per_step": batch * 3 * data.hw * data.hw * 4 + batch * 8, "d2h_bytes_per_step": 4}
    return out


def exchange_selfcheck(rt, engine, ex):
    """One exchange of the whole gradient slab on a known pattern, compared on the device with the fp64
    mixing-matrix oracle (x_r[i] = a_r + b_r*h(i)  =>  result_r[i] = (W a)_r + (W b)_r h(i))."""
    import numpy as np
    import torch
    from ldnn_b200.parallel.exchange import mixing_matrix
    dev, r, ws = rt.device, rt.rank, rt.world_size
    g = engine.flat.grads
    rng = np.random.default_rng(4321)
    av, bv = rng.standard_normal(ws), rng.standard_normal(ws)
    h = (torch.arange(g.numel(), device=dev, dtype=torch.float32) % 1021) / 1021.0
    g.copy_(av[r] + bv[r] * h)
    ex.exchange_grads(engine, apply_update=False)
    if hasattr(ex, "check"):
        ex.check()
    W = mixing_matrix(ex.spec)
    want = float((W @ av)[r]) + float((W @ bv)[r]) * h.double()
    err = float((g.double() - want).abs().max())
    g.zero_()
    t = torch.tensor([err], device=dev, dtype=torch.float64)
    if ws > 1:
        import torch.distributed as dist
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return {"max_abs_err": float(t[0]), "pass": bool(float(t[0]) < 1e-5)}


def run_ours(a):
    import torch
    from ldnn_b200.config import Config, resolve
    from ldnn_b200.engine import make_engine
    from ldnn_b200.launch import init_runtime, shutdown
    from ldnn_b200.parallel import bring_up, make_exchanger
    from ldnn_b200.utils.timing import ClockSampler

    rank, world, local = env_world()
    if world == 1:
        os.environ.setdefault("MASTER_PORT", str(29500 + os.getpid() % 2000))

    def make_cfg(topology, sync, dtype):
        return resolve(Config(topology=topology, sync=sync, model=a.model, batch_size=a.batch, backend="nccl",
                              engine=a.engine, dtype=(dtype if a.engine == "native" else "fp32"),
                              exchange_impl=a.exchange_impl, synthetic=True, quiet=True, seed=0,
                              aggregation_by="gradients", semantics="allreduce", apply_exchanged_grads=None,
                              comm_blocks=a.comm_blocks, aggregation_type=a.aggregation_type))

    cfg = make_cfg(a.topology, a.sync, a.dtype)
    rt = init_runtime("nccl", seed=0)
    dev = rt.device
    ratios = ratios_for(a, world)

    def exchanger_for(cfg_, engine_):
        ex_ = make_exchanger(cfg_, rt, engine_)
        if a.aggregation_type == "shard":
            ex_.set_shard_weights([v / sum(ratios) for v in ratios])
        return ex_

    engine, ex = bring_up(cfg, rt, make_engine)       # falls back to CUDA-IPC slabs if the NVLS bring-up fails
    if a.aggregation_type == "shard":
        ex.set_shard_weights([v / sum(ratios) for v in ratios])
    ex.broadcast_state(engine)
    data = _Data(dev, rank, engine.spec.in_hw, engine.spec.head().num_classes)

    # clocks / throttle reasons are sampled across BOTH timed regions of the headline (device-timed + end-to-end)
    clk = ClockSampler(dev.index or 0, period_s=0.05)
    clk.__enter__()
    head = measure(a, rt, engine, ex, data, a.sync, a.steps, a.warmup, with_e2e=True, ratios=ratios)
    clk.__exit__(None, None, None)
    if hasattr(ex, "check"):
        ex.check()
    hw = engine.spec.in_hw
    dtype_name = a.dtype if a.engine == "native" else "fp32"
    out = {
        "metric": METRIC if a.model == "enhanced_cnn" else METRIC.replace("EnhancedCNN, batch 64/GPU", f"ResNet-50 @{hw}, batch {a.batch}/GPU"),
        "value": head["value"], "unit": "samples/s", "n_gpus": world, "steps": a.steps,
        "warmup": max(a.warmup, 3), "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": dtype_name, "data": "synthetic", "impl": "ours",
        "config": config_block(a, world, round(head["samples"] / a.steps), a.topology, a.sync),
        "clocks": clk.summary(), "e2e": head["e2e"], "gpu_launches": head["launches"],
    }
    extra = {"dtype_detail": {"tf32": "fp32 activations/weights, tcgen05 kind::tf32 MMA, fp32 accumulate (the reference's "
                                      "numerics: fp32 tensors + torch-default TF32 convolutions)",
                              "bf16": "bf16 activations + bf16 weight copy, kind::f16 MMA, fp32 accumulate/master"}[dtype_name]
             if dtype_name in ("tf32", "bf16") else "torch fp32",
             "input": [3, hw, hw], "overlap": head["overlap"], "exchange": a.exchange_impl, "engine": a.engine,
             "launches_per_step": round(head["launches"] / max(1, a.steps), 1)}

    if not a.no_extra and a.engine == "native" and a.model == "enhanced_cnn":
        K2, W2 = max(5, a.extra_steps), 5
        clk2 = ClockSampler(dev.index or 0, period_s=0.05)
        clk2.__enter__()
        runs = {}
        topos = ["allreduce", "ring", "double_ring"] if world > 1 else [a.topology]
        for topo in topos:
            for sync in ("global_epoch", "step"):
                key = f"{topo}/{sync}"
                if (topo, sync) == (a.topology, a.sync):
                    runs[key] = {"value": head["value"], "ms_per_step": head["ms_per_step"], "headline": True}
                    if world > 1:
                        runs[key]["check"] = exchange_selfcheck(rt, engine, ex)
                    continue
                try:
                    ex2 = ex if topo == a.topology else exchanger_for(make_cfg(topo, sync, a.dtype), engine)
                    m = measure(a, rt, engine, ex2, data, sync, K2, W2, with_e2e=False, ratios=ratios)
                    runs[key] = {"value": m["value"], "ms_per_step": m["ms_per_step"], "steps": K2,
                                 "overlap": m["overlap"], "launches_per_step": round(m["launches"] / K2, 1)}
                    if world > 1 and sync == "global_epoch":
                        runs[key]["check"] = exchange_selfcheck(rt, engine, ex2)
                    if ex2 is not ex:
                        if hasattr(ex2, "check"):
                            ex2.check()
                        ex2.close()
                except Exception as e:  # an extra must never take the headline down
                    runs[key] = {"error": f"{type(e).__name__}: {e}"[:300]}
        extra["runs"] = runs
        # second precision: the same model on the other native engine (never the headline)
        other = "bf16" if a.dtype == "tf32" else "tf32"
        try:
            cfg_o = make_cfg(a.topology, a.sync, other)
            eng_o = make_engine(cfg_o, rt)
            eng_o.load_state_dict(engine.state_dict())
            ex_o = exchanger_for(cfg_o, eng_o)
            m = measure(a, rt, eng_o, ex_o, data, a.sync, K2, W2, with_e2e=True, ratios=ratios)
            extra[other] = {"value": m["value"], "ms_per_step": m["ms_per_step"], "steps": K2, "e2e": m["e2e"],
                            "launches_per_step": round(m["launches"] / K2, 1)}
            ex_o.close()
        except Exception as e:
            extra[other] = {"error": f"{type(e).__name__}: {e}"[:300]}
        clk2.__exit__(None, None, None)
        extra["clocks"] = clk2.summary()
    out["extra"] = extra
    if rank == 0:
        emit(out)
    ex.close()
    shutdown(rt)


# =========================================================================== reference
def run_reference(a):
    rank, world, local = env_world()
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    try:
        import install_ref
        note = install_ref.install(verbose=False)
    except Exception as e:  # pragma: no cover
        note = f"unavailable: {e}"
    if not os.path.isdir(os.path.join(ref_dir, "Balanced All-Reduce")):
        return emit({"impl": "reference", "unavailable": note})
    # the reference pins cuda:0 in every process (main.py:25): give each rank its own GPU
    os.environ["CUDA_VISIBLE_DEVICES"] = str(local)
    import torch
    if not torch.cuda.is_available():
        return emit({"impl": "reference", "unavailable": "no CUDA device"})
    import torch.distributed as dist
    import torch.nn as nn
    import torch.optim as optim
    from torch.optim.lr_scheduler import StepLR
    from torch.utils.data import DataLoader, TensorDataset
    import ring_nccl                                             # baseline/ring_nccl.py (restated ring exchange)
    from ldnn_b200.utils.timing import ClockSampler

    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29511")
    dist.init_process_group("nccl", rank=rank, world_size=world)
    device = torch.device("cuda:0")
    disb = bool(a.shard_ratios)
    ratios = ratios_for(a, world)
    share = ratios[rank] / max(ratios)

    def arm(topology, steps, warmup, sample_clocks):
        """One measurement of the reference's loop: its model + train_local_epoch for `steps` batches, then its
        exchange (AR: communication.average_gradients_equal, unmodified; rings: restated on NCCL P2P)."""
        EnhancedCNNModel, train_local_epoch, vname = ring_nccl.import_reference(ref_dir, topology, disb)
        if topology == "allreduce":
            sys.path.insert(0, os.path.join(ref_dir, vname))
            sys.modules.pop("communication", None)
            from communication import average_gradients_equal, average_gradients_weighted   # reference, unmodified
            sys.path.pop(0)
        torch.manual_seed(0)
        model = EnhancedCNNModel().to(device)

        def init_weights(m):                                        # main.py:33-37
            if isinstance(m, (nn.Conv2d, nn.Linear)):
                nn.init.xavier_uniform_(m.weight)
                if m.bias is not None:
                    nn.init.zeros_(m.bias)
        model.apply(init_weights)
        for t in model.state_dict().values():                       # main.py:40-42
            dist.broadcast(t, src=0)
        criterion = nn.CrossEntropyLoss()
        optimizer = optim.Adam(model.parameters(), lr=1e-3)
        scheduler = StepLR(optimizer, step_size=25)

        def loader(k):
            g = torch.Generator().manual_seed(1234 + rank)
            x = torch.randn(k * a.batch, 3, 32, 32, generator=g)
            y = torch.randint(0, 10, (k * a.batch,), generator=g)
            return DataLoader(TensorDataset(x, y), batch_size=a.batch, shuffle=False)

        import inspect
        # four of the six directories take two extra (unused) arguments: (..., device, world_size, scheduler, epoch)
        long_sig = "world_size" in inspect.signature(train_local_epoch).parameters

        def epoch(ld):
            if long_sig:
                train_local_epoch(model, ld, criterion, optimizer, device, world, scheduler, 0)
            else:
                train_local_epoch(model, ld, criterion, optimizer, device, scheduler)
            if topology == "allreduce":
                if a.aggregation_type == "weighted" and world > 1:
                    average_gradients_weighted(model, world, 0.5)
                else:
                    average_gradients_equal(model, world)
            else:
                ring_nccl.exchange_model(model, rank, world, topology, "gradients",
                                         "weighted" if a.aggregation_type == "weighted" else "equal", 0.5)

        my_steps = max(1, round(steps * share))
        warm, timed = loader(max(warmup, 3)), loader(my_steps)
        epoch(warm)
        dist.barrier(); torch.cuda.synchronize()
        clk = ClockSampler(local) if sample_clocks else None
        if clk:
            clk.__enter__()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        w0 = time.perf_counter()
        t0.record()
        epoch(timed)
        t1.record()
        dist.barrier(); torch.cuda.synchronize()
        wall = time.perf_counter() - w0
        ms = t0.elapsed_time(t1)
        if clk:
            clk.__exit__(None, None, None)
        t = torch.tensor([ms, wall * 1e3], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        s = torch.tensor([float(a.batch * my_steps)], dtype=torch.float64, device=device)
        dist.all_reduce(s)
        ms, wall_ms, samples = float(t[0]), float(t[1]), float(s[0])
        del model, optimizer
        torch.cuda.empty_cache()
        return {"value": samples / (ms / 1e3), "ms_per_step": ms / steps, "samples": samples,
                "e2e_value": samples / (wall_ms / 1e3), "clocks": clk.summary() if clk else None, "dir": vname}

    try:
        head = arm(a.topology, a.steps, a.warmup, True)
    except Exception as e:
        if rank == 0:
            emit({"impl": "reference", "unavailable": f"{type(e).__name__}: {e}"[:300]})
        dist.destroy_process_group()
        return
    restated = a.topology != "allreduce"
    path = (f"baseline/_ref/{head['dir']}: train_local_epoch + " +
            ("communication.average_gradients_equal (unmodified, NCCL backend)" if not restated else
             "neighbour exchange RESTATED on dist.batch_isend_irecv/NCCL (baseline/ring_nccl.py; the shipped mpi4py "
             "exchange cannot run here and is a no-op on CUDA tensors)"))
    extra = {"dtype_detail": "fp32 tensors, TF32 convolutions (torch default cudnn.allow_tf32), fp32 matmul",
             "backend": "nccl", "path": path, "restated": restated, "install": note}
    if not a.no_extra and a.topology == "allreduce" and world > 1:
        runs = {}
        for topo in ("ring", "double_ring"):
            try:
                m = arm(topo, max(5, a.extra_steps), 3, False)
                runs[f"{topo}/global_epoch"] = {"value": m["value"], "ms_per_step": m["ms_per_step"], "restated": True,
                                                "steps": max(5, a.extra_steps), "dir": m["dir"]}
            except Exception as e:
                runs[f"{topo}/global_epoch"] = {"error": f"{type(e).__name__}: {e}"[:300]}
        extra["runs"] = runs
    if rank == 0:
        emit({"metric": METRIC, "value": head["value"], "unit": "samples/s", "n_gpus": world,
              "steps": a.steps, "warmup": max(a.warmup, 3), "ms_per_step": head["ms_per_step"], "higher_is_better": True,
              "scaling": "weak", "vs_baseline": None, "dtype": DTYPE_HEADLINE,
              "data": "synthetic", "impl": "reference",
              "config": config_block(a, world, round(head["samples"] / a.steps), a.topology, "global_epoch"),
              "clocks": head["clocks"],
              "e2e": {"value": head["e2e_value"], "unit": "samples/s",
                      "h2d_bytes_per_step": a.batch * 3 * 32 * 32 * 4 + a.batch * 8, "d2h_bytes_per_step": 12},
              "gpu_launches": 0, "extra": extra})
    dist.destroy_process_group()


def _self_launch(a):
    """`python bench.py --gpus N` without a launcher: re-exec under torch.distributed.run."""
    import socket
    import subprocess
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={a.gpus}", "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.abspath(__file__)] + sys.argv[1:]
    raise SystemExit(subprocess.call(cmd))


if __name__ == "__main__":
    args = parse()
    if args.gpus > 1 and "WORLD_SIZE" not in os.environ:
        _self_launch(args)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)

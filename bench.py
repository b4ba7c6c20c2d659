#!/usr/bin/env python
"""Headline benchmark: train samples/s of the repo model (EnhancedCNN, batch 64 per GPU,
Adam) under a variant's schedule, on N GPUs of one node.

    python bench.py --gpus N --steps K --warmup W [--topology allreduce|ring|double_ring]
                    [--sync global_epoch|step] [--impl ours|reference]

Timed region = exactly K training steps (forward + backward + optimizer update on a fresh
batch each step) followed by the variant's exchange — the reference's schedule, in which a
"global epoch" is a run of local steps and ONE exchange (Balanced All-Reduce/trainer.py:46-47,
141-150).  ``--sync step`` measures the product schedule instead (fused exchange+update every
step).  `value` is device-timed (CUDA events, barrier + synchronize on both sides, max over
ranks) on device-resident synthetic data; `e2e` times the same K steps through the public
engine API with a pinned-host → device copy of every step's batch and a device → host read of
every step's loss.  Weak scaling: per-GPU batch is fixed at 64.

`--impl reference` runs the UNMODIFIED reference (baseline/_ref, see baseline/install_ref.py)
through its own `train_local_epoch` + `average_gradients_equal` on the NCCL backend.
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
    return p.parse_args()


def env_world():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


def emit(d):
    print(json.dumps(d), flush=True)


# =========================================================================== ours
def run_ours(a):
    import torch
    import torch.distributed as dist
    from ldnn_b200.config import Config, resolve
    from ldnn_b200.engine import make_engine
    from ldnn_b200.launch import init_runtime, shutdown
    from ldnn_b200.ops import _lib as L
    from ldnn_b200.parallel import make_exchanger
    from ldnn_b200.utils.timing import ClockSampler, barrier_sync, max_over_ranks

    rank, world, local = env_world()
    if world == 1:
        os.environ.setdefault("MASTER_PORT", str(29500 + os.getpid() % 2000))
    cfg = resolve(Config(topology=a.topology, sync=a.sync, model=a.model, batch_size=a.batch, backend="nccl",
                         engine=a.engine, dtype=(a.dtype if a.engine == "native" else "fp32"),
                         exchange_impl=a.exchange_impl, synthetic=True, quiet=True, seed=0,
                         aggregation_by="gradients", semantics="allreduce", apply_exchanged_grads=None,
                         comm_blocks=a.comm_blocks, aggregation_type=a.aggregation_type))
    rt = init_runtime("nccl", seed=0)
    dev = rt.device
    # Disbalanced shards: each rank's share of the work follows its ratio (the reference's uneven
    # random_split, Disbalanced */dataloader.py) and the exchange is weighted by n_r / sum(n).
    ratios = [1.0] * world
    if a.shard_ratios:
        if len(a.shard_ratios) != world:
            raise SystemExit(f"--shard_ratios needs {world} values")
        ratios = [float(v) for v in a.shard_ratios]
    share = ratios[rank] / max(ratios)
    my_steps = a.steps if a.sync == "step" else max(1, round(a.steps * share))
    if a.sync == "step" and a.shard_ratios:
        a.batch = max(1, round(a.batch * share))                # per-rank batch; every rank steps K times
    cfg.batch_size = a.batch
    engine = make_engine(cfg, rt)
    ex = make_exchanger(cfg, rt, engine)
    if a.aggregation_type == "shard":
        ex.set_shard_weights([v / sum(ratios) for v in ratios])
    ex.broadcast_state(engine)
    hw = engine.spec.in_hw
    ncls = engine.spec.head().num_classes
    # pool of distinct device-resident batches; the step's working set (params+moments+grads
    # ≈ 0.9 GB fp32 for the repo model) is far larger than the 126 MB L2 on its own.
    pool = 16
    g = torch.Generator(device="cpu").manual_seed(1234 + rank)
    xs = [torch.randn(a.batch, 3, hw, hw, generator=g).to(dev) for _ in range(pool)]
    ys = [torch.randint(0, ncls, (a.batch,), generator=g).to(dev) for _ in range(pool)]
    xs_pin = [x.cpu().pin_memory() for x in xs]
    ys_pin = [y.cpu().pin_memory() for y in ys]
    sync_step = a.sync == "step"
    overlap = sync_step and not a.no_overlap and hasattr(engine, "train_step_sync") and hasattr(ex, "exchange_bucket")

    def step(i, x, y):
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
            step(i, xs[i % pool], ys[i % pool])
        if not sync_step:
            ex.exchange_grads(engine, apply_update=False)

    epoch(max(a.warmup, 3))
    barrier_sync(dev)
    def launches_now():
        return (L.launch_count() if L.available() else 0) + getattr(engine, "graph_kernel_launches", 0)
    l0 = launches_now()
    # clocks / throttle reasons are sampled across BOTH timed regions (device-timed + end-to-end)
    clk = ClockSampler(dev.index or 0, period_s=0.05)
    clk.__enter__()
    barrier_sync(dev)
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    epoch(my_steps)
    t1.record()
    barrier_sync(dev)
    ms = t0.elapsed_time(t1)
    launches = launches_now() - l0        # eager launches + kernels replayed from the step's CUDA graphs
    ms = max_over_ranks(ms, dev)

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
            loss = step(i, x, y)
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
    clk.__exit__(None, None, None)

    if hasattr(ex, "check"):
        ex.check()
    samples = a.batch * my_steps                               # this rank's; summed over ranks below
    if world > 1:
        t = torch.tensor([float(samples)], device=dev, dtype=torch.float64)
        dist.all_reduce(t)
        samples = float(t[0])
    out = {
        "metric": METRIC if a.model == "enhanced_cnn" else METRIC.replace("EnhancedCNN, batch 64/GPU", f"ResNet-50 @{hw}, batch {a.batch}/GPU"),
        "value": samples / (ms / 1e3), "unit": "samples/s", "n_gpus": world, "steps": a.steps,
        "warmup": max(a.warmup, 3), "ms_per_step": ms / a.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16" if a.engine == "native" else "fp32", "data": "synthetic", "impl": "ours",
        "config": {"model": a.model, "global_batch": int(round(samples / a.steps)), "per_gpu_batch": a.batch,
                   "input": [3, hw, hw], "optimizer": "adam", "parallelism": f"dp{world}", "topology": a.topology,
                   "sync": a.sync, "overlap": bool(overlap), "exchange": a.exchange_impl, "engine": a.engine,
                   "aggregation_type": a.aggregation_type, "shard_ratios": a.shard_ratios,
                   "l2": "per-step working set (params+moments+grads ~0.9 GB) and a 16-batch input pool exceed the 126 MB L2"},
        "clocks": clk.summary(),
        "e2e": {"value": samples / e2e_s, "unit": "samples/s",
                "h2d_bytes_per_step": a.batch * 3 * hw * hw * 4 + a.batch * 8, "d2h_bytes_per_step": 4},
        "gpu_launches": int(launches),
    }
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
    if a.topology != "allreduce":
        return emit({"impl": "reference", "unavailable": "reference ring/double-ring variants hard-wire mpi4py "
                     "(not installed; torch has no MPI backend) and are no-ops on CUDA tensors"})
    vdir = os.path.join(ref_dir, "Balanced All-Reduce")
    if not os.path.isdir(vdir):
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
    sys.path.insert(0, vdir)
    import io
    import contextlib
    try:
        from model import EnhancedCNNModel                      # reference modules, unmodified
        from communication import average_gradients_equal
        import tqdm as _tqdm  # noqa: F401  (trainer imports it)
        import importlib.util
        # trainer.py imports dataloader.py, which imports torchvision (present) — fine.
        from trainer import train_local_epoch
    except Exception as e:
        return emit({"impl": "reference", "unavailable": f"import failed: {type(e).__name__}: {e}"})

    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29511")
    dist.init_process_group("nccl", rank=rank, world_size=world)
    device = torch.device("cuda:0")
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

    def epoch(ld):
        train_local_epoch(model, ld, criterion, optimizer, device, scheduler)
        average_gradients_equal(model, world)

    from ldnn_b200.utils.timing import ClockSampler
    warm, timed = loader(max(a.warmup, 3)), loader(a.steps)
    epoch(warm)
    dist.barrier(); torch.cuda.synchronize()
    with ClockSampler(local) as clk:
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        w0 = time.perf_counter()
        t0.record()
        epoch(timed)
        t1.record()
        dist.barrier(); torch.cuda.synchronize()
        wall = time.perf_counter() - w0
        ms = t0.elapsed_time(t1)
    t = torch.tensor([ms, wall * 1e3], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, wall_ms = float(t[0]), float(t[1])
    samples = a.batch * a.steps * world
    if rank == 0:
        emit({"metric": METRIC, "value": samples / (ms / 1e3), "unit": "samples/s", "n_gpus": world,
              "steps": a.steps, "warmup": max(a.warmup, 3), "ms_per_step": ms / a.steps, "higher_is_better": True,
              "scaling": "weak", "vs_baseline": None, "dtype": "fp32 (TF32 convolutions, torch default)",
              "data": "synthetic", "impl": "reference",
              "config": {"model": "enhanced_cnn", "global_batch": a.batch * world, "per_gpu_batch": a.batch,
                         "optimizer": "adam", "parallelism": f"dp{world}", "topology": "allreduce",
                         "sync": "global_epoch", "backend": "nccl",
                         "path": "baseline/_ref/Balanced All-Reduce: train_local_epoch + average_gradients_equal"},
              "clocks": clk.summary(),
              "e2e": {"value": samples / (wall_ms / 1e3), "unit": "samples/s",
                      "h2d_bytes_per_step": a.batch * 3 * 32 * 32 * 4 + a.batch * 8, "d2h_bytes_per_step": 12},
              "gpu_launches": 0, "install": note})
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

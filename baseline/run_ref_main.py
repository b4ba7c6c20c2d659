"""Runs the UNMODIFIED reference entrypoint (`baseline/_ref/Balanced All-Reduce/main.py::main`) on the NCCL backend
with the two substitutions this box forces, and times its training loop per global epoch.

Substitutions (SURVEY §6.2 allows exactly the first; the second is outside the timed region):
  1. `dataloader.get_loaders`: CIFAR-10 needs a download (no network), so the two `torchvision.datasets.CIFAR10(...)`
     objects become `TensorDataset`s of dataloader.py's shape (float32 [N,3,32,32], int64 labels).  Everything else of
     that function is the reference's own code, called here in the same order: `random_split` 80/20, the timing probe
     `estimate_epoch_duration`, `get_dataset_partition`, `DataLoader(shuffle=False)`.
  2. `vizualizator`: matplotlib is not installed here, so the six plot functions (rank 0, after training) are no-ops.
The reference pins `cuda:0` in every process (main.py:25) and overwrites MASTER_ADDR/PORT at import (main.py:14-15):
each rank gets its own GPU through CUDA_VISIBLE_DEVICES and the launcher's rendezvous is restored after the import.

Timing: hooks on the trainer module's own names — `train_local_epoch` (samples of the loader it was given) and
`get_subset_loaders` (called once at the end of every global epoch, trainer.py:188) — with a device synchronise at
each global-epoch boundary.  Output: one JSON line from rank 0.

    torchrun --nproc-per-node N baseline/run_ref_main.py --train 12800 --epochs_global 2 --epochs_local 2
"""
import argparse
import contextlib
import io
import json
import os
import sys
import time
import types

HERE = os.path.dirname(os.path.abspath(__file__))
PLOTS = ("plot_loss_distribution_by_worker", "plot_loss_distribution_per_epoch", "plot_loss_distribution_per_epoch_global",
         "plot_accuracy_distribution_per_epoch_global", "plot_metrics_global", "plot_metrics_total")


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--variant", default="Balanced All-Reduce")
    ap.add_argument("--train", type=int, default=12800, help="synthetic full training set size PER RANK (before the 80/20 split)")
    ap.add_argument("--test", type=int, default=640)
    ap.add_argument("--epochs_global", type=int, default=2)
    ap.add_argument("--epochs_local", type=int, default=2)
    ap.add_argument("--batch_size", type=int, default=64)
    return ap.parse_args()


def main():
    a = parse()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    os.environ["CUDA_VISIBLE_DEVICES"] = str(local)
    keep = {k: os.environ.get(k) for k in ("MASTER_ADDR", "MASTER_PORT")}
    vdir = os.path.join(HERE, "_ref", a.variant)
    if not os.path.isdir(vdir):
        print(json.dumps({"impl": "reference", "unavailable": f"{vdir} not found (python baseline/install_ref.py)"}))
        return
    import torch
    from torch.utils.data import DataLoader, TensorDataset, random_split
    sys.path.insert(0, vdir)
    viz = types.ModuleType("vizualizator")
    for name in PLOTS:
        setattr(viz, name, lambda *args, **kw: None)
    sys.modules["vizualizator"] = viz
    import dataloader as ref_dl                     # reference module, unmodified
    total = a.train * world                         # the reference shards ONE dataset over the ranks

    def get_loaders(batch_size, world_size, rank_, model, device):
        g = torch.Generator().manual_seed(1234)
        full_trainset = TensorDataset(torch.randn(total, 3, 32, 32, generator=g), torch.randint(0, 10, (total,), generator=g))
        testset = TensorDataset(torch.randn(a.test, 3, 32, 32, generator=g), torch.randint(0, 10, (a.test,), generator=g))
        # from here on: dataloader.py:32-51 with the reference's own helpers
        train_size = int(0.8 * len(full_trainset))
        val_size = len(full_trainset) - train_size
        trainset, valset = random_split(full_trainset, [train_size, val_size])
        trainloader_temp = DataLoader(trainset, batch_size=batch_size, shuffle=True)
        _, efficiency_ratios = ref_dl.estimate_epoch_duration(trainloader_temp, world_size, model, device, num_batches=10)
        part_train, train_indices = ref_dl.get_dataset_partition(trainset, rank_, efficiency_ratios)
        trainloader = DataLoader(part_train, batch_size=batch_size, shuffle=False)
        part_val, val_indices = ref_dl.get_dataset_partition(valset, rank_, efficiency_ratios)
        val_loader = DataLoader(part_val, batch_size=batch_size, shuffle=False)
        test_loader = DataLoader(testset, batch_size=batch_size, shuffle=False)
        return trainloader, val_loader, test_loader, trainset, valset, train_indices, val_indices

    ref_dl.get_loaders = get_loaders
    run_timed(a, rank, world, local, keep, torch)


def run_timed(a, rank, world, local, keep, torch):
    import trainer as ref_tr                        # reference module, unmodified
    marks, samples = [], [0]
    real_tle, real_gsl = ref_tr.train_local_epoch, ref_tr.get_subset_loaders

    def tle(model, trainloader, *rest, **kw):
        if not marks:                               # first step of training: start of global epoch 1
            torch.cuda.synchronize()
            marks.append((time.perf_counter(), 0))
        samples[0] += len(trainloader.dataset)
        return real_tle(model, trainloader, *rest, **kw)

    def gsl(*args, **kw):
        torch.cuda.synchronize()
        marks.append((time.perf_counter(), samples[0]))
        return real_gsl(*args, **kw)

    ref_tr.train_local_epoch, ref_tr.get_subset_loaders = tle, gsl
    import main as ref_main                         # reference entrypoint, unmodified
    for k, v in keep.items():                       # main.py:14-15 overwrote the launcher's rendezvous
        if v is not None:
            os.environ[k] = v
    ref_main.args = argparse.Namespace(local_rank=local, backend="nccl", epochs_local=a.epochs_local,
                                       epochs_global=a.epochs_global, batch_size=a.batch_size, lr=1e-3, time_limit=60,
                                       prev_fraction=0.5, next_fraction=0.5, aggregation_type="equal",
                                       aggregation_by="gradients", local_weight=0.5)
    import torch.distributed as dist
    real_destroy = dist.destroy_process_group
    dist.destroy_process_group = lambda *x, **k: None       # keep the group for the result reduction below
    with contextlib.redirect_stdout(io.StringIO()):          # the reference prints every rank's indices and metrics
        ref_main.main()
    per_epoch = []
    for (t0, s0), (t1, s1) in zip(marks[:-1], marks[1:]):
        t = torch.tensor([t1 - t0, float(s1 - s0)], dtype=torch.float64, device="cuda")
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone(); dist.all_reduce(tsum)
        per_epoch.append(float(tsum[1]) / float(tmax[0]))
    if rank == 0:
        print(json.dumps({"impl": "reference", "entrypoint": f"baseline/_ref/{a.variant}/main.py::main (NCCL backend)",
                          "n_gpus": world, "epochs_global": a.epochs_global, "epochs_local": a.epochs_local,
                          "batch_size": a.batch_size, "train_samples_per_s_per_global_epoch": per_epoch,
                          "substitutions": ["get_loaders: synthetic TensorDataset in place of the CIFAR-10 download",
                                            "vizualizator: no-op plots (matplotlib absent)"]}), flush=True)
    real_destroy()


if __name__ == "__main__":
    main()

"""Six directory entrypoints on synthetic data over gloo (BASELINE config 1 plumbing),
checkpoint round-trip, graphs."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from conftest import ROOT, free_port

VARIANTS = ["Balanced All-Reduce", "Balanced Ring", "Balanced Double-Ring",
            "Disbalanced All-Reduce", "Disbalanced Ring", "Disalanced Double-Ring"]
SMALL = ["--synthetic", "--synthetic_train", "192", "--synthetic_val", "64", "--synthetic_test", "64",
         "--epochs_global", "2", "--epochs_local", "1", "--batch_size", "32", "--probe_batches", "1",
         "--model", "enhanced_cnn", "--seed", "3", "--quiet"]


def launch(variant, nproc, extra, cwd_out):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(free_port()),
           os.path.join(ROOT, variant, "main.py")] + SMALL + ["--output_folder", str(cwd_out)] + extra
    env = dict(os.environ, OMP_NUM_THREADS="2", LDNN_TINY_WIDTHS="1")
    r = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    return r


@pytest.mark.parametrize("variant", VARIANTS)
def test_variant_runs_two_ranks(variant, tmp_path):
    out = tmp_path / "Graphs"
    extra = ["--aggregation_by", "weights"] if "Ring" in variant else []
    launch(variant, 2, extra, out)
    for f in ["loss_distribution_by_worker.png", "loss_distribution_per_epoch.png",
              "loss_distribution_per_epoch_global.png", "accuracy_distribution_per_epoch_global.png",
              "training_metrics.png", "training_metrics_0.png", "time_throughput.png", "metrics.json"]:
        assert (out / f).exists() and (out / f).stat().st_size > 0, f


def test_step_sync_shard_weighted_and_resume(tmp_path):
    out, ck = tmp_path / "G", tmp_path / "ck.pt"
    launch("Disbalanced Ring", 2, ["--sync", "step", "--aggregation_type", "shard", "--shard_ratios", "3", "1",
                                    "--save", str(ck)], out)
    blob = torch.load(ck, weights_only=False)
    assert len(blob["model"]) == 128 and blob["epoch"] == 2
    assert blob["model"]["layer4.1.conv2.weight"].dtype == torch.float32
    assert len(blob["indices"]["train"]) > 0
    per_rank = blob["indices_per_rank"]                       # every rank's own shard travels with the checkpoint
    assert len(per_rank) == 2 and len(per_rank[0]["train"]) > 2 * len(per_rank[1]["train"])      # ratios 3:1
    # resume: one more global epoch
    launch("Disbalanced Ring", 2, ["--sync", "step", "--aggregation_type", "shard", "--shard_ratios", "3", "1",
                                    "--resume", str(ck), "--epochs_global", "3"], out)


def test_time_limit_cuts_the_straggler_short(tmp_path):
    """--time_limit as intended upstream (trainer.py:112-119,134-139; unreachable there): with shards 6:1 the small
    rank finishes every local epoch first and announces it; the big rank stops `time_limit` seconds later instead of
    finishing its shard, and nobody deadlocks (the epoch's collective sits after the loop)."""
    import json
    out = tmp_path / "G"
    args = ["--shard_ratios", "6", "1", "--aggregation_type", "shard", "--synthetic_train", "448", "--batch_size", "8",
            "--epochs_local", "2"]
    launch("Disbalanced All-Reduce", 2, args + ["--time_limit", "0"], out)
    m = json.loads((out / "metrics.json").read_text())
    assert len(m["straggler_cuts"]) == 4 and all(c == [0] for c in m["straggler_cuts"]), m["straggler_cuts"]
    full = 448 * 0.8 * 6 / 7 / 8                        # steps of rank 0's whole shard
    done = [len(v) for v in m["all_epochs_losses"]]     # pooled per local epoch: rank 0's steps + rank 1's
    assert all(d < 0.8 * (full + full / 6) for d in done), (done, full)
    # with a generous limit nobody is cut and every step of both shards runs
    out2 = tmp_path / "G2"
    launch("Disbalanced All-Reduce", 2, args + ["--time_limit", "600"], out2)
    m2 = json.loads((out2 / "metrics.json").read_text())
    assert all(c == [] for c in m2["straggler_cuts"])
    assert sum(len(v) for v in m2["all_epochs_losses"]) > sum(done)

#!/usr/bin/env python
"""Throughput of the REAL entrypoints — `"<variant>/main.py"` of this repo vs the reference's own `main.py::main` — on
the same synthetic data, same schedule (2 global x 2 local epochs by default), each side timed by its own global-epoch
boundaries (ours: `global_epoch_samples_per_s` of metrics.json, i.e. trainer.train_global's timers; reference:
baseline/run_ref_main.py's hooks on its trainer).  Unlike bench.py (engine API, training steps only) a global epoch
here contains everything the programs do: data feeding, training, per-local-epoch validation, metric collectives, the
exchange and the re-shard.

    python bench/entrypoint_bench.py --gpus 2 [--variants "Balanced All-Reduce" "Balanced Ring"] [--train 12800]

One JSON line per run, then a summary table.  The reference arm exists for the All-Reduce variant only (the ring
variants' main.py needs mpi4py).
"""
import argparse
import json
import os
import socket
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ALL = ["Balanced All-Reduce", "Balanced Ring", "Balanced Double-Ring", "Disbalanced All-Reduce", "Disbalanced Ring",
       "Disalanced Double-Ring"]


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def torchrun(n, script, args, timeout=1800):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr",
           "127.0.0.1", "--master-port", str(free_port()), script] + args
    return subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, cwd=ROOT)


def run_ours(variant, a):
    out = tempfile.mkdtemp(prefix="ldnn_entry_")
    total = a.train * a.gpus
    args = ["--synthetic", "--synthetic_train", str(int(0.8 * total)), "--synthetic_val", str(total - int(0.8 * total)),
            "--synthetic_test", "640", "--epochs_global", str(a.epochs_global), "--epochs_local", str(a.epochs_local),
            "--batch_size", "64", "--backend", "nccl", "--quiet", "--seed", "0", "--output_folder", out] + a.extra
    r = torchrun(a.gpus, os.path.join(ROOT, variant, "main.py"), args)
    if r.returncode != 0:
        return {"impl": "ours", "variant": variant, "error": (r.stderr or r.stdout)[-600:]}
    m = json.load(open(os.path.join(out, "metrics.json")))
    return {"impl": "ours", "entrypoint": f"{variant}/main.py", "variant": variant, "n_gpus": a.gpus,
            "epochs_global": a.epochs_global, "epochs_local": a.epochs_local, "batch_size": 64,
            "train_samples_per_s_per_global_epoch": m["global_epoch_samples_per_s"],
            "global_epoch_seconds": m["global_epoch_seconds"], "exchange_ms": m.get("exchange_ms")}


def run_reference(a):
    r = torchrun(a.gpus, os.path.join(ROOT, "baseline", "run_ref_main.py"),
                 ["--train", str(a.train), "--epochs_global", str(a.epochs_global), "--epochs_local", str(a.epochs_local)])
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    if r.returncode != 0 or not lines:
        return {"impl": "reference", "error": (r.stderr or r.stdout)[-600:]}
    return json.loads(lines[-1])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--variants", nargs="+", default=ALL)
    ap.add_argument("--train", type=int, default=12800, help="synthetic samples per rank before the 80/20 split")
    ap.add_argument("--epochs_global", type=int, default=2)
    ap.add_argument("--epochs_local", type=int, default=2)
    ap.add_argument("--no_reference", action="store_true")
    ap.add_argument("extra", nargs="*", help="extra flags for our main.py (after --)")
    a = ap.parse_args()
    rows = []
    if not a.no_reference:
        rows.append(run_reference(a))
        print(json.dumps(rows[-1]), flush=True)
    for v in a.variants:
        rows.append(run_ours(v, a))
        print(json.dumps(rows[-1]), flush=True)
    print(f"# entrypoints, {a.gpus} GPU(s), {a.epochs_global} global x {a.epochs_local} local epochs, "
          f"{a.train} samples/rank: train samples/s per global epoch")
    for r in rows:
        name = r.get("entrypoint", r.get("variant", r["impl"]))
        if "error" in r:
            print(f"{name:<60} ERROR {r['error'][-200:]!r}")
        else:
            print(f"{name:<60} " + "  ".join(f"{v:10.0f}" for v in r["train_samples_per_s_per_global_epoch"]))


if __name__ == "__main__":
    main()

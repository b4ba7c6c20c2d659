"""The six reference graphs (same file names) + a time/throughput graph (S-VIZ).

Reference: ``Balanced All-Reduce/vizualizator.py:9-133`` — four box plots and two
two-panel line charts written under ``Graphs/``.  matplotlib (Agg, file output only)
is used when present; otherwise the built-in rasteriser in ``utils/raster.py``
produces the PNGs, and the raw series always go to ``metrics.json`` beside them.
"""
from __future__ import annotations

import json
import os
from pathlib import Path
from typing import List, Sequence

from . import raster


def _mpl():
    try:
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
        return plt
    except Exception:
        return None


def ensure_directory_exists(directory) -> None:
    os.makedirs(directory, exist_ok=True)


def _boxplot(groups: Sequence[Sequence[float]], labels: List[str], title, xlabel, ylabel, path):
    plt = _mpl()
    if plt is not None:
        fig = plt.figure()
        fig.set_size_inches(16, 10)
        plt.boxplot([list(g) for g in groups], labels=labels)
        plt.title(title); plt.xlabel(xlabel); plt.ylabel(ylabel); plt.xticks(rotation=45); plt.grid(True)
        plt.savefig(path)
        plt.close(fig)
        return
    cv = raster.Canvas()
    ax = raster.Axes(cv, 140, 70, 1560, 900, title, xlabel, ylabel)
    ax.boxplot(groups, labels)
    cv.save(path)


def _two_panel(xs, top, bottom, path, top_title="Individual Loss", bottom_title="Individual Accuracy",
               xlabel="Epochs", top_ylabel="Loss", bottom_ylabel="Accuracy"):
    """top/bottom: lists of (label, ys)."""
    plt = _mpl()
    if plt is not None:
        fig = plt.figure()
        fig.set_size_inches(16, 10)
        for k, (series, title, yl) in enumerate([(top, top_title, top_ylabel), (bottom, bottom_title, bottom_ylabel)]):
            fig.add_subplot(2, 1, k + 1)
            for label, ys in series:
                plt.plot(xs[:len(ys)], ys, "o-", label=label)
            plt.title(title); plt.xlabel(xlabel); plt.ylabel(yl); plt.legend()
        plt.tight_layout()
        plt.savefig(path)
        plt.close(fig)
        return
    cv = raster.Canvas()
    for k, (series, title, yl) in enumerate([(top, top_title, top_ylabel), (bottom, bottom_title, bottom_ylabel)]):
        ax = raster.Axes(cv, 140, 70 + 480 * k, 1560, 430 + 480 * k, title, xlabel, yl)
        vals = [v for _, ys in series for v in ys]
        ax.set_range((min(xs) if xs else 0, max(xs) if xs else 1), (min(vals) if vals else 0, max(vals) if vals else 1))
        ax.xticks(xs, [str(x) for x in xs])
        for i, (label, ys) in enumerate(series):
            ax.plot(xs[:len(ys)], ys, raster.COLORS[i % len(raster.COLORS)], label, i)
    cv.save(path)


# ---- the reference's six entry points (same names, same output files) ------------
def plot_loss_distribution_by_worker(loss_data, output_folder="Graphs"):
    ensure_directory_exists(output_folder)
    _boxplot(loss_data, [f"Worker {i}" for i in range(len(loss_data))], "Loss Distribution per Worker",
             "Worker", "Loss", Path(output_folder) / "loss_distribution_by_worker.png")


def plot_loss_distribution_per_epoch(loss_data, output_folder="Graphs"):
    ensure_directory_exists(output_folder)
    _boxplot(loss_data, [f"Epoch {i + 1}" for i in range(len(loss_data))],
             "Loss Distribution Across All Workers Per Epoch", "Epoch", "Loss",
             Path(output_folder) / "loss_distribution_per_epoch.png")


def plot_loss_distribution_per_epoch_global(loss_data, output_folder="Graphs"):
    ensure_directory_exists(output_folder)
    _boxplot(loss_data, [f"Epoch {i + 1}" for i in range(len(loss_data))],
             "Loss Distribution Across All Workers Per Epoch", "Epoch", "Loss",
             Path(output_folder) / "loss_distribution_per_epoch_global.png")


def plot_accuracy_distribution_per_epoch_global(acc_data, output_folder="Graphs"):
    ensure_directory_exists(output_folder)
    _boxplot(acc_data, [f"Epoch {i + 1}" for i in range(len(acc_data))],
             "Accuracy Distribution Across All Workers Per Epoch", "Epoch", "Accuracy",
             Path(output_folder) / "accuracy_distribution_per_epoch_global.png")


def plot_metrics_global(epochs, train_loss, train_accuracy, val_loss, val_accuracy, output_folder="Graphs"):
    ensure_directory_exists(output_folder)
    xs = list(range(1, epochs + 1))
    _two_panel(xs, [("Train Loss", train_loss), ("Validation Loss", val_loss)],
               [("Worker Train Accuracy", train_accuracy), ("Worker Val Accuracy", val_accuracy)],
               Path(output_folder) / "training_metrics.png")


def plot_metrics_total(epochs, train_loss, train_accuracy, val_loss, val_accuracy, rank, output_folder="Graphs"):
    ensure_directory_exists(output_folder)
    xs = list(range(1, epochs + 1))
    _two_panel(xs, [(f"Worker {rank} Train Loss", train_loss), (f"Worker {rank} Validation Loss", val_loss)],
               [(f"Worker {rank} Train Accuracy", train_accuracy), (f"Worker {rank} Val Accuracy", val_accuracy)],
               Path(output_folder) / f"training_metrics_{rank}.png")


# ---- new: time / throughput ---------------------------------------------------------
def plot_time_throughput(epoch_seconds, samples_per_s, exchange_ms=None, output_folder="Graphs"):
    ensure_directory_exists(output_folder)
    xs = list(range(1, len(epoch_seconds) + 1))
    top = [("Global epoch seconds", epoch_seconds)]
    if exchange_ms:
        top.append(("Exchange seconds", [m / 1e3 for m in exchange_ms]))
    _two_panel(xs, top, [("Samples per second (all workers)", samples_per_s)],
               Path(output_folder) / "time_throughput.png", "Time per Global Epoch", "Training Throughput",
               "Global Epochs", "Seconds", "Samples/s")


def write_all(history, cfg, rank=0, output_folder=None):
    """Everything rank 0 plots at the end of a run (``main.py:65-77``) + JSON dump."""
    out = output_folder or cfg.output_folder
    ensure_directory_exists(out)
    h = history
    plot_metrics_global(len(h.global_train_losses), h.global_train_losses, h.global_train_accuracies,
                        h.global_val_losses, h.global_val_accuracies, out)
    plot_metrics_total(len(h.worker_specific_train_losses), h.worker_specific_train_losses,
                       h.worker_specific_train_accuracies, h.worker_specific_val_losses,
                       h.worker_specific_val_accuracies, rank, out)
    plot_loss_distribution_by_worker(h.all_workers_losses, out)
    plot_loss_distribution_per_epoch(h.all_epochs_losses, out)
    plot_loss_distribution_per_epoch_global(h.global_epoch_losses, out)
    plot_accuracy_distribution_per_epoch_global(h.global_epoch_accuracies, out)
    plot_time_throughput(h.global_epoch_seconds, h.global_epoch_samples_per_s, h.exchange_ms, out)
    with open(Path(out) / "metrics.json", "w") as f:
        json.dump({k: v for k, v in vars(h).items()}, f)

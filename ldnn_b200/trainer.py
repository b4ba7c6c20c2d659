"""Global/local epoch loop (S-LOOP), validation and test evaluation (S-EVAL).

Structure follows the reference's ``train_global`` / ``train_local_epoch``
(``Balanced All-Reduce/trainer.py:11-223``) — local epochs, one exchange per global
epoch, re-shard, the 12 metric series — with these deliberate differences:

* no host sync inside a step: per-batch loss and correct-count stay on the device
  and are read back once per local epoch (upstream: three ``.item()`` per batch,
  ``trainer.py:212-216``);
* the 5+ tiny collectives per local epoch (C3–C6; the loss gather is duplicated
  upstream, ``trainer.py:55-76`` and ``:78-103``) are one fixed-size all-gather of a
  per-rank record;
* ``--sync step`` adds the product schedule: every step's gradients go through the
  fused exchange+update kernel instead of a local optimizer step;
* the straggler limit (``--time_limit``) works: a rank that finishes a local epoch announces it through a
  peer-visible flag word, the ranks still stepping stop ``time_limit`` s later (``parallel/straggler.py``);
  upstream's version is unreachable and would deadlock if it fired.
"""
from __future__ import annotations

import time
from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np
import torch
import torch.distributed as dist

from . import data as D
from .utils.timing import nvtx_range


@dataclass
class TrainHistory:
    """The reference's 12 returned lists (``trainer.py:13-25,192``) + timing extras."""
    all_workers_losses: List[List[float]] = field(default_factory=list)
    all_epochs_losses: List[List[float]] = field(default_factory=list)
    global_epoch_losses: List[List[float]] = field(default_factory=list)
    global_epoch_accuracies: List[List[float]] = field(default_factory=list)
    global_train_losses: List[float] = field(default_factory=list)
    global_train_accuracies: List[float] = field(default_factory=list)
    global_val_losses: List[float] = field(default_factory=list)
    global_val_accuracies: List[float] = field(default_factory=list)
    worker_specific_train_losses: List[float] = field(default_factory=list)
    worker_specific_train_accuracies: List[float] = field(default_factory=list)
    worker_specific_val_losses: List[float] = field(default_factory=list)
    worker_specific_val_accuracies: List[float] = field(default_factory=list)
    # extras (north star: "loss/accuracy/time graphs")
    global_epoch_seconds: List[float] = field(default_factory=list)
    global_epoch_samples_per_s: List[float] = field(default_factory=list)
    exchange_ms: List[float] = field(default_factory=list)
    straggler_cuts: List[List[int]] = field(default_factory=list)   # per local epoch: ranks the time limit cut short

    def as_tuple(self):
        return (self.all_workers_losses, self.all_epochs_losses, self.global_epoch_losses,
                self.global_epoch_accuracies, self.global_train_losses, self.global_train_accuracies,
                self.global_val_losses, self.global_val_accuracies, self.worker_specific_train_losses,
                self.worker_specific_train_accuracies, self.worker_specific_val_losses,
                self.worker_specific_val_accuracies)


# ---------------------------------------------------------------------------
# control-plane gather (K-MET): one record per rank per local epoch
# ---------------------------------------------------------------------------
class RecordGather:
    HEADER = 8  # [n_losses, train_loss, train_acc, val_loss, val_acc, cut short by --time_limit, (unused), samples]

    def __init__(self, rt, exchanger=None):
        self.rt = rt
        # on GPU the record all-gather is the K-MET P2P kernel; torch.distributed otherwise
        self.p2p = exchanger if hasattr(exchanger, "allgather_floats") else None

    def max_int(self, v: int) -> int:
        if self.rt.world_size == 1:
            return int(v)
        t = torch.tensor([v], dtype=torch.int64, device=self._dev())
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return int(t.item())

    def sum_float(self, v: float) -> float:
        if self.rt.world_size == 1:
            return float(v)
        t = torch.tensor([v], dtype=torch.float64, device=self._dev())
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    def _dev(self):
        return self.rt.device if self.rt.backend == "nccl" else torch.device("cpu")

    def gather(self, header: List[float], losses: np.ndarray, capacity: int) -> np.ndarray:
        rec = np.full(self.HEADER + capacity, -1.0, dtype=np.float32)
        rec[:len(header)] = header
        rec[self.HEADER:self.HEADER + len(losses)] = losses
        if self.rt.world_size == 1:
            return rec[None, :]
        if self.p2p is not None and len(rec) <= self.p2p.METRIC_CAP:
            return self.p2p.allgather_floats(rec)
        t = torch.from_numpy(rec).to(self._dev())
        out = [torch.empty_like(t) for _ in range(self.rt.world_size)]
        dist.all_gather(out, t)
        return torch.stack(out).cpu().numpy()


# ---------------------------------------------------------------------------
# local epoch / validation / evaluation
# ---------------------------------------------------------------------------
def train_local_epoch(engine, loader: D.Loader, rt, exchanger=None, sync_step: bool = False,
                      num_steps: Optional[int] = None, by_weights: bool = False, straggler=None):
    """One pass over the rank's shard (``trainer.py:194-223``).
    Returns (mean loss, accuracy %, per-batch losses, samples).

    ``sync_step``: exchange every step — the mean gradient feeds the (fused) update, or with ``by_weights``
    (``--aggregation_by weights``) every rank updates locally and the parameters are averaged after the step."""
    n_steps = len(loader) if num_steps is None else num_steps
    if n_steps == 0:
        engine.scheduler_step()
        return 0.0, 0.0, np.zeros(0, np.float32), 0
    dev = rt.device
    losses = torch.zeros(n_steps, dtype=torch.float32, device=dev)
    correct = torch.zeros((), dtype=torch.float32, device=dev)
    total, i, stopped = 0, 0, False
    while i < n_steps:                       # wraps when a common step count exceeds the shard
        for x, y in D.DeviceFeeder(loader, dev):
            if i >= n_steps:
                break
            if sync_step and exchanger is not None and by_weights:
                loss, corr = engine.train_step(x, y, update=True)
                exchanger.exchange_weights(engine)
            elif sync_step and exchanger is not None:
                if hasattr(engine, "train_step_sync") and hasattr(exchanger, "exchange_bucket") \
                        and not exchanger.spec.is_gossip():
                    loss, corr = engine.train_step_sync(x, y, exchanger)      # bucketed, overlapped
                else:
                    loss, corr = engine.train_step(x, y, update=False)
                    exchanger.exchange_grads(engine, apply_update=True)
            else:
                loss, corr = engine.train_step(x, y, update=True)
            losses[i] = loss
            correct += corr
            total += int(y.shape[0])
            i += 1
            if straggler is not None and i < n_steps and straggler.should_stop():
                stopped = True               # a peer finished this epoch time_limit seconds ago: move on
                break
        if len(loader) == 0 or stopped:
            break
    if straggler is not None and not stopped:
        straggler.announce()
    engine.scheduler_step()
    losses = losses[:i]
    host = losses.cpu().numpy()              # the one sync of the epoch
    n_ok = float(correct.item())
    return float(host.mean()), 100.0 * n_ok / max(total, 1), host, total


def validate(engine, loader: D.Loader, rt):
    """Eval-mode loss/accuracy over the rank's validation shard (``validator.py:3-23``)."""
    n = len(loader)
    if n == 0:
        return 0.0, 0.0
    dev = rt.device
    loss_sum = torch.zeros((), dtype=torch.float32, device=dev)
    correct = torch.zeros((), dtype=torch.float32, device=dev)
    total = 0
    for x, y in D.DeviceFeeder(loader, dev):
        l, c, _ = engine.eval_step(x, y)
        loss_sum += l
        correct += c
        total += int(y.shape[0])
    return float(loss_sum.item()) / n, 100.0 * float(correct.item()) / max(total, 1)


def prf_scores(labels: np.ndarray, preds: np.ndarray, num_classes: int):
    """macro / weighted / micro precision, recall, F1 from the confusion matrix — the
    numbers upstream gets from sklearn (``evaluator.py:41-53``), zero_division=0."""
    cm = np.zeros((num_classes, num_classes), dtype=np.int64)
    np.add.at(cm, (labels, preds), 1)
    present = np.union1d(labels, preds)
    tp = np.diag(cm).astype(np.float64)
    pred_n, true_n = cm.sum(0).astype(np.float64), cm.sum(1).astype(np.float64)
    with np.errstate(divide="ignore", invalid="ignore"):
        prec = np.where(pred_n > 0, tp / pred_n, 0.0)
        rec = np.where(true_n > 0, tp / true_n, 0.0)
        f1 = np.where(prec + rec > 0, 2 * prec * rec / (prec + rec), 0.0)
    w = true_n[present] / max(true_n[present].sum(), 1.0)
    micro = tp.sum() / max(cm.sum(), 1)
    return {
        "macro": (float(prec[present].mean()), float(rec[present].mean()), float(f1[present].mean())),
        "weighted": (float((prec[present] * w).sum()), float((rec[present] * w).sum()), float((f1[present] * w).sum())),
        "micro": (float(micro), float(micro), float(micro)),
    }


def evaluate(engine, loader: D.Loader, rt, num_classes: int = 10, verbose: bool = True):
    """Rank-0 test evaluation (``evaluator.py:6-61``): returns
    (loss, accuracy %, predictions, labels) and prints the same three summary lines."""
    dev = rt.device
    n = len(loader)
    loss_sum = torch.zeros((), dtype=torch.float32, device=dev)
    preds, labels = [], []
    for x, y in D.DeviceFeeder(loader, dev):
        l, _, p = engine.eval_step(x, y)
        loss_sum += l
        preds.append(p.clone())              # the native engine hands out a view of its step-static buffer
        labels.append(y)
    preds = torch.cat(preds).cpu().numpy() if preds else np.zeros(0, np.int64)
    labels = torch.cat(labels).cpu().numpy() if labels else np.zeros(0, np.int64)
    loss = float(loss_sum.item()) / max(n, 1)
    acc = 100.0 * float((preds == labels).sum()) / max(len(labels), 1)
    s = prf_scores(labels, preds, num_classes)
    if verbose:
        pw, rw, fw = s["weighted"]
        print(f"Worker {rt.rank}, Test Loss: {loss:.4f}, Test Accuracy: {acc:.2f}%, Weighted Precision: {pw:.2f}, "
              f"Weighted Recall: {rw:.2f}, Weighted F1 Score: {fw:.2f}")
        print("Precision: {:.2f}, Recall: {:.2f}, F1 Score: {:.2f}".format(*s["macro"]))
        print("Micro precision: {:.2f}, Micro recall: {:.2f}, Micro F1 Score: {:.2f}".format(*s["micro"]))
    return loss, acc, preds, labels


def refresh_shard_weights(cfg, rt, shards: D.ShardState, exchanger):
    """Shard-weighted averaging uses w_r = n_r / sum(n) of the shards the ranks ACTUALLY hold: call after the
    initial sharding, after every re-shard (sizes follow ratio x total, which moves with the timing-derived ratios)
    and after a checkpoint restored different shards.  Collective; returns the weights (None if not shard mode)."""
    if cfg.aggregation_type != "shard":
        return None
    sizes = D.all_gather_floats(float(len(shards.train_indices)), rt)
    w = [s / max(sum(sizes), 1.0) for s in sizes]
    exchanger.set_shard_weights(w)
    return w


# ---------------------------------------------------------------------------
# the global loop
# ---------------------------------------------------------------------------
def train_global(engine, bundle: D.DataBundle, shards: D.ShardState, cfg, rt, exchanger,
                 history: Optional[TrainHistory] = None, start_epoch: int = 0, on_epoch_end=None
                 ) -> TrainHistory:
    ws, rank = rt.world_size, rt.rank
    h = history or TrainHistory()
    if not h.all_workers_losses:
        h.all_workers_losses = [[] for _ in range(ws)]
    g = RecordGather(rt, exchanger)
    from .parallel.straggler import make_straggler, ticket
    straggler = make_straggler(cfg, rt, exchanger)
    rng = np.random.default_rng(None if cfg.seed is None else cfg.seed + 1000 + rank)
    sync_step = cfg.sync == "step"

    for ge in range(start_epoch, cfg.epochs_global):
        train_loader = D.Loader(bundle.trainset, shards.train_indices, cfg.batch_size,
                                D.chunk_batches_for(cfg, bundle.trainset))
        val_loader = D.Loader(bundle.valset, shards.val_indices, cfg.batch_size,
                              D.chunk_batches_for(cfg, bundle.valset))
        capacity = g.max_int(len(train_loader))
        steps = capacity if sync_step else None      # common step count for the per-step schedule
        if sync_step and ws > 1 and -g.max_int(-len(train_loader)) == 0:
            # every step of this schedule is a collective: a rank with nothing to feed it would leave its peers
            # spinning in the exchange until the watchdog fires.  All ranks see the same minimum and stop together.
            raise RuntimeError(f"--sync step: at least one rank holds an empty training shard in global epoch {ge + 1} "
                               f"(this rank: {len(shards.train_indices)} samples, batch {cfg.batch_size}); use a larger "
                               "dataset, a smaller batch size or less skewed --shard_ratios")
        sums = np.zeros(4, dtype=np.float64)         # train loss/acc, val loss/acc over local epochs
        cur_losses: List[float] = []
        cur_accs: List[float] = []
        t_start = time.time()
        samples = 0

        for le in range(cfg.epochs_local):
            if straggler is not None:
                straggler.begin(ticket(ge, le))
            with nvtx_range(f"train g{ge} l{le}"):
                tl, ta, batch_losses, n = train_local_epoch(engine, train_loader, rt, exchanger, sync_step, steps,
                                                            by_weights=cfg.aggregation_by == "weights",
                                                            straggler=straggler)
            cut = 1.0 if (straggler is not None and straggler.cut_short) else 0.0
            samples += n
            with nvtx_range("validate"):
                vl, va = validate(engine, val_loader, rt)
            recs = g.gather([len(batch_losses), tl, ta, vl, va, cut, 0.0, n], batch_losses, capacity)
            if rank == 0:
                pooled = []
                for r in range(ws):
                    k = int(recs[r, 0])
                    v = recs[r, g.HEADER:g.HEADER + k].tolist()
                    h.all_workers_losses[r].extend(v)
                    pooled.extend(v)
                h.all_epochs_losses.append(pooled)
                cur_losses.extend(pooled)
                cur_accs.append(float(recs[:, 2].mean()))
                h.worker_specific_train_losses.append(tl)
                h.worker_specific_train_accuracies.append(ta)
                h.worker_specific_val_losses.append(vl)
                h.worker_specific_val_accuracies.append(va)
            if not cfg.quiet:
                print(f"Rank {rank}, Global Epoch {ge + 1}, Local Epoch {le + 1}, Loss: {tl}, Accuracy: {ta}")
                print(f"Worker {rank}, Global Epoch {ge + 1}, Validation Loss: {vl:.4f}, Validation Accuracy: {va:.2f}%")
            sums += np.array([tl, ta, vl, va])
            if rank == 0:
                h.straggler_cuts.append([int(r) for r in range(ws) if recs[r, 5] >= 1.0])

        g.max_int(1)                                   # post-loop finish flag (C6)

        # ---- the exchange (trainer.py:141-150) --------------------------------
        t_x = time.time()
        if not sync_step:
            with nvtx_range("exchange"):
                if cfg.aggregation_by == "gradients":
                    exchanger.exchange_grads(engine, apply_update=bool(cfg.apply_exchanged_grads))
                else:
                    exchanger.exchange_weights(engine)
        if rt.is_cuda:
            torch.cuda.synchronize(rt.device)
        h.exchange_ms.append(1e3 * (time.time() - t_x))

        # ---- global metrics (C11) ----------------------------------------------
        tot = np.array([g.sum_float(v) for v in sums]) if ws > 1 else sums
        avg = tot / ws / max(cfg.epochs_local, 1)
        h.global_train_losses.append(float(avg[0]))
        h.global_train_accuracies.append(float(avg[1]))
        h.global_val_losses.append(float(avg[2]))
        h.global_val_accuracies.append(float(avg[3]))
        h.global_epoch_losses.append(cur_losses)
        h.global_epoch_accuracies.append(cur_accs)

        if ws > 1:
            dist.barrier()                             # C12
        dur = time.time() - t_start
        tot_dur = g.sum_float(dur)                     # C13
        tot_samples = g.sum_float(float(samples))
        h.global_epoch_seconds.append(dur)
        h.global_epoch_samples_per_s.append(tot_samples / max(dur, 1e-9))
        shards = D.reshard(bundle, shards, cfg, rt, dur, tot_dur, rng)
        refresh_shard_weights(cfg, rt, shards, exchanger)      # the exchange weights follow the new shard sizes
        if on_epoch_end is not None:
            on_epoch_end(ge, h, shards)
    return h

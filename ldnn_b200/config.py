"""Command-line surface.

Every flag of every reference ``main.py`` keeps its name and default
(``Balanced All-Reduce/main.py:83-97``, ``Balanced Ring/main.py:77-92``,
``Balanced Double-Ring/main.py:77-93``, ``Disbalanced */main.py``).  The variant
(topology × balance) is no longer "which directory you cd into" only: the six
directory entrypoints pin it, and ``--topology/--balance`` expose it directly.

Flags the reference parses but never reads (``--local-rank``, ``--gpu_weight``,
``--dist-url``; ``--backend`` in the MPI variants — SURVEY §2.3 B10) are still
accepted so existing launch lines keep working.
"""
from __future__ import annotations

import os

import argparse
from dataclasses import dataclass, field, asdict
from typing import List, Optional

TOPOLOGIES = ("allreduce", "ring", "double_ring")
BALANCES = ("balanced", "disbalanced")


@dataclass
class Config:
    # variant
    topology: str = "allreduce"
    balance: str = "balanced"
    # reference flags (names + defaults preserved)
    local_rank: Optional[int] = None
    backend: str = "gloo"
    epochs_local: int = 5
    epochs_global: int = 20
    batch_size: int = 64
    lr: float = 1e-3
    time_limit: float = 60.0
    prev_fraction: float = 0.5
    next_fraction: float = 0.5
    aggregation_type: str = "equal"          # equal | weighted (self-weight) | shard (n_r/Σn)
    aggregation_by: str = "gradients"        # gradients | weights
    local_weight: float = 0.5
    fixed_ratio: float = 0.5
    gpu_weight: float = 10.0                 # parsed, unused (as upstream)
    dist_url: Optional[str] = None           # parsed, unused (as upstream)
    # new surface
    model: str = "enhanced_cnn"
    num_classes: int = 10
    sync: str = "global_epoch"               # global_epoch (reference schedule) | step (product path)
    semantics: str = "auto"                  # auto | gossip (reference ring math) | allreduce (true RS+AG rings)
    exchange_impl: str = "auto"              # auto | fused (P2P sm_100a kernels) | dist (torch.distributed)
    engine: str = "auto"                     # auto | native (sm_100a kernels + CUDA graph) | torch
    optimizer: str = "adam"                  # adam (reference) | sgd (momentum; north star)
    momentum: float = 0.9
    weight_decay: float = 0.0
    lr_step: int = 25
    lr_gamma: float = 0.1
    dtype: str = "tf32"                      # GPU compute: tf32 (fp32 storage + TF32 MMA = the reference's
                                             # numerics, default) | bf16 (native, faster) | fp32 (torch engine)
    apply_exchanged_grads: Optional[bool] = None  # None → product=True, compat(global_epoch)=False
    synthetic: bool = False
    synthetic_train: int = 4096
    synthetic_val: int = 1024
    synthetic_test: int = 1024
    data_root: str = "data"
    augment: str = "batched"                 # batched | per_image | none  (AutoAugment on the on-disk dataset)
    data_on_device: str = "auto"             # auto (uint8 on-disk dataset on a GPU) | on | off
    chunk_batches: int = 0                   # batches gathered+augmented per call; 0 = 64 on device, 16 on host
    trace: Optional[str] = None              # write a Chrome-trace phase timeline to <trace>.rank<r>.json
    shard_ratios: Optional[List[float]] = None  # explicit per-rank sizes, e.g. 4 2 1 1
    probe_batches: int = 10
    replace: Optional[bool] = None           # re-shard sampling with replacement (None → variant default)
    seed: Optional[int] = None
    save: Optional[str] = None
    resume: Optional[str] = None
    output_folder: str = "Graphs"
    comm_blocks: int = 0                     # SM budget of the exchange kernel (0 → default)
    quiet: bool = False
    watchdog_ms: int = 20000                 # device-side spin-wait timeout

    def to_dict(self):
        return asdict(self)


def build_parser(topology: Optional[str] = None, balance: Optional[str] = None) -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description="B200-native data-parallel trainer "
                                            "(all-reduce / ring / double-ring; balanced / disbalanced)")
    d = Config()
    p.add_argument("--topology", choices=TOPOLOGIES, default=topology or d.topology)
    p.add_argument("--balance", choices=BALANCES, default=balance or d.balance)
    # --- reference flags -------------------------------------------------
    p.add_argument("--local-rank", "--local_rank", type=int, dest="local_rank", default=None,
                   help="accepted for torch.distributed.launch; LOCAL_RANK env wins")
    p.add_argument("--backend", type=str, default=d.backend, choices=["nccl", "gloo"])
    p.add_argument("--epochs_local", type=int, default=d.epochs_local)
    p.add_argument("--epochs_global", type=int, default=d.epochs_global)
    p.add_argument("--batch_size", type=int, default=d.batch_size)
    p.add_argument("--lr", type=float, default=d.lr)
    p.add_argument("--time_limit", type=float, default=d.time_limit)
    p.add_argument("--prev_fraction", type=float, default=d.prev_fraction)
    p.add_argument("--next_fraction", type=float, default=d.next_fraction)
    p.add_argument("--aggregation_type", type=str, default=d.aggregation_type,
                   choices=["equal", "weighted", "shard"])
    p.add_argument("--aggregation_by", type=str, default=d.aggregation_by, choices=["gradients", "weights"])
    p.add_argument("--local_weight", type=float, default=d.local_weight)
    p.add_argument("--fixed_ratio", type=float, default=d.fixed_ratio)
    p.add_argument("--gpu_weight", type=float, default=d.gpu_weight)
    p.add_argument("--dist-url", "--dist_url", type=str, dest="dist_url", default=None)
    # --- new flags ---------------------------------------------------------
    p.add_argument("--model", choices=["enhanced_cnn", "resnet50"], default=d.model)
    p.add_argument("--num_classes", type=int, default=d.num_classes)
    p.add_argument("--sync", choices=["global_epoch", "step"], default=d.sync)
    p.add_argument("--semantics", choices=["auto", "gossip", "allreduce"], default=d.semantics)
    p.add_argument("--exchange_impl", choices=["auto", "fused", "dist"], default=d.exchange_impl)
    p.add_argument("--engine", choices=["auto", "native", "torch"], default=d.engine)
    p.add_argument("--optimizer", choices=["adam", "sgd"], default=d.optimizer)
    p.add_argument("--momentum", type=float, default=d.momentum)
    p.add_argument("--weight_decay", type=float, default=d.weight_decay)
    p.add_argument("--lr_step", type=int, default=d.lr_step)
    p.add_argument("--lr_gamma", type=float, default=d.lr_gamma)
    p.add_argument("--dtype", choices=["tf32", "bf16", "fp32"], default=d.dtype)
    p.add_argument("--apply_exchanged_grads", type=lambda s: s.lower() in ("1", "true", "yes"), default=None)
    p.add_argument("--synthetic", action="store_true")
    p.add_argument("--synthetic_train", type=int, default=d.synthetic_train)
    p.add_argument("--synthetic_val", type=int, default=d.synthetic_val)
    p.add_argument("--synthetic_test", type=int, default=d.synthetic_test)
    p.add_argument("--data_root", type=str, default=d.data_root)
    p.add_argument("--augment", type=str, default=d.augment, choices=["batched", "per_image", "none"],
                   help="AutoAugment on the on-disk dataset: whole batches at once (default), upstream's one image "
                        "at a time, or off")
    p.add_argument("--data_on_device", type=str, default=d.data_on_device, choices=["auto", "on", "off"],
                   help="keep the dataset resident in GPU memory (auto: the uint8 on-disk dataset only)")
    p.add_argument("--trace", type=str, default=os.environ.get("LDNN_TRACE") or d.trace,
                   help="write a Chrome-trace timeline of the phases (train / validate / exchange per epoch) to "
                        "<PATH>.rank<r>.json")
    p.add_argument("--chunk_batches", type=int, default=d.chunk_batches,
                   help="batches gathered / augmented per call (0 = 64 when resident on the device, else 16)")
    p.add_argument("--shard_ratios", type=float, nargs="+", default=None)
    p.add_argument("--probe_batches", type=int, default=d.probe_batches)
    p.add_argument("--replace", type=lambda s: s.lower() in ("1", "true", "yes"), default=None)
    p.add_argument("--seed", type=int, default=None)
    p.add_argument("--save", type=str, default=None)
    p.add_argument("--resume", type=str, default=None)
    p.add_argument("--output_folder", type=str, default=d.output_folder)
    p.add_argument("--comm_blocks", type=int, default=0)
    p.add_argument("--quiet", action="store_true")
    p.add_argument("--watchdog_ms", type=int, default=d.watchdog_ms)
    return p


def parse_args(argv=None, topology: Optional[str] = None, balance: Optional[str] = None) -> Config:
    ns = build_parser(topology, balance).parse_args(argv)
    cfg = Config(**vars(ns))
    return resolve(cfg)


def resolve(cfg: Config) -> Config:
    """Fill the ``auto`` fields from the variant + schedule."""
    if cfg.semantics == "auto":
        # per-global-epoch schedule keeps the reference's neighbour-gossip math for the
        # ring variants; the per-step product path runs true RS+AG rings.
        cfg.semantics = "gossip" if (cfg.sync == "global_epoch" and cfg.topology != "allreduce") else "allreduce"
    if cfg.apply_exchanged_grads is None:
        # upstream's exchanged gradients are dropped by the next zero_grad.
        cfg.apply_exchanged_grads = (cfg.sync == "step")
    if cfg.replace is None:
        # BalAR/BalR sample without replacement; BalDR and all Dis* with.
        cfg.replace = not (cfg.balance == "balanced" and cfg.topology in ("allreduce", "ring"))
    if cfg.model == "resnet50" and cfg.num_classes == 10:
        cfg.num_classes = 1000
    return cfg

"""Orchestration (L5): the body of every variant's ``main.py``.

Sequence mirrors ``Balanced All-Reduce/main.py:17-79``: init process group → build
model → Xavier init → rank-0 broadcast → loaders (timing probe + sharding) →
optimizer/scheduler → ``train_global`` → rank-0 test evaluation + graphs → teardown;
with save/resume and the fused-exchange bring-up added.
"""
from __future__ import annotations

import sys
from typing import Optional

import numpy as np

from . import data as D
from .config import Config, parse_args
from .engine import make_engine
from .launch import init_runtime, shutdown
from .parallel import bring_up
from .trainer import TrainHistory, evaluate, refresh_shard_weights, train_global
from .utils import checkpoint as ckpt
from .utils import viz


def run(cfg: Config) -> TrainHistory:
    rt = init_runtime(cfg.backend, cfg.seed)
    try:
        return _run(cfg, rt)
    finally:
        shutdown(rt)


def _run(cfg: Config, rt) -> TrainHistory:
    from .utils.timing import TRACE, nvtx_range
    if cfg.trace:
        TRACE.enable()
    try:
        return _run_traced(cfg, rt, nvtx_range)
    finally:
        if cfg.trace:
            TRACE.dump(cfg.trace, rt.rank)


def _run_traced(cfg: Config, rt, phase) -> TrainHistory:
    with phase("bring-up"):
        # engine + exchange together: the broadcast is the first thing that uses the exchange, and a failed NVLS
        # bring-up falls back (collectively) to CUDA-IPC slabs
        engine, exchanger = bring_up(cfg, rt, make_engine)
    bundle = D.get_data(cfg, rt.rank)
    D.place_bundle(bundle, cfg, rt)
    exchanger.broadcast_state(engine, src=0)

    shards = D.initial_shards(bundle, cfg, rt, engine)
    refresh_shard_weights(cfg, rt, shards, exchanger)      # shard-weighted averaging: w_r = n_r / Σn

    history = TrainHistory()
    start_epoch = 0
    if cfg.resume:
        start_epoch = ckpt.load_checkpoint(cfg.resume, engine, history, shards, rank=rt.rank, world_size=rt.world_size)
        engine.params_changed()
        refresh_shard_weights(cfg, rt, shards, exchanger)  # the checkpoint may have restored different shards

    def on_epoch_end(ge, h, sh):
        if cfg.save:
            exchanger.gather_moments(engine)       # collective: fused per-step updates shard the moments by owner
            everyone = ckpt.gather_shards(sh, rt)  # collective: each rank trains on its own shard
            if rt.rank == 0:
                ckpt.save_checkpoint(cfg.save, engine, h, sh, epoch=ge + 1, cfg=cfg, all_shards=everyone)
            else:                                  # replicas differ between exchanges (and BN buffers always do)
                ckpt.save_rank_state(cfg.save, engine, rt.rank, rt.world_size, epoch=ge + 1)

    history = train_global(engine, bundle, shards, cfg, rt, exchanger, history, start_epoch, on_epoch_end)

    if rt.rank == 0:
        test_loader = D.Loader(bundle.testset, None, cfg.batch_size)
        with phase("test evaluation"):
            evaluate(engine, test_loader, rt, bundle.num_classes, verbose=not cfg.quiet)
        with phase("graphs"):
            viz.write_all(history, cfg, rank=rt.rank)
    exchanger.close()
    return history


def main(argv=None, topology: Optional[str] = None, balance: Optional[str] = None) -> TrainHistory:
    cfg = parse_args(argv, topology, balance)
    return run(cfg)


if __name__ == "__main__":
    main(sys.argv[1:])

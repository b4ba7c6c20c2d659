"""Data-parallel exchange layer: semantics, torch.distributed baseline, fused P2P kernels."""
from .exchange import DistExchanger, ExchangeSpec, gossip_coefficients, mixing_matrix, reduce_form  # noqa: F401


def make_exchanger(cfg, rt, engine, shard_weights=None):
    """Pick the exchange realisation: the fused sm_100a P2P kernels on GPU (default),
    ``torch.distributed`` (gloo on CPU; NCCL = the measured baseline) otherwise."""
    spec = ExchangeSpec.from_config(cfg, rt.world_size, shard_weights)
    impl = cfg.exchange_impl
    one_node = rt.single_node() if hasattr(rt, "single_node") else True
    if impl == "auto":
        # the fused kernels load/store peer memory over one NVSwitch / CUDA-IPC domain; a job that spans hosts
        # runs the same semantics over torch.distributed (NCCL) instead
        impl = "fused" if (rt.is_cuda and one_node) else "dist"
    if impl == "fused":
        if not rt.is_cuda:
            raise RuntimeError("--exchange_impl fused needs CUDA devices with peer access")
        if not one_node:
            raise RuntimeError("--exchange_impl fused needs every rank on one host (one NVSwitch / CUDA-IPC domain); "
                               "use --exchange_impl dist (or auto) for multi-node jobs")
        from .fused import FusedExchanger
        return FusedExchanger(spec, rt, engine, comm_blocks=cfg.comm_blocks, watchdog_ms=cfg.watchdog_ms)
    ex = DistExchanger(spec, rt.rank)
    return ex


def bring_up(cfg, rt, make_engine):
    """(engine, exchanger) with one collective fallback: if the exchange bring-up fails on ANY rank while the slabs came
    from torch's symmetric-memory allocator (the in-switch NVLS path, on by default for multi-rank GPU jobs), every rank
    rebuilds its engine on plain CUDA-IPC slabs (`LDNN_NVLS=0`: two-shot / ring kernels) instead of aborting the job."""
    import os
    import warnings
    import torch
    import torch.distributed as dist
    engine = make_engine(cfg, rt)
    err = None
    try:
        ex = make_exchanger(cfg, rt, engine)
    except Exception as e:                       # noqa: BLE001 — reported below if there is no fallback
        ex, err = None, e
    failed = 0 if err is None else 1
    if rt.world_size > 1 and dist.is_initialized():
        t = torch.tensor([failed], dtype=torch.int32, device=rt.device if rt.backend == "nccl" else "cpu")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        failed = int(t.item())
    if not failed:
        return engine, ex
    from ..ops import native_ops as K
    if rt.is_cuda and os.environ.get("LDNN_NVLS", "") != "0" and K.nvls_requested():
        warnings.warn(f"exchange bring-up on multicast-capable slabs failed ({type(err).__name__ if err else 'on a peer'}: "
                      f"{err}); falling back to CUDA-IPC slabs (LDNN_NVLS=0)")
        os.environ["LDNN_NVLS"] = "0"
        if ex is not None:
            ex.close()
        del engine
        engine = make_engine(cfg, rt)
        return engine, make_exchanger(cfg, rt, engine)
    raise err if err is not None else RuntimeError("exchange bring-up failed on a peer rank")

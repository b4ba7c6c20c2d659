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

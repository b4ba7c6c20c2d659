"""Worker for the multi-GPU tests (one process per GPU, launched by torchrun): FusedExchanger over real
CUDA-IPC peer mappings vs the fp64 mixing-matrix oracle, for every topology / aggregation / target."""
import itertools
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import torch.distributed as dist
    from ldnn_b200.config import Config, resolve
    from ldnn_b200.engine import OptimConfig, TorchEngine
    from ldnn_b200.launch import init_runtime, shutdown
    from ldnn_b200.parallel.exchange import ExchangeSpec, mixing_matrix
    from ldnn_b200.parallel.fused import FusedExchanger
    from test_flat_optimizer import tiny_spec
    from ldnn_b200.ops import _lib as L

    rt = init_runtime("nccl", seed=0)
    ws, r, dev = rt.world_size, rt.rank, rt.device
    eng = TorchEngine(tiny_spec(), dev, OptimConfig(kind="adam", lr=1e-2))
    n = eng.flat.total
    sw = [float(i + 1) for i in range(ws)]

    def all_data(seed):
        g = torch.Generator().manual_seed(seed)
        return torch.randn(ws, n, generator=g, dtype=torch.float64)

    checked = 0
    combos = [("allreduce", "allreduce"), ("ring", "allreduce"), ("double_ring", "allreduce"), ("ring", "gossip"),
              ("double_ring", "gossip")]
    for (topo, sem), agg, target in itertools.product(combos, ["equal", "shard", "weighted"], ["grads", "params"]):
        sp = ExchangeSpec(topo, sem, agg, 0.7, sw, ws)
        ex = FusedExchanger(sp, rt, eng, comm_blocks=8, watchdog_ms=5000)
        for rep in range(2):
            X = all_data(100 * checked + rep)
            buf = eng.flat.grads if target == "grads" else eng.flat.params
            buf.copy_(X[r].float())
            if target == "grads":
                ex.exchange_grads(eng, apply_update=False)
            else:
                ex.exchange_weights(eng)
            ex.check()
            want = (torch.from_numpy(mixing_matrix(sp)) @ X)[r].to(dev)
            err = float((buf.double() - want).abs().max())
            assert err < 1e-5, (topo, sem, agg, target, rep, err)
        ex.close()
        checked += 1

    # fused update tail: identical parameters everywhere, equal to the local oracle
    for topo in ["allreduce", "ring", "double_ring"]:
        sp = ExchangeSpec(topo, "allreduce", "equal", 0.5, None, ws)
        ex = FusedExchanger(sp, rt, eng, comm_blocks=8, watchdog_ms=5000)
        ref = TorchEngine(tiny_spec(), dev, OptimConfig(kind="adam", lr=1e-2))
        eng.flat.params.copy_(all_data(5)[0].float())          # same starting point on every rank
        ref.flat.params.copy_(eng.flat.params)
        eng.flat.m.zero_(); eng.flat.v.zero_(); eng.flat.step = 0
        for step in range(3):
            X = all_data(7000 + step)
            eng.flat.grads.copy_(X[r].float())
            ex.exchange_grads(eng, apply_update=True)
            ex.check()
            ref.flat.optimizer_step(X.mean(0).float().to(dev))
        diff = (eng.flat.params - ref.flat.params).abs()
        assert float((diff > 1e-5).float().mean()) < 1e-3 and float(diff.max()) < 3e-2, (topo, float(diff.max()))
        gathered = [torch.empty_like(eng.flat.params) for _ in range(ws)]
        dist.all_gather(gathered, eng.flat.params)
        assert all(torch.equal(g, gathered[0]) for g in gathered), topo
        # the fused tail keeps moments only for the slice a rank owns; the gather rebuilds the full state
        assert ex.gather_moments(eng), topo
        assert torch.allclose(eng.flat.m, ref.flat.m, atol=1e-6, rtol=1e-3), (topo, "m")
        assert torch.allclose(eng.flat.v, ref.flat.v, atol=1e-8, rtol=1e-3), (topo, "v")
        assert not ex.gather_moments(eng)            # nothing sharded since the last gather
        ex.close()

    # stress: ranks arrive at every exchange with a different, random skew (up to ~1 ms of device-side sleep), the
    # slab is rewritten between calls — flag epochs, write-after-read on the slabs and the staging buffer must hold
    import numpy as np
    rng = np.random.default_rng(1234 + r)
    for topo in ["allreduce", "ring", "double_ring"]:
        sp = ExchangeSpec(topo, "allreduce", "shard", 0.5, sw, ws)
        ex = FusedExchanger(sp, rt, eng, comm_blocks=8, watchdog_ms=5000)
        M = torch.from_numpy(mixing_matrix(sp))
        for it in range(24):
            X = all_data(9000 + it)
            torch.cuda._sleep(int(rng.integers(0, 2_000_000)))
            eng.flat.grads.copy_(X[r].float(), non_blocking=True)
            ex.exchange_grads(eng, apply_update=False)
            got = eng.flat.grads.double().clone()
            want = (M @ X)[r].to(dev)
            assert float((got - want).abs().max()) < 1e-5, (topo, it)
        ex.check()
        ex.close()

    # reference-signature communication.py functions on CUDA tensors of assorted sizes: they run on the fused P2P
    # kernels (staged through a symmetric slab), and — unlike upstream's ring functions on GPU — they
    # change the caller's tensor.  Oracle: the variant's mixing matrix.
    from ldnn_b200 import compat
    from ldnn_b200.parallel.fused import StagedFusedExchange
    launched0 = L.launch_count()
    cases = [("allreduce", "equal", lambda t: compat._ex("allreduce", "equal", ws).exchange_(t)),
             ("allreduce", "weighted", lambda t: compat._ex("allreduce", "weighted", ws, 0.7).exchange_(t)),
             ("ring", "equal", lambda t: compat.ring_all_reduce_equal(t, r, ws)),
             ("ring", "weighted", lambda t: compat.ring_all_reduce_weighted(t, r, ws, 0.7)),
             ("double_ring", "equal", lambda t: compat.double_ring_all_reduce(t, r, ws)),
             ("double_ring", "weighted", lambda t: compat.double_ring_all_reduce_weighted(t, r, ws, 0.7))]
    for topo, kind, fn in cases:
        sem = "allreduce" if topo == "allreduce" else "gossip"
        M = torch.from_numpy(mixing_matrix(ExchangeSpec(topo, sem, kind, 0.7, None, ws)))
        for shape in [(10,), (64, 3, 3, 3), (1024, 1027)]:
            numel = 1
            for v in shape:
                numel *= v
            g = torch.Generator().manual_seed(numel + len(topo))
            X = torch.randn(ws, numel, generator=g, dtype=torch.float64)
            t = X[r].float().view(shape).to(dev)
            assert StagedFusedExchange.usable(None, t)
            fn(t)
            want = (M @ X)[r].view(shape).to(dev)
            assert float((t.double() - want).abs().max()) < 1e-5, (topo, kind, shape)
    assert L.launch_count() > launched0, "compat exchange did not launch our kernels"
    compat.shutdown()                                   # frees the staging slab and pads (collective)
    t = torch.ones(8, device=dev)
    compat.ring_all_reduce_equal(t, r, ws)              # ... and the next call simply builds them again
    assert float((t - 1).abs().max()) < 1e-6

    # broadcast
    sp = ExchangeSpec("allreduce", "allreduce", "equal", 0.5, None, ws)
    ex = FusedExchanger(sp, rt, eng, comm_blocks=8, watchdog_ms=5000)
    eng.flat.params.fill_(float(r + 1))
    ex.broadcast_state(eng, src=0)
    assert float(eng.flat.params.min()) == 1.0 == float(eng.flat.params.max())
    ex.close()
    if r == 0:
        # LDNN_NVLS=1: the allreduce/equal cases above ran on the in-switch (multimem) kernel when `nvls` is True
        print(f"MGPU_OK world={ws} combos={checked} nvls={bool(getattr(ex, 'nvls', False))}")
    shutdown(rt)


if __name__ == "__main__":
    main()

"""`FusedExchanger.gather_moments` on CPU / gloo: the reduction logic (probe marks → cover check → sum of the
owned parts) with the exchange kernels replaced by a host model of slice ownership.  The kernels' own
ownership is what the probe reads on a GPU (tests/mgpu_worker.py)."""
import ctypes as C
import types

import numpy as np
import pytest
import torch


def _owner_two_shot(n4, world, rank):
    per = (n4 + world - 1) // world
    lo = min(rank * per, n4)
    return [(lo, min(lo + per, n4))]


def _owner_ring(n4, world, rank, rings):
    out = []
    for ring in range(rings):
        d = 1 if ring == 0 else -1
        reg_lo, reg_hi = n4 * ring // rings, n4 * (ring + 1) // rings
        chunk = (reg_hi - reg_lo + world - 1) // world
        c = (rank + d) % world                                    # csrc/comm.cu: message k = W-1 finishes chunk r+dir
        lo = min(reg_lo + c * chunk, reg_hi)
        out.append((lo, min(lo + chunk, reg_hi)))
    return out


def _make_exchanger(rank, world, n, mode, buckets, m, v):
    import torch.distributed as dist  # noqa: F401
    from ldnn_b200.ops import _lib as L
    from ldnn_b200.parallel.fused import FusedExchanger

    class HostModel(FusedExchanger):
        def __init__(self):                                       # no CUDA / IPC bring-up
            self.rank, self.world, self.n = rank, world, n
            self.rt = types.SimpleNamespace(device=torch.device("cpu"), rank=rank, world_size=world)
            self._update_buckets = list(buckets)
            self.launched = []

        def _exchange_struct(self, data, mode_, opt=None, lo=0, hi=None):
            assert mode_ == L.X_UPDATE and opt.kind == 1 and opt.lr == 0.0 and opt.momentum == 0.0
            e = types.SimpleNamespace(lo=lo, hi=self.n if hi is None else hi, m=0, v=0, shadow=0)
            return e

        def _launch_reduce(self, e):
            n_el = e.hi - e.lo
            assert n_el % 4 == 0 and e.v == 0 and e.shadow == 0
            probe = np.ctypeslib.as_array((C.c_float * n_el).from_address(e.m))       # e.m already points at the bucket
            n4 = n_el // 4
            if mode == "replicated":
                owned = [(0, n4)]
            elif mode == "two_shot":
                owned = _owner_two_shot(n4, world, rank)
            elif mode == "clash":
                owned = [(0, n4)] if rank == 0 else _owner_two_shot(n4, world, rank)
            else:
                owned = _owner_ring(n4, world, rank, 1 if mode == "ring" else 2)
            for lo4, hi4 in owned:
                probe[4 * lo4:4 * hi4] = 1.0                       # m <- g on what this rank updates
            self.launched.append((e.lo, e.hi))

        def check(self):
            pass

    ex = HostModel()
    ex.engine = types.SimpleNamespace(flat=types.SimpleNamespace(m=m, v=v, grads=torch.zeros(n)))
    return ex


def _worker(rank, world):
    import torch.distributed as dist
    from ldnn_b200.ops import _lib as L
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n = 4 * 1003 + 64                                             # buckets below leave 64 elements of padding uncovered
    buckets = [(0, 4 * 400), (4 * 400, 4 * 1003)]
    g = torch.Generator().manual_seed(7)
    full_m, full_v = torch.randn(n, generator=g), torch.rand(n, generator=g)
    for mode in ["two_shot", "ring", "double_ring"]:
        # what each rank holds before the gather: the true value on its own slices, stale junk elsewhere
        ex = _make_exchanger(rank, world, n, mode, buckets, None, None)
        own = torch.zeros(n, dtype=torch.bool)
        for lo, hi in buckets:
            n4 = (hi - lo) // 4
            rs = _owner_two_shot(n4, world, rank) if mode == "two_shot" else _owner_ring(n4, world, rank, 1 if mode == "ring" else 2)
            for a, b in rs:
                own[lo + 4 * a: lo + 4 * b] = True
        pad = torch.zeros(n, dtype=torch.bool)
        pad[4 * 1003:] = True
        stale = torch.full((n,), 1e9 * (rank + 1))
        m = torch.where(own | pad, full_m, stale).clone()
        v = torch.where(own | pad, full_v, stale).clone()
        ex.engine.flat.m, ex.engine.flat.v = m, v
        assert ex.gather_moments(ex.engine) is True, mode
        assert ex.launched == buckets
        assert torch.allclose(m, full_m, atol=1e-6), (mode, rank, float((m - full_m).abs().max()))
        assert torch.allclose(v, full_v, atol=1e-6), (mode, rank)
        assert float(ex.engine.flat.grads.abs().max()) == 0.0      # the probe gradient is cleared again
        assert ex.gather_moments(ex.engine) is False               # nothing sharded since the gather
    # every rank updates everything (self-weighted form): moments are replicated, nothing is reduced
    m = full_m.clone()
    ex = _make_exchanger(rank, world, n, "replicated", [(0, 4 * 1003)], m, full_v.clone())
    if world > 1:
        assert ex.gather_moments(ex.engine) is False
        assert torch.equal(m, full_m)
    # SGD-momentum: no second moment
    ex = _make_exchanger(rank, world, n, "two_shot", buckets, full_m.clone(), None)
    assert ex.gather_moments(ex.engine) is True
    # two ranks claiming one slice must be reported, not summed
    if world > 2:
        ex = _make_exchanger(rank, world, n, "clash", buckets, full_m.clone(), full_v.clone())
        with pytest.raises(L.LdnnError):
            ex.gather_moments(ex.engine)
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_gather_moments_rebuilds_full_state(dist_runner, world):
    dist_runner(_worker, world)


def test_gather_moments_is_a_noop_without_peers_or_updates():
    ex = _make_exchanger(0, 1, 64, "two_shot", [(0, 64)], torch.zeros(64), torch.zeros(64))
    assert ex.gather_moments(ex.engine) is False
    ex = _make_exchanger(0, 2, 64, "two_shot", [], torch.zeros(64), torch.zeros(64))
    assert ex.gather_moments(ex.engine) is False

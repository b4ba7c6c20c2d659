"""Mixing-matrix oracle and the torch.distributed exchanger."""
import itertools

import numpy as np
import pytest
import torch

from ldnn_b200.parallel.exchange import DistExchanger, ExchangeSpec, gossip_coefficients, mixing_matrix, reduce_form


@pytest.mark.parametrize("ws", [1, 2, 3, 4, 5, 8])
def test_rows_sum_to_one(ws):
    for topo, sem, agg in itertools.product(["allreduce", "ring", "double_ring"], ["allreduce", "gossip"],
                                            ["equal", "weighted", "shard"]):
        sw = list(np.arange(1, ws + 1, dtype=float))
        W = mixing_matrix(ExchangeSpec(topo, sem, agg, 0.7, sw, ws))
        assert np.allclose(W.sum(1), 1.0), (topo, sem, agg)


def test_reference_coefficients():
    W = mixing_matrix(ExchangeSpec("ring", "gossip", "equal", 0.5, None, 5))
    assert W[2, 2] == 0.5 and W[2, 1] == 0.5 and W[0, 4] == 0.5
    W = mixing_matrix(ExchangeSpec("double_ring", "gossip", "equal", 0.5, None, 5))
    assert np.allclose([W[3, 3], W[3, 2], W[3, 1]], 1 / 3)
    W = mixing_matrix(ExchangeSpec("double_ring", "gossip", "weighted", 0.6, None, 5))
    assert np.allclose([W[0, 0], W[0, 4], W[0, 3]], [0.6, 0.2, 0.2])
    W = mixing_matrix(ExchangeSpec("allreduce", "allreduce", "weighted", 0.6, None, 5))
    assert np.allclose(W[1], [0.1, 0.6, 0.1, 0.1, 0.1])
    # double ring at ws=2 → (2·self + other)/3
    W = mixing_matrix(ExchangeSpec("double_ring", "gossip", "equal", 0.5, None, 2))
    assert np.allclose(W, [[1 / 3, 2 / 3], [2 / 3, 1 / 3]]) or np.allclose(W, [[2 / 3, 1 / 3], [1 / 3, 2 / 3]])
    # r-1 and r-2 at ws=2: r-2 == r → self gets 2/3
    assert np.isclose(W[0, 0], 2 / 3)
    # shard weights 4:2:1:1
    W = mixing_matrix(ExchangeSpec("ring", "allreduce", "shard", 0.5, [4, 2, 1, 1], 4))
    assert np.allclose(W[0], [0.5, 0.25, 0.125, 0.125]) and np.allclose(W[0], W[3])


def test_reduce_form_matches_matrix():
    for agg in ["equal", "weighted", "shard"]:
        sp = ExchangeSpec("allreduce", "allreduce", agg, 0.3, [3, 1, 2, 2], 4)
        W = mixing_matrix(sp)
        for r in range(4):
            a, b = reduce_form(sp, r)
            row = np.array(b)
            row[r] += a
            assert np.allclose(row, W[r])


COMBOS = [("allreduce", "allreduce"), ("ring", "gossip"), ("double_ring", "gossip"), ("ring", "allreduce"),
          ("double_ring", "allreduce")]


def _exchange_worker(rank, ws):
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    sw = [float(i + 1) for i in range(ws)]
    g = torch.Generator().manual_seed(1234)
    X = torch.randn(ws, 1001, generator=g, dtype=torch.float64)
    for (topo, sem), agg in itertools.product(COMBOS, ["equal", "weighted", "shard"]):
        sp = ExchangeSpec(topo, sem, agg, 0.7, sw, ws)
        t = X[rank].clone().float()
        DistExchanger(sp, rank).exchange_(t)
        want = torch.from_numpy(mixing_matrix(sp)) @ X
        assert torch.allclose(t.double(), want[rank], atol=1e-5), (topo, sem, agg, rank)
    dist.destroy_process_group()


@pytest.mark.parametrize("ws", [2, 3])
def test_dist_exchanger_matches_oracle(dist_runner, ws):
    dist_runner(_exchange_worker, ws)


def test_ring_chain_plan_and_flag_free_pad_sizes():
    """Host-side arithmetic of the ring kernels (native library, no GPU needed): the bulk ring's chains cover a rank's
    chunk, respect the 16 KB floor and the flag table; the flag-free ring's slots per phase fit the pad the exchanger
    allocates for its largest payload."""
    import ctypes as C
    from ldnn_b200.ops import _lib as L
    lib = L.lib()
    lib.ldnn_ring_plan.argtypes = [C.c_longlong, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_longlong)]
    for n in (4, 4 * 1000, 1 << 20, 44_595_904, 1 << 28):
        for world in (1, 2, 3, 8):
            for rings in (1, 2):
                for blocks, chains in ((1, 1), (8, 1), (64, 1), (64, 2), (64, 8)):
                    J, sub4 = C.c_int(0), C.c_longlong(0)
                    assert lib.ldnn_ring_plan(n, world, rings, blocks, chains, C.byref(J), C.byref(sub4)) == 0
                    chunk4 = -(-(n // 4 // rings + (1 if (n // 4) % rings else 0)) // world)      # ceil(ceil(n4/rings)/world)
                    assert J.value >= 1 and J.value * sub4.value >= chunk4, (n, world, rings, blocks, chains)
                    assert J.value <= (L.SIGNAL_WORDS - 4096 - 64) // 2
                    assert sub4.value >= 1024 or J.value == (L.SIGNAL_WORDS - 4096 - 64) // 2
                    groups = blocks * 4 // rings
                    assert J.value <= max(1, chains * groups) + 1                               # ~`chains` per group
    assert lib.ldnn_ring_plan(6, 2, 1, 8, 1, C.byref(J), C.byref(sub4)) != 0                     # n % 4
    ll_max = 64 << 20
    phase4 = 2 * (ll_max // 16) + 4096                                                           # FusedExchanger._ll_buf
    for world in (2, 3, 4, 8):
        for rings in (1, 2):
            for n in (4 * world, 4 * 12345, ll_max // 4):
                used = lib.ldnn_ring_ll_slots(n, world, rings)
                assert 0 < used <= phase4, (n, world, rings, used)

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

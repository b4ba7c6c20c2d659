"""Property tests (hypothesis) for the pure-python semantics layers: mixing matrices / reduce forms /
gossip coefficients (parallel/exchange.py) and the sharding functions (data.py)."""
import numpy as np
from hypothesis import given, settings, strategies as st

from ldnn_b200 import data as D
from ldnn_b200.parallel.exchange import ExchangeSpec, gossip_coefficients, mixing_matrix, reduce_form

topologies = st.sampled_from(["allreduce", "ring", "double_ring"])
semantics = st.sampled_from(["allreduce", "gossip"])
aggs = st.sampled_from(["equal", "weighted", "shard"])
worlds = st.integers(min_value=1, max_value=8)
lw = st.floats(min_value=0.05, max_value=0.95)


def _spec(topo, sem, agg, w, ws, sizes):
    sw = [s / sum(sizes[:ws]) for s in sizes[:ws]]
    return ExchangeSpec(topo, sem, agg, w, sw, ws)


sizes8 = st.lists(st.integers(min_value=1, max_value=1000), min_size=8, max_size=8)


@settings(max_examples=200, deadline=None)
@given(topologies, semantics, aggs, lw, worlds, sizes8)
def test_mixing_matrix_is_row_stochastic_and_nonnegative(topo, sem, agg, w, ws, sizes):
    M = mixing_matrix(_spec(topo, sem, agg, w, ws, sizes))
    assert M.shape == (ws, ws)
    assert np.all(M >= -1e-15)
    assert np.allclose(M.sum(axis=1), 1.0, atol=1e-12)
    if ws == 1:
        assert np.allclose(M, 1.0)                      # identity at world size 1


@settings(max_examples=200, deadline=None)
@given(topologies, aggs, lw, worlds, sizes8)
def test_allreduce_semantics_reduce_form_reproduces_matrix(topo, agg, w, ws, sizes):
    sp = _spec(topo, "allreduce", agg, w, ws, sizes)
    M = mixing_matrix(sp)
    for r in range(ws):
        a, b = reduce_form(sp, r)
        row = np.asarray(b, dtype=np.float64).copy()
        row[r] += a
        assert np.allclose(row, M[r], atol=1e-12), (r, row, M[r])
    if agg != "weighted":
        assert np.allclose(M, M[0])                     # rank-independent result: every rank gets the same mean
    if agg == "shard":
        assert np.allclose(M[0], np.asarray(sp.shard_weights), atol=1e-12)
    if agg == "equal":
        assert np.allclose(M, 1.0 / ws)


@settings(max_examples=200, deadline=None)
@given(st.sampled_from(["ring", "double_ring"]), st.sampled_from(["equal", "weighted"]), lw, worlds)
def test_gossip_coefficients_reproduce_matrix_rows(topo, agg, w, ws):
    sp = ExchangeSpec(topo, "gossip", agg, w, None, ws)
    M = mixing_matrix(sp)
    for r in range(ws):
        c_self, nbrs = gossip_coefficients(sp, r)
        row = np.zeros(ws)
        row[r] += c_self
        for d, c in nbrs:
            row[(r - d) % ws] += c
        assert np.allclose(row, M[r], atol=1e-12), (topo, agg, ws, r)


@settings(max_examples=200, deadline=None)
@given(st.integers(min_value=1, max_value=5000), st.lists(st.floats(min_value=0.05, max_value=10), min_size=1, max_size=8))
def test_balanced_ranges_are_disjoint_contiguous_and_drop_less_than_one_sample_per_rank(total, ratios):
    """`int(total*ratio)` sizes (dataloader.py:62-72): consecutive, disjoint; only the rounding remainder is unused."""
    ratios = D.normalise_ratios(ratios)
    seen = np.zeros(total, dtype=np.int64)
    prev_hi = 0
    for r in range(len(ratios)):
        lo, hi = D.range_for_rank(total, r, ratios)
        assert lo == prev_hi and lo <= hi <= total
        assert hi - lo == int(total * ratios[r])
        prev_hi = hi
        seen[D.partition_balanced(total, r, ratios)] += 1
    assert total - prev_hi <= len(ratios)
    assert np.all(seen <= 1)                            # no sample is given to two ranks


@settings(max_examples=100, deadline=None)
@given(st.integers(min_value=2, max_value=6), st.floats(min_value=0.0, max_value=1.0), st.integers(min_value=0, max_value=10_000))
def test_class_skew_partition_has_the_ranks_size_and_reaches_the_fixed_share(ws, fixed_ratio, seed):
    rng = np.random.default_rng(seed)
    labels = rng.integers(0, 10, size=2000)
    ratios = [1.0 / ws] * ws
    for r in range(ws):
        lo, hi = D.range_for_rank(len(labels), r, ratios)
        idx, fixed = D.partition_class_skew(labels, r, ratios, fixed_ratio, rng=np.random.default_rng(seed + r))
        assert fixed == D.fixed_classes_for(r)
        assert len(idx) == hi - lo                      # shard keeps its size whatever the skew
        assert idx.min() >= 0 and idx.max() < len(labels)
        frac = np.isin(labels[idx], fixed).mean()
        assert frac >= fixed_ratio - 0.01               # upstream draws extra fixed-class samples (with replacement)

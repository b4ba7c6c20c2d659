"""Sharding / re-sharding invariants (S-DATA)."""
import numpy as np

from ldnn_b200 import data as D


def test_balanced_ranges_tile_the_dataset():
    ratios = [0.4, 0.3, 0.2, 0.1]
    parts = [D.partition_balanced(1000, r, ratios) for r in range(4)]
    assert [len(p) for p in parts] == [400, 300, 200, 100]
    assert np.array_equal(np.concatenate(parts), np.arange(1000))


def test_explicit_ratios_4211():
    r = D.normalise_ratios([4, 2, 1, 1])
    assert [len(D.partition_balanced(800, i, r)) for i in range(4)] == [400, 200, 100, 100]


def test_class_skew_reaches_fixed_ratio():
    rng = np.random.default_rng(0)
    labels = rng.integers(0, 10, size=5000)
    for rank in range(5):
        idx, fixed = D.partition_class_skew(labels, rank, [0.2] * 5, 0.5, 10, rng)
        assert fixed == [(2 * rank) % 10, (2 * rank + 1) % 10]
        assert len(idx) == 1000
        frac = np.isin(labels[idx], fixed).mean()
        assert 0.48 <= frac <= 0.52, frac


def test_next_partition_sizes_and_overlap():
    rng = np.random.default_rng(1)
    prev = np.arange(100, 300)
    nxt = D.next_partition(1000, prev, 0.5, 0.5, ratio=0.2, replace=False, rng=rng)
    assert len(nxt) == 200 and len(np.unique(nxt)) == 200
    assert np.isin(nxt, prev).sum() >= 100          # at least prev_fraction comes from the old shard
    nxt = D.next_partition(1000, prev, 0.25, 0.5, ratio=0.2, replace=True, rng=rng)
    assert len(nxt) == 150


def test_next_partition_restores_class_skew():
    rng = np.random.default_rng(2)
    labels = rng.integers(0, 10, size=4000)
    prev, fixed = D.partition_class_skew(labels, 1, [0.25] * 4, 0.5, 10, rng)
    nxt = D.next_partition(4000, prev, 0.5, 0.5, 0.25, True, rng, labels, fixed, 0.5)
    assert len(nxt) == 1000
    assert np.isin(labels[nxt], fixed).mean() >= 0.49


def test_loader_keeps_partial_batch_and_order():
    b = D.make_synthetic(100, 10, 10)
    ld = D.Loader(b.trainset, np.arange(10, 80), 32)
    sizes = [len(y) for _, y in ld]
    assert sizes == [32, 32, 6] and len(ld) == 3
    x, y = next(iter(ld))
    assert x.shape == (32, 3, 32, 32) and x.dtype.is_floating_point and y.dtype.__str__() == "torch.int64"
    assert np.array_equal(y.numpy(), b.trainset.labels[10:42])

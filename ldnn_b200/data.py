"""Datasets, sharding and re-sharding (S-DATA).

Behavioural contract comes from the reference dataloaders
(``Balanced All-Reduce/dataloader.py:9-153``, ``Disbalanced Ring/dataloader.py:54-165``)
but the mechanism is different: everything operates on a **label array** and index
arrays with vectorised numpy, instead of indexing ``dataset[i][1]`` (which upstream
runs the full PIL transform pipeline just to read a label, and makes the re-shard
O(N·M)).  Batches are staged through pinned memory and copied on a side stream.

Sharding modes
* balanced      — contiguous ``[start,end)`` range sized by per-rank ratios
* disbalanced   — same range, then class-skewed so that two "fixed classes"
                  ``[(2r)%C, (2r+1)%C]`` make up ``fixed_ratio`` of the shard
* shard_ratios  — explicit sizes such as 4:2:1:1 (north-star "uneven shards");
                  the same ratios feed the shard-weighted averaging ``w_r = n_r/Σn``
Ratios otherwise come from the timing probe (``estimate_epoch_duration``) exactly
as upstream, including its direction (slower rank ⇒ larger shard).
"""
from __future__ import annotations

import math
import os
import time
from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.distributed as dist


# ---------------------------------------------------------------------------
# datasets
# ---------------------------------------------------------------------------
class ArrayDataset:
    """Images as one tensor + labels as one numpy array.

    ``images`` is either float32 NCHW already normalised, or uint8 NCHW with
    ``mean/std`` applied at batch time (optionally after augmentation)."""

    def __init__(self, images: torch.Tensor, labels: np.ndarray, mean=None, std=None, augment=None):
        assert images.shape[0] == len(labels)
        self.images = images
        self.labels = np.asarray(labels, dtype=np.int64)
        self.mean = None if mean is None else torch.as_tensor(mean, dtype=torch.float32).view(1, -1, 1, 1)
        self.std = None if std is None else torch.as_tensor(std, dtype=torch.float32).view(1, -1, 1, 1)
        self.augment = augment

    def __len__(self) -> int:
        return len(self.labels)

    def __getitem__(self, i: int):
        """(image, label) like a torch Dataset — for code that indexes samples one by one."""
        x, y = self.batch(np.asarray([int(i)]))
        return x[0], int(y[0])

    def to(self, device) -> "ArrayDataset":
        """Keep the images resident on `device` (a CIFAR-10 shard is ~120 MB of uint8 against 180 GB of HBM):
        batches are then gathered, augmented and normalised on the device and no per-step H2D copy is left."""
        device = torch.device(device)
        self.images = self.images.to(device)
        if self.mean is not None:
            self.mean, self.std = self.mean.to(device), self.std.to(device)
        return self

    @property
    def device(self) -> torch.device:
        return self.images.device

    def batch(self, idx: np.ndarray) -> Tuple[torch.Tensor, torch.Tensor]:
        it = torch.as_tensor(np.asarray(idx, dtype=np.int64)).to(self.images.device)
        x = self.images.index_select(0, it)
        if x.dtype == torch.uint8:
            if self.augment is not None:
                if getattr(self.augment, "batched", False):
                    x = self.augment(x)                       # whole uint8 batch at once, on the data's device
                else:
                    x = torch.stack([self.augment(im) for im in x])
            x = x.float().div_(255.0)
            if self.mean is not None:
                x = (x - self.mean) / self.std
        y = torch.from_numpy(self.labels[np.asarray(idx, dtype=np.int64)])
        if x.device.type != "cpu":
            y = y.to(x.device, non_blocking=True)
        return x, y

    def subset_view(self, idx: np.ndarray) -> "ArrayDataset":
        """Index-remapped view (used for the train/val split)."""
        it = torch.as_tensor(np.asarray(idx, dtype=np.int64))
        ds = ArrayDataset(self.images.index_select(0, it), self.labels[np.asarray(idx)], None, None, self.augment)
        ds.mean, ds.std = self.mean, self.std
        return ds


class SubsetView:
    """Index view of an :class:`ArrayDataset` with the surface of ``torch.utils.data.Subset``
    (``.dataset``, ``.indices``, ``len``, ``[i]``) — what upstream's ``get_dataset_partition`` returns."""

    def __init__(self, dataset: ArrayDataset, indices):
        self.dataset, self.indices = dataset, np.asarray(indices, dtype=np.int64)
        self.labels = dataset.labels[self.indices]

    def __len__(self) -> int:
        return len(self.indices)

    def __getitem__(self, i: int):
        return self.dataset[int(self.indices[i])]

    def batch(self, idx: np.ndarray):
        return self.dataset.batch(self.indices[np.asarray(idx, dtype=np.int64)])


@dataclass
class DataBundle:
    trainset: ArrayDataset
    valset: ArrayDataset
    testset: ArrayDataset
    num_classes: int
    source: str

    def to(self, device) -> "DataBundle":
        for ds in (self.trainset, self.valset, self.testset):
            ds.to(device)
        return self

    def nbytes(self) -> int:
        return sum(ds.images.numel() * ds.images.element_size() for ds in (self.trainset, self.valset, self.testset))


def make_synthetic(n_train=4096, n_val=1024, n_test=1024, num_classes=10, hw=32, channels=3,
                   seed=0, signal: float = 0.25) -> DataBundle:
    """Synthetic data of the reference's batch shape (float32 ``[B,3,32,32]``, int64
    labels in ``[0,10)``; ``dataloader.py:14-20,42``).  Images are a class-dependent
    low-frequency pattern plus unit noise so that loss curves are meaningful."""
    g = torch.Generator().manual_seed(seed)
    protos = torch.randn(num_classes, channels, 4, 4, generator=g)
    protos = torch.nn.functional.interpolate(protos, size=(hw, hw), mode="bilinear", align_corners=False)

    def draw(n, s):
        gg = torch.Generator().manual_seed(seed * 1000 + s)
        y = torch.randint(0, num_classes, (n,), generator=gg)
        # `signal` sets the difficulty: 0.25 is learnt to ~100 % within an epoch (plumbing tests); ~0.03 leaves the
        # Bayes accuracy well below 100 %, so accuracy can discriminate between engines (test_gpu_training_quality)
        x = signal * protos[y] + torch.randn(n, channels, hw, hw, generator=gg)
        return ArrayDataset(x.contiguous(), y.numpy())

    return DataBundle(draw(n_train, 1), draw(n_val, 2), draw(n_test, 3), num_classes, "synthetic")


def load_cifar10(root="data", augment="batched", rng: Optional[np.random.Generator] = None) -> DataBundle:
    """CIFAR-10 with the reference pipeline: dataset mean/std, AutoAugment on the
    training images, 80/20 train/val split (``dataloader.py:10-35``).  Needs the
    archive on disk (no network here) — callers fall back to synthetic.

    ``augment``: "batched" (default; `ops.augment.BatchedAutoAugment`, same pixel math applied to whole
    batches on the data's device), "per_image" / True (torchvision's transform one image at a time, as
    upstream), "none" / False."""
    import torchvision
    from torchvision import transforms
    full = torchvision.datasets.CIFAR10(root, train=True, download=False)
    test = torchvision.datasets.CIFAR10(root, train=False, download=False)
    arr = full.data.astype(np.float32) / 255.0
    mean, std = arr.mean(axis=(0, 1, 2)), arr.std(axis=(0, 1, 2))
    rng = rng or np.random.default_rng()
    perm = rng.permutation(len(full))
    n_tr = int(0.8 * len(full))
    imgs = torch.from_numpy(full.data).permute(0, 3, 1, 2).contiguous()
    labels = np.asarray(full.targets, dtype=np.int64)
    if augment in ("batched",):
        from .ops.augment import BatchedAutoAugment
        aug = BatchedAutoAugment("cifar10")
    elif augment in (True, "per_image"):
        aug = transforms.AutoAugment(transforms.AutoAugmentPolicy.CIFAR10)
    elif augment in (False, None, "none"):
        aug = None
    else:
        raise ValueError(f"augment={augment!r}: expected batched | per_image | none")
    train = ArrayDataset(imgs[perm[:n_tr]], labels[perm[:n_tr]], mean, std, aug)
    val = ArrayDataset(imgs[perm[n_tr:]], labels[perm[n_tr:]], mean, std, aug)  # upstream: valset shares the train transform
    timgs = torch.from_numpy(test.data).permute(0, 3, 1, 2).contiguous()
    tst = ArrayDataset(timgs, np.asarray(test.targets, dtype=np.int64), mean, std, None)
    return DataBundle(train, val, tst, 10, "cifar10")


def get_data(cfg, rank=0) -> DataBundle:
    if not cfg.synthetic:
        try:
            return load_cifar10(cfg.data_root, augment=getattr(cfg, "augment", "batched"))
        except Exception as e:  # no archive / no torchvision → synthetic, loudly
            if rank == 0 and not cfg.quiet:
                print(f"[data] CIFAR-10 unavailable ({type(e).__name__}: {e}); using synthetic data")
    seed = 0 if cfg.seed is None else cfg.seed
    hw = 224 if cfg.model == "resnet50" else 32
    return make_synthetic(cfg.synthetic_train, cfg.synthetic_val, cfg.synthetic_test, cfg.num_classes,
                          hw=hw, seed=seed)


def place_bundle(bundle: DataBundle, cfg, rt, cap_bytes: int = 16 << 30) -> bool:
    """Device residency of the dataset (``--data_on_device``): ``auto`` moves the uint8 on-disk dataset onto the
    rank's GPU (gather + augmentation + normalisation then run there and the per-step H2D copy disappears),
    ``on`` moves any dataset that fits, ``off`` keeps the pinned double-buffer path.  Returns True if moved."""
    mode = getattr(cfg, "data_on_device", "auto")
    if mode == "off" or not getattr(rt, "is_cuda", False) or bundle.nbytes() > cap_bytes:
        return False
    if mode == "auto" and bundle.trainset.images.dtype != torch.uint8:
        return False
    bundle.to(rt.device)
    return True


def chunk_batches_for(cfg, dataset: ArrayDataset) -> int:
    """How many batches a Loader gathers per call: only augmented datasets gain from large chunks."""
    k = int(getattr(cfg, "chunk_batches", 0) or 0)
    if k > 0:
        return k
    if not getattr(dataset.augment, "batched", False):
        return 1
    return 64 if dataset.images.device.type == "cuda" else 16


# ---------------------------------------------------------------------------
# sharding (pure functions on sizes / label arrays)
# ---------------------------------------------------------------------------
def fixed_classes_for(rank: int, num_classes: int = 10) -> List[int]:
    return [(rank * 2) % num_classes, (rank * 2 + 1) % num_classes]


def range_for_rank(total: int, rank: int, ratios: Sequence[float]) -> Tuple[int, int]:
    """Contiguous range; sizes ``int(total*ratio_i)`` (``dataloader.py:62-72``)."""
    start = 0
    for i, r in enumerate(ratios):
        n = int(total * r)
        if i == rank:
            return start, start + n
        start += n
    raise ValueError(f"rank {rank} outside ratios of length {len(ratios)}")


def partition_balanced(total: int, rank: int, ratios: Sequence[float]) -> np.ndarray:
    s, e = range_for_rank(total, rank, ratios)
    return np.arange(s, e, dtype=np.int64)


def partition_class_skew(labels: np.ndarray, rank: int, ratios: Sequence[float], fixed_ratio: float,
                         num_classes: int = 10, rng: Optional[np.random.Generator] = None
                         ) -> Tuple[np.ndarray, List[int]]:
    """Class-skewed shard (``Disbalanced Ring/dataloader.py:54-103``)."""
    rng = rng or np.random.default_rng()
    total = len(labels)
    s, e = range_for_rank(total, rank, ratios)
    fixed = fixed_classes_for(rank, num_classes)
    in_range = np.arange(s, e, dtype=np.int64)
    is_fixed = np.isin(labels[s:e], fixed)
    fixed_idx, other_idx = in_range[is_fixed], in_range[~is_fixed]
    want_fixed = int(round((e - s) * fixed_ratio))
    if len(fixed_idx) < want_fixed:
        pool_mask = np.isin(labels, fixed)
        pool_mask[fixed_idx] = False
        pool = np.nonzero(pool_mask)[0]
        if len(pool):
            fixed_idx = np.concatenate([fixed_idx, rng.choice(pool, size=want_fixed - len(fixed_idx), replace=True)])
    excess = len(fixed_idx) + len(other_idx) - (e - s)
    if excess > 0:
        other_idx = other_idx[:len(other_idx) - excess] if excess < len(other_idx) else other_idx[:0]
    out = np.concatenate([fixed_idx, other_idx]).astype(np.int64)
    rng.shuffle(out)
    return out, fixed


def next_partition(total: int, prev_indices: np.ndarray, prev_fraction: float, next_fraction: float,
                   ratio: float, replace: bool, rng: Optional[np.random.Generator] = None,
                   labels: Optional[np.ndarray] = None, fixed_classes: Optional[Sequence[int]] = None,
                   fixed_ratio: float = 0.5) -> np.ndarray:
    """Per-global-epoch re-shard: keep ``prev_fraction`` of a freshly sized shard from
    the old indices and draw ``next_fraction`` from the rest of the dataset
    (``Balanced All-Reduce/dataloader.py:77-104``); with ``fixed_classes`` the class
    skew is restored afterwards (``Disbalanced Ring/dataloader.py:105-153``)."""
    rng = rng or np.random.default_rng()
    prev_indices = np.asarray(prev_indices, dtype=np.int64)
    node_points = int(total * ratio)
    n_prev, n_next = int(node_points * prev_fraction), int(node_points * next_fraction)
    if not replace:
        n_prev = min(n_prev, len(prev_indices))
    prev_sel = rng.choice(prev_indices, size=n_prev, replace=replace) if n_prev and len(prev_indices) else np.empty(0, np.int64)
    mask = np.ones(total, dtype=bool)
    mask[prev_sel] = False
    remaining = np.nonzero(mask)[0]
    if not replace:
        n_next = min(n_next, len(remaining))
    next_sel = rng.choice(remaining, size=n_next, replace=replace) if n_next and len(remaining) else np.empty(0, np.int64)
    final = np.concatenate([prev_sel, next_sel]).astype(np.int64)

    if fixed_classes is not None and labels is not None and len(final):
        is_fixed = np.isin(labels[final], list(fixed_classes))
        want = int((n_prev + n_next) * fixed_ratio)
        short = want - int(is_fixed.sum())
        if short > 0:
            pool_mask = np.isin(labels, list(fixed_classes))
            pool_mask[final] = False
            pool = np.nonzero(pool_mask)[0]
            replaceable = np.nonzero(~is_fixed)[0]
            short = min(short, len(replaceable))
            if len(pool) and short > 0:
                final[replaceable[len(replaceable) - short:]] = rng.choice(pool, size=short, replace=True)
    rng.shuffle(final)
    return final


def normalise_ratios(r: Sequence[float]) -> List[float]:
    s = float(sum(r))
    return [float(x) / s for x in r]


# ---------------------------------------------------------------------------
# loaders
# ---------------------------------------------------------------------------
class Loader:
    """``shuffle=False``, keeps the last partial batch (``dataloader.py:42,46,49``)."""

    def __init__(self, dataset: ArrayDataset, indices: Optional[np.ndarray], batch_size: int, chunk_batches: int = 1):
        self.dataset = dataset
        self.indices = np.arange(len(dataset), dtype=np.int64) if indices is None else np.asarray(indices, np.int64)
        self.batch_size = batch_size
        # gather / augment / normalise `chunk_batches` batches in one call (amortises the grouped augmentation),
        # then hand them out one by one — same batches, same order
        self.chunk_batches = max(1, int(chunk_batches))

    def __len__(self) -> int:
        return math.ceil(len(self.indices) / self.batch_size)

    def num_samples(self) -> int:
        return len(self.indices)

    def __iter__(self):
        bs, span = self.batch_size, self.batch_size * self.chunk_batches
        for i in range(0, len(self.indices), span):
            x, y = self.dataset.batch(self.indices[i:i + span])
            if self.chunk_batches == 1:
                yield x, y
                continue
            for j in range(0, x.shape[0], bs):
                yield x[j:j + bs], y[j:j + bs]


class DeviceFeeder:
    """Pinned double-buffer + copy stream in front of a :class:`Loader`.

    Replaces the reference's synchronous pageable ``inputs.to(device)``
    (``trainer.py:203``): batch *i+1* is copied while batch *i* computes."""

    def __init__(self, loader: Loader, device: torch.device):
        self.loader, self.device = loader, device
        self.cuda = device.type == "cuda"
        self.stream = torch.cuda.Stream(device) if self.cuda else None
        # two pinned staging slots, reused (a fresh pin_memory() per batch costs a cudaHostAlloc each time)
        self._pin = [dict(), dict()]
        self._done = [None, None]
        self._slot = 0

    def __len__(self):
        return len(self.loader)

    def _pinned(self, slot: int, t: torch.Tensor) -> torch.Tensor:
        key = (tuple(t.shape), t.dtype)
        buf = self._pin[slot].get(key)
        if buf is None:
            buf = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
            self._pin[slot][key] = buf
        return buf

    def _stage(self, x, y):
        if not self.cuda:
            return x, y, None
        if x.device == self.device:                  # dataset resident on the device: nothing to copy
            return x, y.to(self.device, non_blocking=True), None
        slot = self._slot
        self._slot ^= 1
        if self._done[slot] is not None:
            self._done[slot].synchronize()          # the copy that last used this slot has finished
        xp, yp = self._pinned(slot, x), self._pinned(slot, y)
        xp.copy_(x)
        yp.copy_(y)
        with torch.cuda.stream(self.stream):
            xd = xp.to(self.device, non_blocking=True)
            yd = yp.to(self.device, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self.stream)
        self._done[slot] = ev
        return xd, yd, (ev, xp, yp)

    def __iter__(self):
        it = iter(self.loader)
        nxt = None
        try:
            nxt = self._stage(*next(it))
        except StopIteration:
            return
        while nxt is not None:
            xd, yd, keep = nxt
            try:
                upcoming = self._stage(*next(it))
            except StopIteration:
                upcoming = None
            if keep is not None:
                torch.cuda.current_stream(self.device).wait_event(keep[0])
                xd.record_stream(torch.cuda.current_stream(self.device))
                yd.record_stream(torch.cuda.current_stream(self.device))
            yield xd, yd
            nxt = upcoming


# ---------------------------------------------------------------------------
# timing probe + shard bookkeeping
# ---------------------------------------------------------------------------
def all_gather_floats(value: float, rt) -> List[float]:
    if rt.world_size == 1:
        return [float(value)]
    t = torch.tensor([value], dtype=torch.float64, device=rt.device if rt.backend == "nccl" else "cpu")
    out = [torch.zeros_like(t) for _ in range(rt.world_size)]
    dist.all_gather(out, t)
    return [float(o.item()) for o in out]


def estimate_epoch_duration(engine, loader: Loader, rt, num_batches: int = 10) -> Tuple[float, List[float]]:
    """Timing probe (``dataloader.py:119-153``): ``num_batches`` forward/backward passes,
    durations all-gathered, ``ratio_i = d_i / Σd`` (C2)."""
    t0 = time.time()
    feeder = DeviceFeeder(loader, rt.device)
    for i, (x, y) in enumerate(feeder):
        if i >= num_batches:
            break
        engine.probe_step(x, y)
    if rt.is_cuda:
        torch.cuda.synchronize(rt.device)
    d = max(time.time() - t0, 1e-9)
    alld = all_gather_floats(d, rt)
    tot = sum(alld)
    return d / tot, [x / tot for x in alld]


@dataclass
class ShardState:
    """What a rank needs to carry between global epochs (and into a checkpoint)."""
    train_indices: np.ndarray
    val_indices: np.ndarray
    ratios: List[float]
    fixed_classes: Optional[List[int]] = None


def initial_shards(bundle: DataBundle, cfg, rt, engine=None, rng=None) -> ShardState:
    """``get_loaders`` minus the loaders (``dataloader.py:37-46``)."""
    rng = rng or np.random.default_rng(None if cfg.seed is None else cfg.seed + 17 * rt.rank)
    if cfg.shard_ratios is not None:
        assert len(cfg.shard_ratios) == rt.world_size, "--shard_ratios needs one entry per rank"
        ratios = normalise_ratios(cfg.shard_ratios)
    elif engine is not None and cfg.probe_batches > 0:
        tmp = Loader(bundle.trainset, rng.permutation(len(bundle.trainset))[: cfg.batch_size * cfg.probe_batches],
                     cfg.batch_size)
        _, ratios = estimate_epoch_duration(engine, tmp, rt, cfg.probe_batches)
    else:
        ratios = [1.0 / rt.world_size] * rt.world_size
    fixed = None
    if cfg.balance == "disbalanced":
        tr, fixed = partition_class_skew(bundle.trainset.labels, rt.rank, ratios, cfg.fixed_ratio, bundle.num_classes, rng)
        va, _ = partition_class_skew(bundle.valset.labels, rt.rank, ratios, cfg.fixed_ratio, bundle.num_classes, rng)
    else:
        tr = partition_balanced(len(bundle.trainset), rt.rank, ratios)
        va = partition_balanced(len(bundle.valset), rt.rank, ratios)
    return ShardState(tr, va, list(ratios), fixed)


def reshard(bundle: DataBundle, st: ShardState, cfg, rt, epoch_duration: float, total_duration: float,
            rng=None) -> ShardState:
    """``get_subset_loaders`` (``dataloader.py:107-117``): ratio = own duration / Σ durations,
    or the explicit shard ratio when ``--shard_ratios`` is given."""
    rng = rng or np.random.default_rng()
    ratio = st.ratios[rt.rank] if cfg.shard_ratios is not None else epoch_duration / max(total_duration, 1e-12)
    kw = dict(prev_fraction=cfg.prev_fraction, next_fraction=cfg.next_fraction, ratio=ratio,
              replace=cfg.replace, rng=rng, fixed_ratio=cfg.fixed_ratio)
    tr = next_partition(len(bundle.trainset), st.train_indices, labels=bundle.trainset.labels,
                        fixed_classes=st.fixed_classes, **kw)
    va = next_partition(len(bundle.valset), st.val_indices, labels=bundle.valset.labels,
                        fixed_classes=st.fixed_classes, **kw)
    return ShardState(tr, va, st.ratios, st.fixed_classes)

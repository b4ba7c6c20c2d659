"""ops/augment.BatchedAutoAugment: grouped application == torchvision's per-image AutoAugment for the same
random choices (CIFAR-10 policy, reference dataloader.py:14-20), on uint8 batches."""
import numpy as np
import pytest
import torch

tv = pytest.importorskip("torchvision")

from ldnn_b200.ops.augment import BatchedAutoAugment


def _per_image_reference(aug, x, pid, u, sign):
    from torchvision.transforms.autoaugment import _apply_op
    space = aug._ref._augmentation_space(10, (x.shape[2], x.shape[3]))
    out = []
    for i in range(x.shape[0]):
        img = x[i]
        for k, (name, p, mag_id) in enumerate(aug.policies[int(pid[i])]):
            if float(u[i, k]) <= p:
                mags, signed = space[name]
                m = float(mags[mag_id].item()) if mag_id is not None else 0.0
                if signed and int(sign[i, k]) == 0:
                    m *= -1.0
                img = _apply_op(img, name, m, interpolation=aug.interpolation, fill=None)
        out.append(img)
    return torch.stack(out)


def test_batched_equals_per_image_for_every_subpolicy():
    aug = BatchedAutoAugment("cifar10")
    g = torch.Generator().manual_seed(0)
    n = 25 * 8
    x = torch.randint(0, 256, (n, 3, 32, 32), dtype=torch.uint8, generator=g)
    pid = torch.arange(n) % 25                                   # every sub-policy, 8 images each
    u = torch.zeros(n, 2)                                        # always apply both operations
    sign = torch.randint(0, 2, (n, 2), generator=g)
    got = aug.apply(x, pid, u, sign)
    want = _per_image_reference(aug, x, pid, u, sign)
    assert got.dtype == torch.uint8 and got.shape == x.shape
    assert torch.equal(got, want)
    assert not torch.equal(got, x)


def test_batched_equals_per_image_with_sampled_choices():
    aug = BatchedAutoAugment("cifar10")
    g = torch.Generator().manual_seed(3)
    x = torch.randint(0, 256, (300, 3, 32, 32), dtype=torch.uint8, generator=g)
    pid, u, sign = aug.sample(300, g)
    assert int(pid.min()) >= 0 and int(pid.max()) < 25 and set(np.unique(sign.numpy())) <= {0, 1}
    assert torch.equal(aug.apply(x, pid, u, sign), _per_image_reference(aug, x, pid, u, sign))
    # u above every probability: nothing is applied
    same = aug.apply(x, pid, torch.full((300, 2), 2.0), sign)
    assert torch.equal(same, x)


def test_call_is_seedable_and_changes_most_images():
    aug = BatchedAutoAugment("cifar10")
    x = torch.randint(0, 256, (256, 3, 32, 32), dtype=torch.uint8, generator=torch.Generator().manual_seed(1))
    a = aug(x, torch.Generator().manual_seed(5))
    b = aug(x, torch.Generator().manual_seed(5))
    c = aug(x, torch.Generator().manual_seed(6))
    assert torch.equal(a, b) and not torch.equal(a, c)
    changed = (a != x).flatten(1).any(1).float().mean()
    assert 0.5 < float(changed) <= 1.0                            # most sub-policies fire with p ~ 0.5+


def test_array_dataset_uses_the_batched_path():
    from ldnn_b200.data import ArrayDataset
    aug = BatchedAutoAugment("cifar10")
    imgs = torch.randint(0, 256, (128, 3, 32, 32), dtype=torch.uint8, generator=torch.Generator().manual_seed(2))
    ds = ArrayDataset(imgs, np.arange(128) % 10, [0.5] * 3, [0.25] * 3, aug)
    x, y = ds.batch(np.arange(64))
    assert x.shape == (64, 3, 32, 32) and x.dtype == torch.float32 and torch.isfinite(x).all()
    assert np.array_equal(y.numpy(), np.arange(64) % 10)


def test_chunked_loader_yields_the_same_batches_in_the_same_order():
    from ldnn_b200.data import ArrayDataset, Loader
    g = torch.Generator().manual_seed(4)
    ds = ArrayDataset(torch.randn(203, 3, 8, 8, generator=g), np.arange(203) % 10)
    idx = np.random.default_rng(0).permutation(203)[:150]
    plain = list(Loader(ds, idx, 16))
    for k in (2, 3, 64):
        chunked = list(Loader(ds, idx, 16, chunk_batches=k))
        assert len(chunked) == len(plain) == 10                        # 9 full batches + the partial one
        for (xa, ya), (xb, yb) in zip(plain, chunked):
            assert torch.equal(xa, xb) and torch.equal(ya, yb)
    assert len(Loader(ds, idx, 16, chunk_batches=4)) == 10


def test_place_bundle_and_chunk_policy():
    import types
    from ldnn_b200 import data as D
    from ldnn_b200.config import Config
    u8 = D.ArrayDataset(torch.zeros(8, 3, 4, 4, dtype=torch.uint8), np.zeros(8), [0.5] * 3, [0.2] * 3, BatchedAutoAugment())
    f32 = D.ArrayDataset(torch.zeros(8, 3, 4, 4), np.zeros(8))
    cpu_rt = types.SimpleNamespace(is_cuda=False, device=torch.device("cpu"))
    assert D.place_bundle(D.DataBundle(u8, u8, u8, 10, "cifar10"), Config(), cpu_rt) is False      # no GPU: stays on host
    fake_gpu = types.SimpleNamespace(is_cuda=True, device=torch.device("cpu"))                     # "device" = cpu so .to() works here
    assert D.place_bundle(D.DataBundle(f32, f32, f32, 10, "synthetic"), Config(), fake_gpu) is False   # auto: uint8 only
    assert D.place_bundle(D.DataBundle(f32, f32, f32, 10, "synthetic"), Config(data_on_device="on"), fake_gpu) is True
    assert D.place_bundle(D.DataBundle(u8, u8, u8, 10, "cifar10"), Config(), fake_gpu) is True
    assert D.place_bundle(D.DataBundle(u8, u8, u8, 10, "cifar10"), Config(data_on_device="off"), fake_gpu) is False
    assert D.place_bundle(D.DataBundle(u8, u8, u8, 10, "cifar10"), Config(), fake_gpu, cap_bytes=10) is False
    assert D.chunk_batches_for(Config(), f32) == 1                       # nothing to amortise without augmentation
    assert D.chunk_batches_for(Config(), u8) == 16                       # host-side batched augmentation
    assert D.chunk_batches_for(Config(chunk_batches=5), u8) == 5
    x, y = u8.to("cpu").batch(np.arange(4))
    assert x.shape == (4, 3, 4, 4) and y.shape == (4,)


def test_batched_equalize_is_bit_identical_to_torchvision():
    from torchvision.transforms import functional as F
    from ldnn_b200.ops.augment import batched_equalize
    g = torch.Generator().manual_seed(9)
    x = torch.randint(0, 256, (40, 3, 32, 32), dtype=torch.uint8, generator=g)
    x[0] = 7                                                       # constant plane: step == 0 -> unchanged
    x[1, 0] = torch.randint(100, 104, (32, 32), dtype=torch.uint8, generator=g)    # very narrow histogram
    x[2] = (x[2] // 64) * 64                                       # four levels
    x[3, 1, :, :16] = 0; x[3, 1, :, 16:] = 255                      # two extreme levels
    want = torch.stack([F.equalize(im) for im in x])
    got = batched_equalize(x)
    assert got.dtype == torch.uint8 and torch.equal(got, want)
    small = torch.randint(0, 256, (5, 1, 4, 4), dtype=torch.uint8, generator=g)      # fewer pixels than bins
    assert torch.equal(batched_equalize(small), torch.stack([F.equalize(im) for im in small]))

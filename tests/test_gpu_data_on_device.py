"""`--data_on_device`: the uint8 dataset lives in HBM, the grouped AutoAugment + normalisation run on the GPU and the
loader hands device tensors to the engine (no per-step H2D copy).  The CPU suite covers the grouping logic
(tests/test_batched_augment.py); this file runs the same code on CUDA tensors and trains from it."""
import time
import types

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
tv = pytest.importorskip("torchvision")


def test_batched_autoaugment_on_cuda_matches_cpu():
    from ldnn_b200.ops.augment import BatchedAutoAugment, batched_equalize
    aug = BatchedAutoAugment("cifar10")
    g = torch.Generator().manual_seed(0)
    n = 25 * 16
    x = torch.randint(0, 256, (n, 3, 32, 32), dtype=torch.uint8, generator=g)
    pid = torch.arange(n) % 25                                   # every sub-policy
    u = torch.zeros(n, 2)
    sign = torch.randint(0, 2, (n, 2), generator=g)
    want = aug.apply(x, pid, u, sign)
    got = aug.apply(x.cuda(), pid, u, sign)
    assert got.is_cuda and got.dtype == torch.uint8
    # integer colour operations are bit-identical; the affine ones resample through grid_sample, whose float
    # rounding may move a border pixel between the CPU and the CUDA implementation
    diff = (got.cpu().int() - want.int()).abs()
    assert float((diff > 0).float().mean()) < 5e-3, float((diff > 0).float().mean())
    assert torch.equal(batched_equalize(x.cuda()).cpu(), batched_equalize(x))


def test_training_from_a_device_resident_uint8_dataset():
    from ldnn_b200 import data as D
    from ldnn_b200.config import Config
    from ldnn_b200.engine import OptimConfig
    from ldnn_b200.models import enhanced_cnn_spec
    from ldnn_b200.models.native import NativeEngine
    from ldnn_b200.ops.augment import BatchedAutoAugment
    g = torch.Generator().manual_seed(1)
    n = 64 * 40
    labels = (np.arange(n) % 10).astype(np.int64)
    # class-dependent brightness so that a few steps of training visibly lower the loss
    imgs = (torch.randint(0, 96, (n, 3, 32, 32), generator=g) + torch.from_numpy(labels).view(-1, 1, 1, 1) * 16).to(torch.uint8)
    mk = lambda aug: D.ArrayDataset(imgs.clone(), labels, [0.45] * 3, [0.25] * 3, aug)
    bundle = D.DataBundle(mk(BatchedAutoAugment("cifar10")), mk(None), mk(None), 10, "cifar10-like")
    rt = types.SimpleNamespace(is_cuda=True, device=torch.device("cuda", 0))
    cfg = Config(data_on_device="auto")
    assert D.place_bundle(bundle, cfg, rt) is True and bundle.trainset.images.is_cuda
    loader = D.Loader(bundle.trainset, np.arange(n), 64, D.chunk_batches_for(cfg, bundle.trainset))
    assert D.chunk_batches_for(cfg, bundle.trainset) == 64
    eng = NativeEngine(enhanced_cnn_spec(), "cuda:0", OptimConfig(lr=1e-3), batch_size=64)
    losses = []
    for x, y in D.DeviceFeeder(loader, rt.device):
        assert x.is_cuda and x.dtype == torch.float32 and y.is_cuda and x.shape == (64, 3, 32, 32)
        loss, _ = eng.train_step(x, y)
        losses.append(loss.detach().reshape(()).clone())
    losses = torch.stack(losses).cpu()
    assert torch.isfinite(losses).all() and float(losses[-8:].mean()) < 0.85 * float(losses[:4].mean()), losses.tolist()
    # loader alone: gather + grouped augmentation + normalisation on the GPU
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    k = 0
    for x, y in loader:
        k += x.shape[0]
    torch.cuda.synchronize()
    rate = k / (time.perf_counter() - t0)
    print(f"[data_on_device] GPU-side chunked AutoAugment loader: {rate:.0f} images/s")
    assert rate > 20000

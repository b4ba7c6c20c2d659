"""The real-data path (`data.load_cifar10`, reference `dataloader.py:10-35`) against a tiny fake
CIFAR-10 archive written in the python-pickle layout torchvision reads (no network here)."""
import os
import pickle

import numpy as np
import pytest
import torch

tv = pytest.importorskip("torchvision")

from ldnn_b200 import data as D
from ldnn_b200.config import Config


def _write_fake_cifar(root, per_batch=20, seed=0):
    rng = np.random.default_rng(seed)
    base = os.path.join(root, "cifar-10-batches-py")
    os.makedirs(base)
    names = [f"data_batch_{i}" for i in range(1, 6)] + ["test_batch"]
    for n in names:
        blob = {"data": rng.integers(0, 256, size=(per_batch, 3072), dtype=np.uint8),
                "labels": rng.integers(0, 10, size=per_batch).tolist()}
        with open(os.path.join(base, n), "wb") as f:
            pickle.dump(blob, f)
    with open(os.path.join(base, "batches.meta"), "wb") as f:
        pickle.dump({"label_names": [f"c{i}" for i in range(10)]}, f)


@pytest.fixture()
def fake_cifar(tmp_path, monkeypatch):
    _write_fake_cifar(str(tmp_path))
    # the fake archive cannot match the published md5s: integrity = "file exists"
    monkeypatch.setattr(tv.datasets.CIFAR10, "_check_integrity", lambda self: True)
    monkeypatch.setattr(tv.datasets.cifar, "check_integrity", lambda path, md5=None: os.path.exists(path))
    return str(tmp_path)


def test_load_cifar10_split_normalisation_and_batches(fake_cifar):
    b = D.load_cifar10(fake_cifar, augment=False, rng=np.random.default_rng(3))
    assert b.source == "cifar10" and b.num_classes == 10
    assert len(b.trainset) == 80 and len(b.valset) == 20 and len(b.testset) == 20      # 80/20 split of 100
    x, y = b.trainset.batch(np.arange(16))
    assert x.shape == (16, 3, 32, 32) and x.dtype == torch.float32 and y.dtype == torch.int64
    # dataset statistics are applied: the normalised training pixels are ~zero-mean / unit-std per channel
    allx, _ = b.trainset.batch(np.arange(80))
    assert allx.mean(dim=(0, 2, 3)).abs().max() < 0.1
    assert (allx.std(dim=(0, 2, 3)) - 1).abs().max() < 0.1
    # test images are not augmented and use the training statistics
    tx, ty = b.testset.batch(np.arange(20))
    assert tx.shape == (20, 3, 32, 32) and int(ty.min()) >= 0 and int(ty.max()) < 10


def test_load_cifar10_autoaugment_keeps_shape_and_labels(fake_cifar):
    b = D.load_cifar10(fake_cifar, augment=True, rng=np.random.default_rng(3))
    torch.manual_seed(0)
    x, y = b.trainset.batch(np.arange(8))
    assert x.shape == (8, 3, 32, 32) and torch.isfinite(x).all()
    assert np.array_equal(y.numpy(), b.trainset.labels[:8])
    img, lab = b.trainset[0]
    assert img.shape == (3, 32, 32) and lab == int(b.trainset.labels[0])


def test_get_data_prefers_disk_and_falls_back_to_synthetic(fake_cifar, tmp_path):
    cfg = Config(synthetic=False, data_root=fake_cifar, quiet=True)
    assert D.get_data(cfg).source == "cifar10"
    cfg2 = Config(synthetic=False, data_root=str(tmp_path / "nowhere"), quiet=True,
                  synthetic_train=64, synthetic_val=32, synthetic_test=32)
    b = D.get_data(cfg2)
    assert b.source == "synthetic" and len(b.trainset) == 64


def test_loader_and_sharding_on_cifar_bundle(fake_cifar):
    b = D.load_cifar10(fake_cifar, augment=False, rng=np.random.default_rng(1))
    idx = D.partition_balanced(len(b.trainset), 1, [0.5, 0.5])
    assert len(idx) == 40
    ld = D.Loader(b.trainset, idx, batch_size=16)
    assert len(ld) == 3 and ld.num_samples() == 40            # last partial batch kept (dataloader.py:42)
    seen = 0
    for x, y in ld:
        assert x.shape[1:] == (3, 32, 32)
        seen += len(y)
    assert seen == 40


def test_cli_trains_on_disk_dataset_single_process(fake_cifar, tmp_path, monkeypatch):
    """Whole program on the real-data path (no --synthetic): uint8 images, AutoAugment, dataset
    statistics, 80/20 split, world_size 1 (exchange is the identity)."""
    monkeypatch.setenv("LDNN_TINY_WIDTHS", "1")
    monkeypatch.setenv("MASTER_ADDR", "127.0.0.1")
    monkeypatch.setenv("MASTER_PORT", str(29900 + os.getpid() % 90))
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        monkeypatch.delenv(k, raising=False)
    from ldnn_b200 import cli
    out = tmp_path / "graphs"
    h = cli.main(["--backend", "gloo", "--data_root", fake_cifar, "--epochs_global", "2", "--epochs_local", "1",
                  "--batch_size", "16", "--output_folder", str(out), "--quiet", "--seed", "5"],
                 topology="allreduce", balance="balanced")
    assert len(h.global_train_losses) == 2 and all(np.isfinite(h.global_train_losses))
    assert os.path.exists(out / "training_metrics.png") and os.path.exists(out / "metrics.json")


def test_phase_trace_is_written_in_chrome_format(tmp_path, monkeypatch):
    import json
    monkeypatch.setenv("LDNN_TINY_WIDTHS", "1")
    monkeypatch.setenv("MASTER_ADDR", "127.0.0.1")
    monkeypatch.setenv("MASTER_PORT", str(29800 + os.getpid() % 90))
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        monkeypatch.delenv(k, raising=False)
    from ldnn_b200 import cli
    cli.main(["--backend", "gloo", "--synthetic", "--synthetic_train", "64", "--synthetic_val", "32", "--synthetic_test", "32",
              "--epochs_global", "2", "--epochs_local", "1", "--batch_size", "32", "--probe_batches", "1", "--quiet",
              "--output_folder", str(tmp_path / "g"), "--trace", str(tmp_path / "tl")], topology="allreduce", balance="balanced")
    blob = json.load(open(tmp_path / "tl.rank0.json"))
    names = [e["name"] for e in blob["traceEvents"]]
    assert names.count("exchange") == 2 and names.count("validate") == 2 and "bring-up" in names and "graphs" in names
    assert sum(n.startswith("train g") for n in names) == 2
    assert all(e["ph"] == "X" and e["dur"] >= 0 and e["pid"] == 0 for e in blob["traceEvents"])

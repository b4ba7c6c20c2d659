"""Upstream-signature API exported by the six variant directories (CPU, single process)."""
import importlib.util
import os

import numpy as np
import pytest
import torch
import torch.nn as nn

from conftest import ROOT

VARIANTS = {"Balanced All-Reduce": ("allreduce", False), "Balanced Ring": ("ring", False),
            "Balanced Double-Ring": ("double_ring", False), "Disbalanced All-Reduce": ("allreduce", True),
            "Disbalanced Ring": ("ring", True), "Disalanced Double-Ring": ("double_ring", True)}
FILES = ["main", "trainer", "communication", "dataloader", "model", "validator", "evaluator", "vizualizator"]


def load(variant, name):
    path = os.path.join(ROOT, variant, name + ".py")
    spec = importlib.util.spec_from_file_location(f"{variant.replace(' ', '_').replace('-', '_')}_{name}", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.mark.parametrize("variant", list(VARIANTS))
def test_every_variant_has_the_eight_modules(variant):
    topo, dis = VARIANTS[variant]
    for f in FILES:
        assert os.path.exists(os.path.join(ROOT, variant, f + ".py")), (variant, f)
    comm = load(variant, "communication")
    want = {"allreduce": ["average_gradients_equal", "average_gradients_weighted", "average_weights_equal",
                          "average_weights_weighted"],
            "ring": ["ring_all_reduce_equal", "ring_all_reduce_weighted"],
            "double_ring": ["double_ring_all_reduce", "double_ring_all_reduce_weighted"]}[topo]
    for fn in want:
        assert callable(getattr(comm, fn))
    tr = load(variant, "trainer")
    assert callable(tr.train_global) and callable(tr.train_local_epoch)
    m = load(variant, "model")
    assert len(m.EnhancedCNNModel().state_dict()) == 128
    main = load(variant, "main")
    assert (main.TOPOLOGY, main.BALANCE) == (topo, "disbalanced" if dis else "balanced")


def test_single_process_exchange_is_identity():
    comm = load("Balanced Ring", "communication")
    t = torch.arange(10.0)
    comm.ring_all_reduce_equal(t, 0, 1)
    comm.ring_all_reduce_weighted(t, 0, 1, 0.3)
    assert torch.equal(t, torch.arange(10.0))
    ar = load("Balanced All-Reduce", "communication")
    lin = nn.Linear(3, 2)
    lin.weight.grad = torch.ones_like(lin.weight)
    ar.average_gradients_weighted(lin, 1, 0.5)        # upstream divides by zero here
    assert torch.equal(lin.weight.grad, torch.ones_like(lin.weight))


def test_reference_style_script_flow(monkeypatch, tmp_path):
    """get_loaders → train_local_epoch → validate → evaluate → plots, with upstream call signatures."""
    monkeypatch.setenv("LDNN_SYNTHETIC", "1")
    from ldnn_b200 import compat
    compat._BUNDLE.clear()
    from ldnn_b200.models import SpecNet
    from test_flat_optimizer import tiny_spec
    dl = load("Disbalanced Ring", "dataloader")
    tr = load("Disbalanced Ring", "trainer")
    va = load("Disbalanced Ring", "validator")
    ev = load("Disbalanced Ring", "evaluator")
    vz = load("Disbalanced Ring", "vizualizator")

    class Net(nn.Module):                     # 32x32 input → tiny_spec expects 8x8: pool first
        def __init__(self):
            super().__init__()
            self.net = SpecNet(tiny_spec())

        def forward(self, x):
            return self.net(torch.nn.functional.avg_pool2d(x, 4))
    model = Net()
    device = torch.device("cpu")
    out = dl.get_loaders(32, 1, 0, model, device, 0.5)
    train_loader, val_loader, test_loader, trainset, valset, tr_idx, va_idx, fixed = out
    assert fixed == [0, 1] and len(tr_idx) > 0
    crit = nn.CrossEntropyLoss()
    opt = torch.optim.Adam(model.parameters(), lr=1e-3)
    sched = torch.optim.lr_scheduler.StepLR(opt, step_size=25)
    small = type(train_loader)(trainset, np.asarray(tr_idx)[:96], 32)
    loss, acc, losses = tr.train_local_epoch(model, small, crit, opt, device, sched)
    assert len(losses) == 3 and np.isfinite(loss) and 0 <= acc <= 100
    vsmall = type(val_loader)(valset, np.asarray(va_idx)[:64], 32)
    vl, vacc = va.validate(model, vsmall, crit, device)
    assert np.isfinite(vl)
    tl, tacc, preds, labels = ev.evaluate(model, type(test_loader)(test_loader.dataset, np.arange(64), 32), crit, device, 0)
    assert len(preds) == len(labels) == 64
    # full train_global with the DisR positional order (comm, local_weight, fixed_classes, …)
    res = tr.train_global(model, small, vsmall, trainset, valset, np.asarray(tr_idx)[:96], np.asarray(va_idx)[:64], crit,
                          opt, sched, device, 0, 1, 1, 2, 60, 32, 0.5, 0.5, object(), 0.5, fixed, "equal", "weights", 0.5)
    assert len(res) == 12 and len(res[4]) == 2
    vz.plot_metrics_global(2, res[4], res[5], res[6], res[7], output_folder=str(tmp_path))
    vz.plot_loss_distribution_by_worker(res[0], output_folder=str(tmp_path))
    assert (tmp_path / "training_metrics.png").stat().st_size > 1000


def test_prf_scores_match_sklearn():
    from sklearn.metrics import f1_score, precision_score, recall_score
    from ldnn_b200.trainer import prf_scores
    rng = np.random.default_rng(0)
    y, p = rng.integers(0, 10, 500), rng.integers(0, 10, 500)
    s = prf_scores(y, p, 10)
    for avg in ("macro", "weighted", "micro"):
        want = (precision_score(y, p, average=avg, zero_division=0), recall_score(y, p, average=avg, zero_division=0),
                f1_score(y, p, average=avg, zero_division=0))
        assert np.allclose(s[avg], want, atol=1e-9), avg


def test_partition_functions_return_subset_like_objects(monkeypatch):
    monkeypatch.setenv("LDNN_SYNTHETIC", "1")
    from ldnn_b200 import compat, data as D
    b = D.make_synthetic(400, 50, 50)
    sub, idx = compat.get_dataset_partition(b.trainset, 1, [0.25] * 4)
    assert len(sub) == 100 == len(idx) and list(sub.indices) == list(range(100, 200))
    x, y = sub[3]
    assert x.shape == (3, 32, 32) and y == int(b.trainset.labels[103])
    sub2, idx2, fixed = compat.get_dataset_partition(b.trainset, 1, [0.25] * 4, 0.5)
    assert fixed == [2, 3] and len(sub2) == len(idx2)
    sub3, idx3 = compat.get_next_dataset_partition(b.trainset, idx, 0.5, 0.5, 1.0, torch.tensor([4.0]))
    assert len(sub3) == 100 and len(idx3) == 100
    ld = D.Loader(sub3, None, 32)                      # a SubsetView is loadable like a dataset
    assert sum(len(yy) for _, yy in ld) == 100


def test_evaluate_copies_predictions_out_of_engine_static_buffers():
    """The native engine returns a view of its step-static prediction buffer; evaluate() must not alias it."""
    import types
    import numpy as np
    import torch
    from ldnn_b200 import data as D
    from ldnn_b200.trainer import evaluate

    class StaticBufEngine:
        def __init__(self):
            self.pred = torch.zeros(8, dtype=torch.int64)

        def eval_step(self, x, y):
            self.pred[: len(y)] = y                          # "predicts" the labels, into the same buffer every step
            return torch.tensor(0.5), (self.pred[: len(y)] == y).sum(), self.pred[: len(y)]

    labels = np.arange(24) % 10
    ds = D.ArrayDataset(torch.zeros(24, 3, 4, 4), labels)
    rt = types.SimpleNamespace(device=torch.device("cpu"), rank=0)
    loss, acc, preds, got_labels = evaluate(StaticBufEngine(), D.Loader(ds, None, 8), rt, 10, verbose=False)
    assert np.array_equal(preds, labels) and np.array_equal(got_labels, labels) and acc == 100.0 and loss == 0.5

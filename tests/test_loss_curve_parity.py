"""Loss-curve parity: the reference's own model + torch.optim.Adam + StepLR trained with its
per-batch sequence (zero_grad / forward / CrossEntropyLoss / backward / step, `Balanced All-Reduce/trainer.py:203-218`)
against this framework's engine on the flat slabs — identical initial state_dict, identical synthetic batches."""
import importlib
import os
import sys

import pytest
import torch

REF_DIRS = [os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "baseline", "_ref", "Balanced All-Reduce"),
            "/root/reference/Balanced All-Reduce"]
REF = next((d for d in REF_DIRS if os.path.isdir(d)), None)


@pytest.mark.skipif(REF is None, reason="reference tree not available")
def test_torch_engine_follows_the_reference_training_loop():
    from ldnn_b200.engine import OptimConfig, TorchEngine
    from ldnn_b200.models import enhanced_cnn_spec
    sys.path.insert(0, REF)
    try:
        ref = importlib.import_module("model").EnhancedCNNModel()
    finally:
        sys.path.pop(0)
        sys.modules.pop("model", None)
    torch.manual_seed(0)
    eng = TorchEngine(enhanced_cnn_spec(), "cpu", OptimConfig(kind="adam", lr=1e-3))
    eng.load_state_dict(ref.state_dict())
    opt = torch.optim.Adam(ref.parameters(), lr=1e-3)                     # main.py:53
    sched = torch.optim.lr_scheduler.StepLR(opt, step_size=25, gamma=0.1)  # main.py:54
    crit = torch.nn.CrossEntropyLoss()                                     # main.py:52
    g = torch.Generator().manual_seed(42)
    ref.train()
    ref_losses, our_losses = [], []
    for step in range(4):
        x = torch.randn(8, 3, 32, 32, generator=g)
        y = torch.randint(0, 10, (8,), generator=g)
        opt.zero_grad()
        loss = crit(ref(x), y)
        loss.backward()
        opt.step()
        ref_losses.append(float(loss.detach()))
        l, _ = eng.train_step(x, y)
        our_losses.append(float(l))
    sched.step()
    eng.scheduler_step()
    assert our_losses == pytest.approx(ref_losses, rel=2e-4, abs=2e-4), (our_losses, ref_losses)
    assert eng.flat.lr == pytest.approx(opt.param_groups[0]["lr"])
    # and the resulting weights / running statistics are the reference's
    sd_ref, sd_ours = ref.state_dict(), eng.state_dict()
    for k in sd_ref:
        if k.endswith("num_batches_tracked"):
            assert int(sd_ref[k]) == int(sd_ours[k])
        else:
            assert torch.allclose(sd_ref[k], sd_ours[k].to(sd_ref[k].dtype), atol=2e-4, rtol=2e-3), k

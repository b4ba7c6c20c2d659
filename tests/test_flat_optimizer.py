"""FlatState optimizer math vs torch.optim (oracle for K-OPT)."""
import pytest
import torch

from ldnn_b200.engine import OptimConfig, TorchEngine
from ldnn_b200.models import SpecNet, enhanced_cnn_spec
from ldnn_b200.models.spec import ConvBN, Head, NetSpec


def tiny_spec():
    s = NetSpec("tiny", 3, 8)
    s.units.append(ConvBN("prep.0", "prep.1", 3, 8, 3, 1, 1, "x", "a0"))
    s.units.append(ConvBN("layer1.0.conv1", "layer1.0.bn1", 8, 16, 3, 2, 1, "a0", "a1"))
    s.units.append(ConvBN("layer1.0.shortcut.0", "layer1.0.shortcut.1", 8, 16, 1, 2, 0, "a0", "s1", relu=False))
    s.units.append(ConvBN("layer1.0.conv2", "layer1.0.bn2", 16, 16, 3, 1, 1, "a1", "a2", residual="s1"))
    s.units.append(Head("fc", 16, 10, "a2"))
    return s


@pytest.mark.parametrize("kind", ["adam", "sgd"])
def test_matches_torch_optim(kind):
    torch.manual_seed(0)
    eng = TorchEngine(tiny_spec(), "cpu", OptimConfig(kind=kind, lr=1e-2, lr_step=2))
    ref = SpecNet(tiny_spec())
    ref.load_state_dict(eng.state_dict())
    opt = (torch.optim.Adam(ref.parameters(), lr=1e-2) if kind == "adam"
           else torch.optim.SGD(ref.parameters(), lr=1e-2, momentum=0.9))
    sched = torch.optim.lr_scheduler.StepLR(opt, step_size=2)
    for it in range(6):
        x, y = torch.randn(8, 3, 8, 8), torch.randint(0, 10, (8,))
        loss, _ = eng.train_step(x, y)
        opt.zero_grad()
        l2 = torch.nn.functional.cross_entropy(ref(x), y)
        l2.backward()
        opt.step()
        assert torch.allclose(loss, l2.detach(), atol=1e-5)
        if it % 2 == 1:
            eng.scheduler_step(); sched.step()
    for (k, a), (_, b) in zip(eng.model.named_parameters(), ref.named_parameters()):
        assert torch.allclose(a, b, atol=2e-5), k

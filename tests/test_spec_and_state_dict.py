"""state_dict layout parity with the reference network."""
import os
import sys

import pytest
import torch

from ldnn_b200.models import EnhancedCNNModel, ResNet50, enhanced_cnn_spec, resnet50_spec


def test_enhanced_cnn_layout():
    m = EnhancedCNNModel()
    sd = m.state_dict()
    assert len(sd) == 128
    params = list(m.named_parameters())
    assert len(params) == 65
    assert sum(p.numel() for _, p in params) == 44_595_786 == enhanced_cnn_spec().num_params()
    assert sum(1 for k in sd if k.endswith("num_batches_tracked")) == 21
    assert sd["prep.0.weight"].shape == (64, 3, 3, 3)
    assert sd["layer4.1.conv2.weight"].shape == (1024, 1024, 3, 3)
    assert sd["layer1.0.shortcut.0.weight"].shape == (128, 64, 1, 1)
    assert sd["fc.weight"].shape == (10, 1024) and sd["fc.bias"].shape == (10,)
    keys = list(sd.keys())
    # shortcut registers after conv2/bn2 inside a block
    assert keys.index("layer1.0.bn2.weight") < keys.index("layer1.0.shortcut.0.weight")
    assert all(v.dtype == torch.float32 for k, v in sd.items() if not k.endswith("num_batches_tracked"))
    assert all(v.dtype == torch.int64 for k, v in sd.items() if k.endswith("num_batches_tracked"))


@pytest.mark.skipif(not os.path.isdir("/root/reference/Balanced All-Reduce"), reason="reference tree not mounted")
def test_matches_reference_module():
    sys.path.insert(0, "/root/reference/Balanced All-Reduce")
    try:
        import importlib
        refm = importlib.import_module("model")
        ref = refm.EnhancedCNNModel()
    finally:
        sys.path.pop(0)
        sys.modules.pop("model", None)
    ours = EnhancedCNNModel()
    assert list(ref.state_dict().keys()) == list(ours.state_dict().keys())
    ours.load_state_dict(ref.state_dict())
    ref.load_state_dict(ours.state_dict())
    x = torch.randn(3, 3, 32, 32)
    ref.eval(); ours.eval()
    assert torch.equal(ref(x), ours(x))
    ref.train(); ours.train()
    assert torch.allclose(ref(x), ours(x))


def test_resnet50_layout():
    spec = resnet50_spec()
    assert spec.num_params() == 25_557_032
    m = ResNet50()
    assert "layer1.0.downsample.0.weight" in m.state_dict()
    y = m(torch.randn(1, 3, 64, 64))
    assert y.shape == (1, 1000)


def test_shape_inference_and_flop_count():
    from ldnn_b200.models import enhanced_cnn_spec, resnet50_spec
    s = enhanced_cnn_spec()
    shp = s.slot_shapes()
    assert shp["x"] == (32, 32, 3) and shp[s.head().src] == (2, 2, 1024)
    assert s.forward_flops(1) == 1_077_301_248 and s.forward_flops(64) == 64 * s.forward_flops(1)
    r = resnet50_spec()
    assert r.slot_shapes()[r.head().src] == (7, 7, 2048)
    assert abs(r.forward_flops(1) / 2e9 - 4.09) < 0.02            # the usual 4.1 GMACs of ResNet-50 at 224


def test_resnet50_matches_torchvision_layout_and_forward():
    """M-R50: same 320-entry state_dict (names, shapes, order) and the same function as torchvision's ResNet-50 v1.5."""
    tvm = pytest.importorskip("torchvision.models")
    ref = tvm.resnet50(weights=None)
    ours = ResNet50()
    rs, os_ = ref.state_dict(), ours.state_dict()
    assert list(rs.keys()) == list(os_.keys())
    assert all(rs[k].shape == os_[k].shape for k in rs)
    ours.load_state_dict(rs)
    x = torch.randn(2, 3, 64, 64)
    ref.eval(); ours.eval()
    with torch.no_grad():
        assert torch.allclose(ref(x), ours(x), atol=1e-5, rtol=1e-5)

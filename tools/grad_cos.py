"""Per-tensor gradient agreement NativeEngine (bf16) vs TorchEngine (fp32) — debugging aid."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from test_gpu_engine import _engines, _batch
from ldnn_b200.models import enhanced_cnn_spec
te, ne = _engines(enhanced_cnn_spec, 64, use_graph=False)
x, y = _batch(64, 1)
lt, _ = te.train_step(x, y, update=False); ln, _ = ne.train_step(x, y, update=False)
print("loss", float(lt), float(ln))
for s in te.flat.slots:
    gt = te.flat.view(te.flat.grads, s.key).float(); gn = ne.export_tensor(ne.flat.grads, s.key).cuda().float()
    cos = float((gt * gn).sum() / (gt.norm() * gn.norm() + 1e-20))
    print(f"{s.key:32s} cos {cos:.4f} ratio {float(gn.norm() / (gt.norm() + 1e-20)):.4f} |g| {float(gt.norm()):.3e}")
from test_gpu_engine import _autocast_grads
amp = _autocast_grads(te, x, y)
print("---- torch bf16 autocast vs fp32")
for s in te.flat.slots:
    gt = te.flat.view(te.flat.grads, s.key).float(); ga = amp[s.key]
    print(f"{s.key:32s} cos {float((gt * ga).sum() / (gt.norm() * ga.norm() + 1e-20)):.4f}")

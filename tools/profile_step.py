"""A few eager (no CUDA graph) training steps of the flagship model, for ncu launch lists / captures."""
import os, sys
os.environ["LDNN_NO_GRAPH"] = "1"
os.environ.setdefault("LDNN_NO_AUTOTUNE", "1")   # event timing under ncu would mis-tune
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ldnn_b200.engine import OptimConfig
from ldnn_b200.models import enhanced_cnn_spec
from ldnn_b200.models.native import NativeEngine
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
dtype = sys.argv[2] if len(sys.argv) > 2 else "tf32"
eng = NativeEngine(enhanced_cnn_spec(), "cuda:0", OptimConfig(), batch_size=64, dtype=dtype)
x = torch.randn(64, 3, 32, 32, device="cuda"); y = torch.randint(0, 10, (64,), device="cuda")
for _ in range(steps):
    loss, _ = eng.train_step(x, y)
torch.cuda.synchronize()
print("loss", float(loss), "launches/step", eng.launches_per_step())

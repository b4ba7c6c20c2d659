import os, sys
sys.path.insert(0, os.getcwd())
import torch
from ldnn_b200.engine import OptimConfig
from ldnn_b200.models import resnet50_spec
from ldnn_b200.models.native import NativeEngine
from ldnn_b200.ops import native_ops as K
B = 32
x = torch.randn(B, 3, 224, 224, device="cuda"); y = torch.randint(0, 1000, (B,), device="cuda")
eng = NativeEngine(resnet50_spec(), "cuda:0", OptimConfig(), batch_size=B)
for _ in range(5): eng.train_step(x, y)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(20): eng.train_step(x, y)
b.record(); torch.cuda.synchronize()
print("ms/step", a.elapsed_time(b) / 20)
names = {0: "fprop", 1: "dgrad", 2: "wgrad"}
nl = sum(1 for v in K._TUNE_CACHE.values() if v[2]); print("tuned entries", len(K._TUNE_CACHE), "lite chosen", nl)
for k, v in list(K._TUNE_CACHE.items())[:60]:
    print(names[k[0]], k[1:], "->", v)

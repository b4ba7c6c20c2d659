"""Per-op time of one ResNet-50 training step (eager, CUDA events around every op of the plan)."""
import os, sys, collections
os.environ["LDNN_NO_GRAPH"] = "1"; os.environ["LDNN_SERIAL"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ldnn_b200.engine import OptimConfig
from ldnn_b200.models import resnet50_spec, enhanced_cnn_spec
from ldnn_b200.models.native import NativeEngine
which = sys.argv[1] if len(sys.argv) > 1 else "resnet50"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 32
spec = resnet50_spec() if which == "resnet50" else enhanced_cnn_spec()
hw = spec.in_hw
eng = NativeEngine(spec, "cuda:0", OptimConfig(), batch_size=B)
x = torch.randn(B, 3, hw, hw, device="cuda"); y = torch.randint(0, 10, (B,), device="cuda")
for _ in range(3): eng.train_step(x, y, update=False)
p = eng._plan(B, True)
agg = collections.defaultdict(lambda: [0, 0.0])
for rep in range(3):
    for fn, side in p.ops:
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        name = getattr(fn, "__qualname__", str(fn))
        if "lambda" in name:
            import inspect
            try: src = inspect.getsource(fn).strip().split("K.")[1].split("(")[0]
            except Exception: src = "other"
            name = src
        elif "ConvPlan.run" in name:
            name = "conv_" + {0: "fprop", 1: "dgrad", 2: "wgrad"}[fn.__self__.mode]
        if rep == 2:
            agg[name][0] += 1; agg[name][1] += a.elapsed_time(b) * 1e3
tot = sum(v[1] for v in agg.values())
print(f"{which} batch {B}: sum of per-op times {tot/1e3:.2f} ms (eager, serial)")
for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{t:9.1f} us {100*t/tot:5.1f}% {c:4d}x avg {t/c:7.1f}  {k}")

#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_conv.py tests/test_gpu_engine.py tests/test_zz_no_alloc.py -q -x --timeout=300 -p no:cacheprovider 2>&1 | tail -3
for v in "LDNN_X=1" "LDNN_NO_AUTOTUNE=1"; do
  echo "== $v"; env $v timeout 300 python bench.py --steps 200 --warmup 20 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print(round(d['value']), 'samples/s', round(d['ms_per_step'], 4), 'ms/step  e2e', round(d['e2e']['value']), 'launches', d['gpu_launches'])"
done
python - <<'PY'
import sys, os
sys.path.insert(0, os.getcwd())
import torch, time
from ldnn_b200.engine import OptimConfig
from ldnn_b200.models import enhanced_cnn_spec
from ldnn_b200.models.native import NativeEngine
from ldnn_b200.ops import native_ops as K
t0 = time.time()
eng = NativeEngine(enhanced_cnn_spec(), "cuda:0", OptimConfig(), batch_size=64)
x = torch.randn(64, 3, 32, 32, device="cuda"); y = torch.randint(0, 10, (64,), device="cuda")
eng.train_step(x, y); torch.cuda.synchronize()
print("engine build + first step", round(time.time() - t0, 2), "s")
names = {0: "fprop", 1: "dgrad", 2: "wgrad"}
for k, v in K._TUNE_CACHE.items():
    if v != (0, 0): print(names[k[0]], k[1:], "->", v)
PY

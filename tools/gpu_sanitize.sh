#!/bin/bash
# full GPU test-suite, then compute-sanitizer (ONE tool per call) on a small conv+bn+head step
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 1500 python -m pytest tests -m gpu -q --timeout=240 -p no:cacheprovider > gpurun_out/pytest_gpu_all.log 2>&1; echo "pytest gpu all exit $?" >> gpurun_out/summary.txt
cat > /tmp/san_step.py <<'PY'
import os, sys
os.environ["LDNN_NO_GRAPH"] = "1"
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
import torch
from ldnn_b200.engine import OptimConfig
from ldnn_b200.models.native import NativeEngine
from test_flat_optimizer import tiny_spec
from ldnn_b200.models.spec import ConvBN, Head, NetSpec
s = NetSpec("mini", 3, 8)
s.units.append(ConvBN("prep.0", "prep.1", 3, 64, 3, 1, 1, "x", "a0"))
s.units.append(ConvBN("layer1.0.conv1", "layer1.0.bn1", 64, 128, 3, 2, 1, "a0", "a1"))
s.units.append(ConvBN("layer1.0.shortcut.0", "layer1.0.shortcut.1", 64, 128, 1, 2, 0, "a0", "s1", relu=False))
s.units.append(ConvBN("layer1.0.conv2", "layer1.0.bn2", 128, 128, 3, 1, 1, "a1", "a2", residual="s1"))
s.units.append(Head("fc", 128, 10, "a2"))
eng = NativeEngine(s, "cuda:0", OptimConfig(), batch_size=8)
x = torch.randn(8, 3, 8, 8, device="cuda"); y = torch.randint(0, 10, (8,), device="cuda")
for _ in range(2):
    loss, _ = eng.train_step(x, y)
torch.cuda.synchronize(); print("loss", float(loss))
PY
python /tmp/san_step.py > gpurun_out/san_plain.log 2>&1 && \
timeout 900 compute-sanitizer --tool ${1:-memcheck} --log-file gpurun_out/sanitizer_${1:-memcheck}.log python /tmp/san_step.py > gpurun_out/san_run.log 2>&1
echo "sanitizer ${1:-memcheck} exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -3 gpurun_out/pytest_gpu_all.log; tail -5 gpurun_out/sanitizer_${1:-memcheck}.log

#!/bin/bash
# First-contact GPU run: diagnostics, then each GPU test file in its own process.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python tools/gpu_diag.py > gpurun_out/diag.log 2>&1
for f in tests/test_gpu_conv.py tests/test_gpu_bn_head_optim.py tests/test_gpu_comm_emulated.py tests/test_gpu_engine.py; do
  timeout 900 python -m pytest $f -q --timeout=240 -p no:cacheprovider > gpurun_out/$(basename $f .py).log 2>&1
  echo "$f exit $?" >> gpurun_out/summary.txt
done
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -5 gpurun_out/diag.log

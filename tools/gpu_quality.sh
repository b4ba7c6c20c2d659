#!/bin/bash
mkdir -p gpurun_out
for s in 0.06 0.08; do
LDNN_QUALITY_SIGNAL=$s LDNN_QUALITY_OUT=gpurun_out/training_quality_$s.json timeout 600 python -m pytest tests/test_gpu_training_quality.py -q -s 2>&1 | grep -E "training quality|passed|failed|Error|assert" | head -8
done

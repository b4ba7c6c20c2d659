#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 600 python -m pytest tests/test_zz_no_alloc.py -q --timeout=120 -p no:cacheprovider > gpurun_out/noalloc_alone.log 2>&1; echo "noalloc alone exit $?" >> gpurun_out/summary.txt
timeout 1500 python -m pytest tests -m gpu -x -q --timeout=240 -p no:cacheprovider > gpurun_out/pytest_gpu_all.log 2>&1; echo "pytest gpu all exit $?" >> gpurun_out/summary.txt
timeout 300 python tools/grad_cos.py > gpurun_out/grad_cos.log 2>&1
cat gpurun_out/summary.txt; tail -3 gpurun_out/noalloc_alone.log; tail -6 gpurun_out/pytest_gpu_all.log

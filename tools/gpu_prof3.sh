#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 900 python -m pytest tests/test_gpu_conv.py tests/test_gpu_bn_head_optim.py tests/test_gpu_engine.py tests/test_zz_no_alloc.py -q -x --timeout=240 -p no:cacheprovider > gpurun_out/pytest_sub.log 2>&1; echo "pytest exit $?" >> gpurun_out/summary.txt
timeout 300 python tools/microbench.py ${1:-all} > gpurun_out/microbench.log 2>&1
timeout 600 python bench.py --steps 100 --warmup 10 > gpurun_out/bench_ours.log 2>&1; echo "bench exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -4 gpurun_out/pytest_sub.log; cat gpurun_out/microbench.log; tail -1 gpurun_out/bench_ours.log | cut -c1-330

#!/bin/bash
# 1 GPU: engine parity with yardsticks, the reworked bench (both arms).
mkdir -p gpurun_out; S=gpurun_out/r2_third_summary.txt; rm -f $S
export LDNN_TUNE_CACHE=$PWD/gpurun_out/tune_r2.json
timeout 1500 python -m pytest tests/test_gpu_engine.py -q -s > gpurun_out/t_engine.log 2>&1; echo "engine exit $?" >> $S
timeout 900 python bench.py > gpurun_out/bench_ours.log 2>&1; echo "bench ours exit $?" >> $S
timeout 600 python bench.py --impl reference > gpurun_out/bench_ref.log 2>&1; echo "bench ref exit $?" >> $S
cat $S; grep "grad parity" gpurun_out/t_engine.log; tail -4 gpurun_out/t_engine.log; tail -c 3000 gpurun_out/bench_ours.log; echo; tail -c 1500 gpurun_out/bench_ref.log

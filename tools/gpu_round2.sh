#!/bin/bash
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
for f in tests/test_gpu_comm_emulated.py tests/test_gpu_engine.py; do
  timeout 900 python -m pytest $f -q --timeout=240 -p no:cacheprovider > gpurun_out/$(basename $f .py).log 2>&1
  echo "$f exit $?" >> gpurun_out/summary.txt
done
timeout 600 python bench.py --steps 50 --warmup 5 > gpurun_out/bench_ours.log 2>&1; echo "bench ours exit $?" >> gpurun_out/summary.txt
timeout 600 python bench.py --impl reference --steps 50 --warmup 5 > gpurun_out/bench_ref.log 2>&1; echo "bench ref exit $?" >> gpurun_out/summary.txt
timeout 600 python bench.py --steps 50 --warmup 5 --sync step > gpurun_out/bench_ours_step.log 2>&1; echo "bench ours step exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -2 gpurun_out/bench_ours.log; tail -2 gpurun_out/bench_ref.log; tail -2 gpurun_out/bench_ours_step.log

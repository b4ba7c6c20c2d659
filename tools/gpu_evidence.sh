#!/bin/bash
# Evidence run: launch list of one step, ncu --set full of the conv kernels, SASS check, R50 bench.
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
python tools/profile_step.py 3 > gpurun_out/prof_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv python tools/profile_step.py 3 > gpurun_out/prof_ncu.log 2>&1
echo "launch list exit $?" >> gpurun_out/summary.txt
python tools/profile_step.py 2 > gpurun_out/prof_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:conv_gemm_kernel -s 70 -c 12 -o gpurun_out/conv_full -f python tools/profile_step.py 2 > gpurun_out/prof_ncu_full.log 2>&1
echo "ncu full exit $?" >> gpurun_out/summary.txt
timeout 900 python bench.py --model resnet50 --batch 32 --steps 20 --warmup 3 > gpurun_out/bench_r50.log 2>&1; echo "bench r50 exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -3 gpurun_out/bench_r50.log | cut -c1-400; ls -la gpurun_out | head -20

#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
python tools/profile_misc.py > gpurun_out/misc_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"bn_apply_sums|bn_bwd_reduce_atomic|bn_bwd_apply_sums|head_sample|head_reduce|optim_kernel|im2col_stem|allreduce_fused|ring_fused" -s 40 -c 14 -o gpurun_out/misc_full -f python tools/profile_misc.py > gpurun_out/misc_ncu.log 2>&1
echo "ncu misc exit $?" >> gpurun_out/summary.txt
python tools/profile_step.py 3 > gpurun_out/prof_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv python tools/profile_step.py 3 > gpurun_out/prof_ncu.log 2>&1
echo "launch list exit $?" >> gpurun_out/summary.txt
timeout 300 python tools/microbench.py all > gpurun_out/microbench.log 2>&1
cat gpurun_out/summary.txt; tail -3 gpurun_out/misc_ncu.log; ls -la gpurun_out/*.ncu-rep

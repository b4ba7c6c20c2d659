#!/bin/bash
# 1 GPU, final: whole GPU suite, smoke, bench both arms, then per-kernel ncu sections of one tf32 / bf16 step exported to
# CSV on the box (the .ncu-rep files are too large to bring back).
mkdir -p gpurun_out; S=gpurun_out/r2_ev3_summary.txt; rm -f $S
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest -m gpu exit $?" >> $S
timeout 600 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> $S
timeout 600 python bench.py --impl reference > gpurun_out/bench_ref_1.log 2>&1; echo "bench ref exit $?" >> $S
timeout 600 python bench.py > gpurun_out/bench_ours_1.log 2>&1; echo "bench ours exit $?" >> $S
K='regex:conv_gemm|bn_|head_|optim_kernel|im2col'
SEC="--section SpeedOfLight --section MemoryWorkloadAnalysis --section ComputeWorkloadAnalysis --section LaunchStats --section Occupancy --metrics sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_tensor.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__warps_active.avg.pct_of_peak_sustained_active"
for dt in tf32 bf16; do
python tools/profile_step.py 3 $dt > gpurun_out/prof_plain_sec_$dt.log 2>&1 && \
timeout 900 ncu $SEC --clock-control none -k "$K" -s 256 -c 128 -f -o /tmp/step_$dt python tools/profile_step.py 3 $dt > gpurun_out/ncu_sections_$dt.log 2>&1
echo "ncu sections $dt exit $?" >> $S
python tools/ncu_summary.py /tmp/step_$dt.ncu-rep > gpurun_out/ncu_key_metrics_$dt.csv 2>> $S
done
cat $S; tail -2 gpurun_out/pytest_gpu.log; tail -1 gpurun_out/smoke.log; wc -l gpurun_out/ncu_key_metrics_*.csv; head -4 gpurun_out/ncu_key_metrics_tf32.csv | cut -c1-400
tail -c 700 gpurun_out/bench_ours_1.log

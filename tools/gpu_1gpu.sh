#!/bin/bash
# One B200: the whole GPU test suite, smoke, bench (both arms), microbench with clock record, entrypoint bench, launch
# lists of one tf32 / bf16 training step, per-kernel ncu sections exported to CSV ON THE BOX (a report of 128 kernels is
# ~150 MB; gpurun brings back at most 64 MiB), and one `--set full --import-source on` capture of the conv kernels.
# Every ncu command runs only after its plain twin exited 0.  Outputs land in gpurun_out/; copy what is to be kept
# into profiles/.
mkdir -p gpurun_out; S=gpurun_out/summary_1gpu.txt; rm -f $S
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest -m gpu exit $?" >> $S
timeout 600 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> $S
timeout 900 python bench.py --impl reference > gpurun_out/bench_ref_1.log 2>&1; echo "bench ref exit $?" >> $S
timeout 900 python bench.py > gpurun_out/bench_ours_1.log 2>&1; echo "bench ours exit $?" >> $S
timeout 900 python tools/microbench.py > gpurun_out/microbench_kernels_r2.txt 2>&1; echo "microbench exit $?" >> $S
timeout 1200 python bench/entrypoint_bench.py --gpus 1 > gpurun_out/entrypoints_1.txt 2>&1; echo "entrypoints exit $?" >> $S
K='regex:conv_gemm|bn_|head_|optim_kernel|im2col'
SEC="--section SpeedOfLight --section MemoryWorkloadAnalysis --section ComputeWorkloadAnalysis --section LaunchStats --section Occupancy --metrics sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_tensor.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__warps_active.avg.pct_of_peak_sustained_active"
for dt in tf32 bf16; do
python tools/profile_step.py 3 $dt > gpurun_out/prof_plain_$dt.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_$dt.csv \
    python tools/profile_step.py 3 $dt > gpurun_out/prof_ncu_$dt.log 2>&1
echo "launch list $dt exit $?" >> $S
python tools/agg_launches.py gpurun_out/launches_$dt.csv > gpurun_out/launch_list_step_$dt.txt 2>&1
python tools/profile_step.py 3 $dt > gpurun_out/prof_plain_sec_$dt.log 2>&1 && \
timeout 900 ncu $SEC --clock-control none -k "$K" -s 256 -c 128 -f -o /tmp/step_$dt python tools/profile_step.py 3 $dt > gpurun_out/ncu_sections_$dt.log 2>&1
echo "ncu sections $dt exit $?" >> $S
python tools/ncu_summary.py /tmp/step_$dt.ncu-rep > gpurun_out/ncu_key_metrics_$dt.csv 2>> $S
done
python tools/profile_step.py 3 tf32 > gpurun_out/prof_plain_src.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:conv_gemm_kernel -s 120 -c 6 -f -o gpurun_out/conv_src_tf32 python tools/profile_step.py 3 tf32 > gpurun_out/ncu_src_tf32.log 2>&1
echo "ncu full + source (6 conv launches) exit $?" >> $S
cat $S; tail -2 gpurun_out/pytest_gpu.log; tail -1 gpurun_out/smoke.log; tail -c 900 gpurun_out/bench_ours_1.log; echo; tail -9 gpurun_out/entrypoints_1.txt

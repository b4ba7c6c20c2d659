#!/bin/bash
# 1 GPU: the whole GPU test suite, bench (both arms), entrypoint bench, microbench with clock record, launch lists and
# ncu --set full of every kernel of one tf32 / bf16 training step (each ncu command only after its plain twin exited 0).
mkdir -p gpurun_out; S=gpurun_out/r2_ev1_summary.txt; rm -f $S
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest -m gpu exit $?" >> $S
timeout 900 python bench.py --impl reference > gpurun_out/bench_ref_1.log 2>&1; echo "bench ref exit $?" >> $S
timeout 900 python bench.py > gpurun_out/bench_ours_1.log 2>&1; echo "bench ours exit $?" >> $S
timeout 900 python tools/microbench.py > gpurun_out/microbench_kernels_r2.txt 2>&1; echo "microbench exit $?" >> $S
timeout 1200 python bench/entrypoint_bench.py --gpus 1 > gpurun_out/entrypoints_1.txt 2>&1; echo "entrypoints exit $?" >> $S
for dt in tf32 bf16; do
python tools/profile_step.py 3 $dt > gpurun_out/prof_plain_$dt.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_$dt.csv \
    python tools/profile_step.py 3 $dt > gpurun_out/prof_ncu_$dt.log 2>&1
echo "launch list $dt exit $?" >> $S
python tools/agg_launches.py gpurun_out/launches_$dt.csv > gpurun_out/launch_list_step_$dt.txt 2>&1
done
python tools/profile_step.py 3 tf32 > gpurun_out/prof_plain2.log 2>&1 && \
ncu --set full --clock-control none -k regex:ldnn -s 266 -c 133 -f -o gpurun_out/step_tf32 python tools/profile_step.py 3 tf32 > gpurun_out/ncu_full_tf32.log 2>&1
echo "ncu full tf32 exit $?" >> $S
python tools/profile_step.py 3 bf16 > gpurun_out/prof_plain3.log 2>&1 && \
ncu --set full --clock-control none -k regex:ldnn -s 266 -c 133 -f -o gpurun_out/step_bf16 python tools/profile_step.py 3 bf16 > gpurun_out/ncu_full_bf16.log 2>&1
echo "ncu full bf16 exit $?" >> $S
python tools/profile_step.py 3 tf32 > gpurun_out/prof_plain4.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:conv_gemm_kernel -s 120 -c 6 -f -o gpurun_out/conv_src_tf32 python tools/profile_step.py 3 tf32 > gpurun_out/ncu_src_tf32.log 2>&1
echo "ncu source tf32 exit $?" >> $S
ls -la gpurun_out/*.ncu-rep >> $S 2>&1
cat $S; tail -3 gpurun_out/pytest_gpu.log; tail -c 1200 gpurun_out/bench_ours_1.log; echo; tail -14 gpurun_out/entrypoints_1.txt; tail -45 gpurun_out/microbench_kernels_r2.txt

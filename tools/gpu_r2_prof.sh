#!/bin/bash
# 1 GPU: launch lists (ncu, per-kernel device time) of one tf32 and one bf16 training step, shipped tunings replayed.
mkdir -p gpurun_out; S=gpurun_out/r2_prof_summary.txt; rm -f $S
timeout 600 python -m pytest tests/test_gpu_comm_emulated.py -q -x > gpurun_out/t_comm_emu.log 2>&1; echo "emulated comm exit $?" >> $S
for dt in tf32 bf16; do
python tools/profile_step.py 3 $dt > gpurun_out/prof_plain_$dt.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_$dt.csv \
    python tools/profile_step.py 3 $dt > gpurun_out/prof_ncu_$dt.log 2>&1
echo "launch list $dt exit $?" >> $S
python tools/agg_launches.py gpurun_out/launches_$dt.csv > gpurun_out/launch_list_step_$dt.txt 2>&1
done
cat $S; tail -2 gpurun_out/t_comm_emu.log; cat gpurun_out/launch_list_step_tf32.txt; cat gpurun_out/launch_list_step_bf16.txt

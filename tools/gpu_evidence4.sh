#!/bin/bash
# Launch list of one training step WITH the autotuner's per-shape choices: a plain run tunes and writes
# LDNN_TUNE_CACHE; the ncu run (where event timing would mis-tune) replays that file with timing disabled.
mkdir -p gpurun_out; rm -f gpurun_out/summary4.txt
export LDNN_TUNE_CACHE=$PWD/gpurun_out/tune.json
LDNN_NO_AUTOTUNE= python tools/profile_step.py 3 > gpurun_out/prof4_plain.log 2>&1 && \
LDNN_NO_AUTOTUNE=1 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_tuned.csv \
    python tools/profile_step.py 3 > gpurun_out/prof4_ncu.log 2>&1
echo "launch list exit $?" >> gpurun_out/summary4.txt
python tools/agg_launches.py gpurun_out/launches_tuned.csv > gpurun_out/launch_list_step_tuned.txt 2>&1
unset LDNN_TUNE_CACHE
python bench.py > gpurun_out/bench_final_1gpu.log 2>&1; echo "bench exit $?" >> gpurun_out/summary4.txt
tail -1 gpurun_out/bench_final_1gpu.log; cat gpurun_out/summary4.txt; cat gpurun_out/launch_list_step_tuned.txt; tail -2 gpurun_out/prof4_plain.log

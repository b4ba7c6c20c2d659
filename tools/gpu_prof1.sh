#!/bin/bash
mkdir -p gpurun_out
python tools/profile_step.py 3 > gpurun_out/prof_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv python tools/profile_step.py 3 > gpurun_out/prof_ncu.log 2>&1
echo "exit $?"; tail -2 gpurun_out/prof_plain.log; wc -l gpurun_out/launches.csv

#!/bin/bash
# BASELINE config 3 at 4 GPUs: uneven shards 4:2:1:1, shard-weighted exchange (n_r / sum n).
mkdir -p gpurun_out
R="--shard_ratios 4 2 1 1 --aggregation_type shard --steps 100 --warmup 5"
timeout 32 python bench.py --gpus 4 --topology double_ring $R > gpurun_out/disb4_dring_epoch.json 2> gpurun_out/disb4_dring_epoch.err; echo "dring/global_epoch exit $?"
timeout 32 python bench.py --gpus 4 --topology ring --sync step $R > gpurun_out/disb4_ring_step.json 2> gpurun_out/disb4_ring_step.err; echo "ring/step exit $?"
tail -1 gpurun_out/disb4_dring_epoch.json; tail -1 gpurun_out/disb4_ring_step.json; tail -3 gpurun_out/disb4_ring_step.err

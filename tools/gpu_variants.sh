#!/bin/bash
# Runs every variant entrypoint on N GPUs (synthetic data) and collects its graphs.
N=${1:-2}
mkdir -p gpurun_out/graphs; rm -f gpurun_out/summary.txt
COMMON="--synthetic --backend nccl --epochs_global 4 --epochs_local 2 --synthetic_train 8192 --synthetic_val 2048 --synthetic_test 2048 --seed 1 --quiet"
i=0
for v in "Balanced All-Reduce" "Balanced Ring" "Balanced Double-Ring" "Disbalanced All-Reduce" "Disbalanced Ring" "Disalanced Double-Ring"; do
  i=$((i+1))
  extra="--aggregation_by weights"
  case "$v" in Disb*|Disa*) extra="$extra --aggregation_type weighted --local_weight 0.6";; esac
  out="gpurun_out/graphs/$v"
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29560+i)) "$v/main.py" $COMMON $extra --output_folder "$out" > "gpurun_out/variant_$i.log" 2>&1
  echo "$v exit $?" >> gpurun_out/summary.txt
done
# product schedule through the CLI: per-step fused ring exchange, shard-weighted, uneven shards, checkpoint
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29570 "Disbalanced Ring/main.py" $COMMON --sync step --aggregation_type shard --shard_ratios 3 1 --save /tmp/ckpt_step.pt --output_folder "gpurun_out/graphs/Disbalanced Ring (sync step)" > gpurun_out/variant_step.log 2>&1
echo "product-schedule run exit $?" >> gpurun_out/summary.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29571 tests/mgpu_worker.py > gpurun_out/mgpu_worker_$N.log 2>&1; echo "mgpu worker exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -3 gpurun_out/variant_1.log; tail -4 gpurun_out/variant_step.log; ls -la /tmp/ckpt_step.pt; du -sh gpurun_out

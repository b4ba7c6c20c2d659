#!/bin/bash
mkdir -p gpurun_out
for v in "Balanced Ring" "Balanced All-Reduce" "Balanced Ring"; do
python "$v/main.py" --synthetic --synthetic_train 10240 --synthetic_val 2560 --synthetic_test 640 --epochs_global 4 --epochs_local 2 --batch_size 64 --quiet --seed 0 --backend nccl --output_folder gpurun_out/ep --trace gpurun_out/trace > gpurun_out/ringslow.log 2>&1
python - <<PY
import json
m = json.load(open('gpurun_out/ep/metrics.json'))
print("$v", "epoch s", [round(x, 3) for x in m['global_epoch_seconds']], "exchange ms", [round(x, 1) for x in m['exchange_ms']])
t = json.load(open('gpurun_out/trace.rank0.json'))['traceEvents']
print("   phases ms:", [(e['name'], round(e['dur'] / 1e3, 1)) for e in t])
PY
done

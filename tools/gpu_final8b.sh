#!/bin/bash
N=${1:-8}
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
run() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
run 29544 bench.py --gpus $N --steps 100 --warmup 10 > gpurun_out/bench_allreduce_epoch_$N.log 2>&1; echo "bench epoch exit $?" >> gpurun_out/summary.txt
for topo in allreduce ring double_ring; do
run 29542 bench.py --gpus $N --steps 100 --warmup 10 --topology $topo --sync step > gpurun_out/bench_${topo}_step_overlap_$N.log 2>&1; echo "bench $topo step overlap exit $?" >> gpurun_out/summary.txt
done
run 29547 bench.py --gpus $N --steps 20 --warmup 3 --model resnet50 --batch 32 --topology double_ring --sync step > gpurun_out/bench_r50_dring_step_$N.log 2>&1; echo "bench r50 exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; grep -h '"metric"' gpurun_out/bench_*_$N.log | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); c = d['config']; print(d['impl'], c.get('model'), c.get('topology'), c.get('sync'), 'overlap' if c.get('overlap') else '', round(d['value']), 'samples/s', round(d['ms_per_step'],3), 'ms/step e2e', round(d['e2e']['value']), d['clocks']['reasons'], 'launches', d['gpu_launches'])
"

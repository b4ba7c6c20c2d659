#!/bin/bash
# Final multi-GPU validation on N GPUs: IPC worker test, headline bench (both arms), per-step product
# schedules for the three topologies, ResNet-50 double-ring, exchange sweep.
N=${1:-8}
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
run() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
run 29541 tests/mgpu_worker.py > gpurun_out/mgpu_worker_$N.log 2>&1; echo "mgpu worker exit $?" >> gpurun_out/summary.txt
run 29544 bench.py --gpus $N --steps 100 --warmup 10 > gpurun_out/bench_allreduce_epoch_$N.log 2>&1; echo "bench epoch exit $?" >> gpurun_out/summary.txt
run 29546 bench.py --gpus $N --steps 100 --warmup 10 --impl reference > gpurun_out/bench_ref_$N.log 2>&1; echo "bench ref exit $?" >> gpurun_out/summary.txt
for topo in allreduce ring double_ring; do
run 29542 bench.py --gpus $N --steps 100 --warmup 10 --topology $topo --sync step > gpurun_out/bench_${topo}_step_overlap_$N.log 2>&1; echo "bench $topo step overlap exit $?" >> gpurun_out/summary.txt
done
run 29547 bench.py --gpus $N --steps 20 --warmup 3 --model resnet50 --batch 32 --topology double_ring --sync step > gpurun_out/bench_r50_dring_step_$N.log 2>&1; echo "bench r50 exit $?" >> gpurun_out/summary.txt
run 29545 bench/exchange_sweep.py --out gpurun_out/exchange_sweep_$N.json > gpurun_out/sweep_$N.log 2>&1; echo "sweep exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -1 gpurun_out/mgpu_worker_$N.log | cut -c1-200; grep -h '"metric"' gpurun_out/bench_*_$N.log | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); c = d['config']; print(d['impl'], c.get('model'), c.get('topology'), c.get('sync'), 'overlap' if c.get('overlap') else '', round(d['value']), 'samples/s', round(d['ms_per_step'],3), 'ms/step e2e', round(d['e2e']['value']), d['clocks'])
"; tail -22 gpurun_out/sweep_$N.log

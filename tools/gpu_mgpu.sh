#!/bin/bash
# usage: gpu_mgpu.sh N   — multi-GPU checks on an N-GPU box
N=${1:-2}
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
nvidia-smi topo -m > gpurun_out/topo_$N.txt 2>&1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 tests/mgpu_worker.py > gpurun_out/mgpu_worker_$N.log 2>&1; echo "mgpu worker exit $?" >> gpurun_out/summary.txt
for topo in allreduce ring double_ring; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus $N --steps 50 --warmup 5 --topology $topo > gpurun_out/bench_ours_${topo}_$N.log 2>&1; echo "bench $topo exit $?" >> gpurun_out/summary.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29543 bench.py --gpus $N --steps 50 --warmup 5 --topology $topo --sync step > gpurun_out/bench_ours_${topo}_step_$N.log 2>&1; echo "bench $topo step exit $?" >> gpurun_out/summary.txt
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus $N --steps 50 --warmup 5 --impl reference > gpurun_out/bench_ref_$N.log 2>&1; echo "bench ref exit $?" >> gpurun_out/summary.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29545 bench/exchange_sweep.py --out gpurun_out/exchange_sweep_$N.json > gpurun_out/sweep_$N.log 2>&1; echo "sweep exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -3 gpurun_out/mgpu_worker_$N.log; grep -h '"metric"' gpurun_out/bench_*_$N.log | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print(d['impl'], d['config'].get('topology'), d['config'].get('sync'), round(d['value']), 'samples/s', round(d['ms_per_step'],3), 'ms/step e2e', round(d['e2e']['value']))
"; tail -24 gpurun_out/sweep_$N.log

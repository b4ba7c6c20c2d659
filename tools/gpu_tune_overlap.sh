#!/bin/bash
N=${1:-4}
mkdir -p gpurun_out
run() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 "${@:2}" 2>/dev/null | grep '"metric"' | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print(round(d['value']), 'samples/s', round(d['ms_per_step'], 3), 'ms/step')"; }
echo "== epoch schedule"; run 29601 bench.py --gpus $N --steps 60 --warmup 5
for topo in allreduce ring; do
for cb in 32 64 128; do for mb in 16 32 64; do
  echo "== $topo comm_blocks=$cb bucket_mb=$mb"; LDNN_BUCKET_MB=$mb run 29602 bench.py --gpus $N --steps 60 --warmup 5 --topology $topo --sync step --comm_blocks $cb
done; done; done

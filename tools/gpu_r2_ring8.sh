#!/bin/bash
mkdir -p gpurun_out
run() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
LDNN_RING_LL_MAX_MB=0 run 29545 bench/exchange_sweep.py --max_mb 512 --min_kb 65536 --impls ring,double_ring,nccl --modes reduce --iters 8 --out gpurun_out/ring_chains1_8.json > gpurun_out/ring_chains1_8gpu.txt 2>&1; echo "exit $?"
grep -v "^\*\|OMP_NUM\|Warning\|warn" gpurun_out/ring_chains1_8gpu.txt | tail -8

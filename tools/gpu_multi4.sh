#!/bin/bash
# 4 GPUs: exchange kernels on real peers, full sweep with clock record, ring chain-count comparison, bench both arms, R50.
N=4
mkdir -p gpurun_out; S=gpurun_out/multi_${N}_summary.txt; rm -f $S
run() { timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
run 29542 tests/mgpu_worker.py > gpurun_out/mgpu_worker_default_$N.log 2>&1; echo "mgpu worker (defaults) exit $?" >> $S
run 29544 bench/exchange_sweep.py --max_mb 512 --out gpurun_out/exchange_sweep_$N.json > gpurun_out/exchange_sweep_${N}gpu.txt 2>&1; echo "sweep exit $?" >> $S
for ch in 1 2 4; do
LDNN_RING_LL_MAX_MB=0 LDNN_RING_CHAINS=$ch run 29545 bench/exchange_sweep.py --max_mb 512 --min_kb 16384 --impls ring,double_ring --modes reduce --iters 8 --out gpurun_out/ring_chains${ch}_$N.json > gpurun_out/ring_chains${ch}_${N}gpu.txt 2>&1; echo "ring chains=$ch exit $?" >> $S
done
run 29546 bench.py --gpus $N --steps 60 --warmup 5 --impl reference > gpurun_out/bench_ref_$N.log 2>&1; echo "bench ref exit $?" >> $S
run 29547 bench.py --gpus $N --steps 60 --warmup 5 > gpurun_out/bench_ours_$N.log 2>&1; echo "bench ours exit $?" >> $S
run 29552 bench.py --gpus $N --steps 30 --warmup 5 --no_extra --model resnet50 --batch 32 --dtype bf16 --topology double_ring --sync step > gpurun_out/bench_r50_dring_$N.log 2>&1; echo "resnet50 double ring exit $?" >> $S
cat $S; tail -1 gpurun_out/mgpu_worker_default_$N.log | cut -c1-200
grep -v "^\*\|OMP_NUM\|Warning\|warn" gpurun_out/exchange_sweep_${N}gpu.txt | tail -48
for ch in 1 2 4; do echo "chains=$ch (bulk kernel forced)"; grep -v "^\*\|OMP_NUM\|Warning\|warn\|^#" gpurun_out/ring_chains${ch}_${N}gpu.txt | tail -6; done
for f in gpurun_out/bench_*_$N.log; do grep -h '"metric"' $f | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); c = d['config']; print('$f', d['impl'], c.get('model'), c.get('topology'), c.get('sync'), round(d['value']), 'samples/s', round(d['ms_per_step'],3), 'ms/step e2e', round(d['e2e']['value']))
    for k, v in (d.get('extra', {}).get('runs') or {}).items(): print('    ', k, {a: (round(b, 3) if isinstance(b, float) else b) for a, b in v.items()})
    if 'bf16' in d.get('extra', {}): print('     bf16', d['extra']['bf16'])
"; done

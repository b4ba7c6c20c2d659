#!/bin/bash
# usage: gpu_multi.sh N [max_mb] [full] — multi-GPU evidence of round 2 on N GPUs: exchange kernels on real peers (bulk
# rings, flag-free rings, two-shot, NVLS, compat functions), exchange sweep with clock record + NVLink byte counters,
# bench both arms (headline + extras), entrypoint bench; with `full`: ResNet-50 @224 double ring and Disbalanced runs.
N=${1:-2}; MAXMB=${2:-256}; FULL=${3:-}
mkdir -p gpurun_out; S=gpurun_out/multi_${N}_summary.txt; rm -f $S
run() { timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
LDNN_RING_LL_MAX_MB=0 LDNN_NVLS=0 run 29541 tests/mgpu_worker.py > gpurun_out/mgpu_worker_bulk_$N.log 2>&1; echo "mgpu worker (bulk rings, IPC slabs) exit $?" >> $S
run 29542 tests/mgpu_worker.py > gpurun_out/mgpu_worker_default_$N.log 2>&1; echo "mgpu worker (defaults: flag-free rings, NVLS) exit $?" >> $S
run 29544 bench/exchange_sweep.py --max_mb $MAXMB --out gpurun_out/exchange_sweep_$N.json > gpurun_out/exchange_sweep_${N}gpu.txt 2>&1; echo "sweep exit $?" >> $S
run 29546 bench.py --gpus $N --steps 60 --warmup 5 --impl reference > gpurun_out/bench_ref_$N.log 2>&1; echo "bench ref exit $?" >> $S
run 29547 bench.py --gpus $N --steps 60 --warmup 5 > gpurun_out/bench_ours_$N.log 2>&1; echo "bench ours exit $?" >> $S
timeout 1500 python bench/entrypoint_bench.py --gpus $N --variants "Balanced All-Reduce" "Balanced Ring" "Balanced Double-Ring" > gpurun_out/entrypoints_$N.txt 2>&1; echo "entrypoints exit $?" >> $S
if [ -n "$FULL" ]; then
R4=$(python -c "print(' '.join((['4','2','1','1']*8)[:$N]))")
run 29548 bench.py --gpus $N --steps 60 --warmup 5 --no_extra --topology ring --shard_ratios $R4 --aggregation_type shard > gpurun_out/bench_disb_ring_epoch_$N.log 2>&1; echo "disbalanced ring (reference schedule) exit $?" >> $S
run 29549 bench.py --gpus $N --steps 60 --warmup 5 --no_extra --topology double_ring --sync step --shard_ratios $R4 --aggregation_type shard > gpurun_out/bench_disb_dring_step_$N.log 2>&1; echo "disbalanced double ring (per step) exit $?" >> $S
run 29550 bench.py --gpus $N --steps 60 --warmup 5 --no_extra --topology allreduce --sync step --shard_ratios $R4 --aggregation_type shard > gpurun_out/bench_disb_ar_step_$N.log 2>&1; echo "disbalanced all-reduce (per step) exit $?" >> $S
run 29551 bench.py --gpus $N --steps 60 --warmup 5 --no_extra --impl reference --shard_ratios $R4 > gpurun_out/bench_disb_ref_$N.log 2>&1; echo "disbalanced reference exit $?" >> $S
fi
run 29552 bench.py --gpus $N --steps 30 --warmup 5 --no_extra --model resnet50 --batch 32 --dtype bf16 --topology double_ring --sync step > gpurun_out/bench_r50_dring_$N.log 2>&1; echo "resnet50 double ring exit $?" >> $S
cat $S
for f in bulk default; do tail -1 gpurun_out/mgpu_worker_${f}_$N.log | cut -c1-300; done
grep -v "^\*\|OMP_NUM\|Warning\|warn" gpurun_out/exchange_sweep_${N}gpu.txt | tail -60
for f in gpurun_out/bench_*_$N.log; do grep -h '"metric"' $f | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); c = d['config']; print('$f', d['impl'], c.get('model'), c.get('topology'), c.get('sync'), round(d['value']), 'samples/s', round(d['ms_per_step'],3), 'ms/step e2e', round(d['e2e']['value']))
    for k, v in (d.get('extra', {}).get('runs') or {}).items(): print('    ', k, {a: (round(b, 3) if isinstance(b, float) else b) for a, b in v.items()})
    if 'bf16' in d.get('extra', {}): print('     bf16', d['extra']['bf16'])
"; done
tail -6 gpurun_out/entrypoints_$N.txt

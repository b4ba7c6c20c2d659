#!/bin/bash
# usage: gpu_r2_mgpu.sh N [max_mb] — multi-GPU checks of round 2 on N GPUs: exchange kernels on real peers (bulk rings,
# flag-free rings, NVLS), exchange sweep with clock record, bench both arms.
N=${1:-2}; MAXMB=${2:-256}
mkdir -p gpurun_out; S=gpurun_out/r2_mgpu_${N}_summary.txt; rm -f $S
run() { timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
timeout 600 python -m pytest tests/test_gpu_comm_emulated.py -q -x > gpurun_out/t_comm_emu.log 2>&1; echo "emulated comm exit $?" >> $S
LDNN_RING_LL_MAX_MB=0 run 29541 tests/mgpu_worker.py > gpurun_out/mgpu_worker_bulk_$N.log 2>&1; echo "mgpu worker (bulk rings) exit $?" >> $S
run 29542 tests/mgpu_worker.py > gpurun_out/mgpu_worker_ll_$N.log 2>&1; echo "mgpu worker (flag-free rings) exit $?" >> $S
LDNN_NVLS=1 run 29543 tests/mgpu_worker.py > gpurun_out/mgpu_worker_nvls_$N.log 2>&1; echo "mgpu worker (NVLS) exit $?" >> $S
run 29544 bench/exchange_sweep.py --max_mb $MAXMB --out gpurun_out/exchange_sweep_$N.json > gpurun_out/sweep_$N.log 2>&1; echo "sweep exit $?" >> $S
LDNN_NVLS=1 run 29545 bench/exchange_sweep.py --max_mb $MAXMB --modes reduce --out gpurun_out/exchange_sweep_nvls_$N.json > gpurun_out/sweep_nvls_$N.log 2>&1; echo "sweep nvls exit $?" >> $S
run 29546 bench.py --gpus $N --steps 60 --warmup 5 > gpurun_out/bench_ours_$N.log 2>&1; echo "bench ours exit $?" >> $S
run 29547 bench.py --gpus $N --steps 60 --warmup 5 --impl reference > gpurun_out/bench_ref_$N.log 2>&1; echo "bench ref exit $?" >> $S
cat $S; tail -3 gpurun_out/t_comm_emu.log
for f in bulk ll nvls; do tail -2 gpurun_out/mgpu_worker_${f}_$N.log | cut -c1-400; done
grep -v "^\*\|OMP_NUM" gpurun_out/sweep_$N.log | tail -50; grep -v "^\*\|OMP_NUM" gpurun_out/sweep_nvls_$N.log | tail -26
tail -c 2500 gpurun_out/bench_ours_$N.log; echo; tail -c 1800 gpurun_out/bench_ref_$N.log

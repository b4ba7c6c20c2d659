#!/bin/bash
# Round 2, second GPU call (1 GPU): MN-major TF32 fix, engine parity, pair MMA in tf32, the delayed-CTA regression test.
mkdir -p gpurun_out; S=gpurun_out/r2_second_summary.txt; rm -f $S
export LDNN_TUNE_CACHE=$PWD/gpurun_out/tune_r2.json
timeout 300 python tools/tf32_probe.py > gpurun_out/tf32_probe2.txt 2>&1; echo "probe exit $?" >> $S
timeout 900 python -m pytest tests/test_gpu_conv.py -q -k "tf32" > gpurun_out/t_conv_tf32.log 2>&1; echo "conv tf32 exit $?" >> $S
LDNN_TEST_PAIR=1 timeout 300 python -m pytest tests/test_gpu_conv.py -q -k "cta_pair and tf32" > gpurun_out/t_pair_tf32.log 2>&1; echo "pair tf32 exit $?" >> $S
timeout 600 python -m pytest tests/test_gpu_bn_head_optim.py -x -q -k "sums" > gpurun_out/t_bn.log 2>&1; echo "bn sums exit $?" >> $S
timeout 1500 python -m pytest tests/test_gpu_engine.py -q -s > gpurun_out/t_engine.log 2>&1; echo "engine exit $?" >> $S
timeout 600 python -m pytest tests/test_gpu_comm_emulated.py tests/test_gpu_comm_nvls_emulated.py -q > gpurun_out/t_comm.log 2>&1; echo "comm exit $?" >> $S
timeout 900 python bench.py --dtype tf32 > gpurun_out/bench_tf32.log 2>&1; echo "bench tf32 exit $?" >> $S
timeout 600 python bench.py --impl reference > gpurun_out/bench_ref.log 2>&1; echo "bench ref exit $?" >> $S
cat $S; cat gpurun_out/tf32_probe2.txt; tail -5 gpurun_out/t_conv_tf32.log; tail -3 gpurun_out/t_bn.log; grep "grad parity" gpurun_out/t_engine.log; tail -8 gpurun_out/t_engine.log; tail -3 gpurun_out/t_pair_tf32.log; tail -4 gpurun_out/t_comm.log; tail -1 gpurun_out/bench_tf32.log; tail -1 gpurun_out/bench_ref.log

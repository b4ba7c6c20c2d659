#!/bin/bash
# Round 2, first GPU call (1 GPU): TF32 path numerics + the tests of everything templated, the staged pair-MMA
# tests in their own process (a trap there must not take the rest down), a first tf32 / bf16 bench line.
mkdir -p gpurun_out; S=gpurun_out/r2_first_summary.txt; rm -f $S
export LDNN_TUNE_CACHE=$PWD/gpurun_out/tune_r2.json
for t in tfloat32 float32; do LDNN_TF32_TMA=$t timeout 300 python tools/tf32_probe.py >> gpurun_out/tf32_probe.txt 2>&1; echo "probe $t exit $?" >> $S; done
timeout 900 python -m pytest tests/test_gpu_conv.py -x -q -k "tf32" > gpurun_out/t_conv_tf32.log 2>&1; echo "conv tf32 exit $?" >> $S
timeout 900 python -m pytest tests/test_gpu_conv.py -x -q -k "bf16" > gpurun_out/t_conv_bf16.log 2>&1; echo "conv bf16 exit $?" >> $S
timeout 600 python -m pytest tests/test_gpu_bn_head_optim.py -x -q > gpurun_out/t_bn.log 2>&1; echo "bn/head exit $?" >> $S
timeout 1500 python -m pytest tests/test_gpu_engine.py -q -s > gpurun_out/t_engine.log 2>&1; echo "engine exit $?" >> $S
LDNN_TEST_PAIR=1 timeout 300 python -m pytest tests/test_gpu_conv.py -q -k "cta_pair and bf16" > gpurun_out/t_pair_bf16.log 2>&1; echo "pair bf16 exit $?" >> $S
LDNN_TEST_PAIR=1 timeout 300 python -m pytest tests/test_gpu_conv.py -q -k "cta_pair and tf32" > gpurun_out/t_pair_tf32.log 2>&1; echo "pair tf32 exit $?" >> $S
timeout 900 python bench.py --dtype tf32 > gpurun_out/bench_tf32.log 2>&1; echo "bench tf32 exit $?" >> $S
timeout 900 python bench.py --dtype bf16 > gpurun_out/bench_bf16.log 2>&1; echo "bench bf16 exit $?" >> $S
timeout 600 python -m pytest tests/test_zz_no_alloc.py tests/test_gpu_comm_emulated.py tests/test_gpu_comm_nvls_emulated.py -q > gpurun_out/t_rest.log 2>&1; echo "rest exit $?" >> $S
cat $S; cat gpurun_out/tf32_probe.txt; tail -3 gpurun_out/t_conv_tf32.log; tail -3 gpurun_out/t_bn.log; tail -15 gpurun_out/t_engine.log; tail -3 gpurun_out/t_pair_bf16.log; tail -3 gpurun_out/t_pair_tf32.log; tail -1 gpurun_out/bench_tf32.log; tail -1 gpurun_out/bench_bf16.log

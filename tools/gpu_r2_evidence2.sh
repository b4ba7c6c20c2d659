#!/bin/bash
# 1 GPU: emulated exchange tests after the ring change, then ncu --set full of every kernel of one tf32 / bf16 step.
mkdir -p gpurun_out; S=gpurun_out/r2_ev2_summary.txt; rm -f $S
timeout 600 python -m pytest tests/test_gpu_comm_emulated.py tests/test_gpu_comm_nvls_emulated.py -q -x > gpurun_out/t_comm_emu.log 2>&1; echo "emulated comm exit $?" >> $S
K='regex:conv_gemm|bn_|head_|optim_kernel|im2col'
for dt in tf32 bf16; do
python tools/profile_step.py 3 $dt > gpurun_out/prof_plain_full_$dt.log 2>&1 && \
ncu --set full --clock-control none -k "$K" -s 256 -c 128 -f -o gpurun_out/step_$dt python tools/profile_step.py 3 $dt > gpurun_out/ncu_full_$dt.log 2>&1
echo "ncu full $dt exit $?" >> $S
done
ls -la gpurun_out/*.ncu-rep >> $S 2>&1
cat $S; tail -2 gpurun_out/t_comm_emu.log; tail -4 gpurun_out/ncu_full_tf32.log

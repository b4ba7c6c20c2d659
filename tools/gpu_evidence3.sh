#!/bin/bash
# Final evidence: launch list (heuristic tiling, eager), ncu --set full per kernel family.
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
export LDNN_NO_AUTOTUNE=1
python tools/profile_step.py 3 > gpurun_out/prof_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv python tools/profile_step.py 3 > gpurun_out/prof_ncu.log 2>&1
echo "launch list exit $?" >> gpurun_out/summary.txt
python tools/profile_misc.py > gpurun_out/misc_plain.log 2>&1 && {
for fam in "conv_gemm_kernel" "bn_apply_sums|bn_bwd" "head_sample|head_reduce|im2col_stem|optim_kernel" "allreduce_fused|ring_fused"; do
  tag=$(echo "$fam" | tr -c 'a-z_' '_' | cut -c1-24)
  ncu --set full --clock-control none --import-source on -k regex:"$fam" -s 2 -c 4 -o gpurun_out/full_$tag -f python tools/profile_misc.py > gpurun_out/ncu_$tag.log 2>&1
  echo "ncu $tag exit $?" >> gpurun_out/summary.txt
done; }
cat gpurun_out/summary.txt; ls -la gpurun_out/*.ncu-rep

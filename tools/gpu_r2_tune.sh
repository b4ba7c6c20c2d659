#!/bin/bash
# 1 GPU: re-tune every conv shape with the pair / multicast variants among the candidates, then bench with the result.
mkdir -p gpurun_out; S=gpurun_out/r2_tune_summary.txt; rm -f $S
timeout 300 python -m pytest tests/test_gpu_conv.py -q -k "cta_pair" > gpurun_out/t_pair.log 2>&1; echo "pair tests exit $?" >> $S
export LDNN_TUNE_CACHE=$PWD/gpurun_out/tune_r2_pair.json LDNN_NO_TUNE_DEFAULTS=1
timeout 1200 python bench.py --steps 100 > gpurun_out/bench_tuned.log 2>&1; echo "bench (fresh tuning) exit $?" >> $S
unset LDNN_TUNE_CACHE LDNN_NO_TUNE_DEFAULTS
timeout 600 python bench.py --steps 100 --no_extra > gpurun_out/bench_defaults.log 2>&1; echo "bench (shipped defaults) exit $?" >> $S
cat $S; tail -2 gpurun_out/t_pair.log
for f in tuned defaults; do tail -1 gpurun_out/bench_$f.log | python -c "
import sys, json
d = json.loads(sys.stdin.read()); print('$f', round(d['value']), 'samples/s', round(d['ms_per_step'],4), 'ms/step e2e', round(d['e2e']['value']), 'bf16:', d.get('extra',{}).get('bf16'))"; done
python - <<'PY'
import json
t = json.load(open('gpurun_out/tune_r2_pair.json'))
for k, v in sorted(t['configs'].items()):
    if v[2] >= 2: print(k, v)
PY

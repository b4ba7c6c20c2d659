#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_conv.py tests/test_gpu_bn_head_optim.py tests/test_gpu_engine.py tests/test_zz_no_alloc.py -q -x --timeout=240 -p no:cacheprovider 2>&1 | tail -3
for v in "LDNN_PDL=1" "LDNN_PDL=0" "LDNN_PDL=1 LDNN_SERIAL=1"; do
  echo "== $v"; env $v timeout 300 python bench.py --steps 200 --warmup 20 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print(round(d['value']), 'samples/s', round(d['ms_per_step'], 4), 'ms/step  e2e', round(d['e2e']['value']), 'launches', d['gpu_launches'])"
done

#!/bin/bash
mkdir -p gpurun_out
for v in "" "LDNN_SERIAL=1" "LDNN_BUCKET_MB=64" "LDNN_BUCKET_MB=16"; do
  echo "== $v"; env $v timeout 300 python bench.py --steps 200 --warmup 20 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print(round(d['value']), 'samples/s', round(d['ms_per_step'], 4), 'ms/step  e2e', round(d['e2e']['value']), 'launches', d['gpu_launches'])"
done

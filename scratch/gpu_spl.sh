#!/bin/bash
for spl in 2 1 3 4 6 2; do
  B200PT_SPL=$spl timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('spl $spl', round(d['value']/1e9,4), 'G; e2e', round(d['e2e']['value']/1e9,4), 'trace ms', round(d['config']['trace_ms_per_step'],1), 'nif ms', round(d['config']['nif_ms_per_step'],1), 'TF', round(d['roofline']['achieved'],1), 'launches', d['gpu_launches'])"
done

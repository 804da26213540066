#!/bin/bash
for spl in 27 9 18 13; do
  B200PT_SPL=$spl timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 2960$((spl%10)) bench.py --gpus 8 --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | grep '^{' | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('N=8 spl $spl', round(d['value']/1e9,4), 'G; e2e', round(d['e2e']['value']/1e9,4), 'ms/step', round(d['ms_per_step'],2), 'nif ms', round(d['config']['nif_ms_per_step'],1), 'TF', round(d['roofline']['achieved'],1), 'launches', d['gpu_launches'])"
done

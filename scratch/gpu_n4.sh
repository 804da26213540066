#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29641 bench.py --gpus 4 --steps 5 --warmup 3 > gpurun_out/bench_g4.json 2> gpurun_out/bench_g4.err; echo "N=4 exit $?"
grep '^{' gpurun_out/bench_g4.json | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('N=4', round(d['value']/1e9,4), 'G; e2e', round(d['e2e']['value']/1e9,4), 'ms/step', round(d['ms_per_step'],2), 'TF', round(d['roofline']['achieved'],1), 'launches', d['gpu_launches'])"

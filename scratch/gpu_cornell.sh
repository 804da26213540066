#!/bin/bash
mkdir -p gpurun_out
python scratch/parity_probe.py 2>&1 | tail -4 | tee gpurun_out/parity_probe.log
timeout 1200 python -m pytest tests -q -m gpu 2>&1 | tail -8 | tee gpurun_out/pytest_gpu.log
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('bench', round(d['value']/1e9,4), 'G; trace ms', d['config']['trace_ms_per_step'], 'nif ms', d['config']['nif_ms_per_step'], 'TF', d['roofline']['achieved'])"

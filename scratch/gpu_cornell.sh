#!/bin/bash
mkdir -p gpurun_out
python scratch/cornell_diag.py 2>&1 | tail -5
timeout 900 python -m pytest tests/test_cornell.py tests/test_gpu_trace.py tests/test_gpu_kat.py tests/test_gpu_render.py -q -m gpu 2>&1 | tail -25 | tee gpurun_out/cornell_test.log
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('bench', round(d['value']/1e9,4), 'G; trace ms', d['config']['trace_ms_per_step'], 'nif ms', d['config']['nif_ms_per_step'], 'TF', d['roofline']['achieved'])"
B200PT_OVERLAP=0 timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('no-overlap bench', round(d['value']/1e9,4), 'G; trace ms', d['config']['trace_ms_per_step'], 'nif ms', d['config']['nif_ms_per_step'])"

#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_trace.py tests/test_gpu_kat.py tests/test_gpu_render.py tests/test_cornell.py -q -m gpu 2>&1 | tail -6
python scratch/parity_probe.py 2>&1 | tail -2
for rep in 1 2; do
for lib in libb200pt_prev.so libb200pt.so; do
  B200PT_LIB=$PWD/ipu_path_trace_b200/$lib timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$lib overlap   ', round(d['value']/1e9,4), 'G; trace ms', round(d['config']['trace_ms_per_step'],1), 'nif ms', round(d['config']['nif_ms_per_step'],1), 'TF', round(d['roofline']['achieved'],1), 'mhz', d['clocks']['sm_mhz'])"
done; done
for lib in libb200pt_prev.so libb200pt.so; do
  B200PT_OVERLAP=0 B200PT_LIB=$PWD/ipu_path_trace_b200/$lib timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$lib no-overlap', round(d['value']/1e9,4), 'G; trace ms', round(d['config']['trace_ms_per_step'],1), 'nif ms', round(d['config']['nif_ms_per_step'],1))"
done

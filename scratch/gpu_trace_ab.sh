#!/bin/bash
# trace kernel v2 (live-ray compaction + regeneration) vs v1: parity, stand-alone time, step time
one() { python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value']/1e9,4), round(d['e2e']['value']/1e9,4), round(d['roofline']['achieved'],1), d['clocks']['sm_mhz'], 'nif', round(d['config']['nif_ms_per_step'],2), 'trace', round(d['config']['trace_ms_per_step'],2), 'step', round(d['ms_per_step'],2))"; }
timeout 900 python -m pytest tests/test_gpu_baseline_configs.py tests/test_gpu_render.py tests/test_gpu_trace.py -m gpu -x -q 2>&1 | tail -3
for tv in v1 v2 v2p4 v2p2; do
  echo -n "$tv overlap=0: "; B200PT_TRACE=$tv B200PT_OVERLAP=0 timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline | one
done
for i in 1 2; do for tv in v1 v2 v2p4; do
  echo -n "$tv overlap=1: "; B200PT_TRACE=$tv timeout 300 python bench.py --steps 4 --warmup 3 --no-cpu-baseline | one
done; done

#!/bin/bash
# A/B on one box: $1 = variant library (relative to ipu_path_trace_b200/), default library is B
A=$PWD/ipu_path_trace_b200/$1; B=$PWD/ipu_path_trace_b200/libb200pt.so
timeout 600 python -m pytest tests/test_gpu_nif.py -m gpu -x -q 2>&1 | tail -2
for i in 1 2; do for lib in $A $B; do echo -n "$(basename $lib) C4 "; B200PT_LIB=$lib timeout 120 python scratch/nif_bench.py 4194304 20; done; done 2>&1
for i in 1 2; do for lib in $A $B; do echo -n "$(basename $lib) C2 "; B200PT_LIB=$lib timeout 300 python bench.py --steps 4 --warmup 3 --no-cpu-baseline | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value']/1e9, d['e2e']['value']/1e9, d['roofline']['achieved'], d['clocks']['sm_mhz'], d['config']['nif_ms_per_step'], d['config']['trace_ms_per_step'])"; done; done 2>&1

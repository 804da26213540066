#!/bin/bash
# A/B of two library builds on one box: NIF parity tests with the new build, C4 microbenchmark and bench.py for both.
mkdir -p gpurun_out
A=$PWD/ipu_path_trace_b200/libb200pt_prev.so; B=$PWD/ipu_path_trace_b200/libb200pt.so
timeout 600 python -m pytest tests/test_gpu_nif.py tests/test_gpu_render.py -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/pytest_ab.log
for i in 1 2; do for lib in $A $B; do echo -n "$(basename $lib) "; B200PT_LIB=$lib timeout 120 python scratch/nif_bench.py 4194304 20; done; done 2>&1 | tee gpurun_out/ab.log
timeout 120 python scratch/nif_prof.py 4194304 2>&1 | tee gpurun_out/prof.log
run() { echo "== $*"; env "$@" timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); c=d['config']; print(round(d['value']/1e6,1),'Msamples/s ms',round(d['ms_per_step'],1),'trace',round(c['trace_ms_per_step'],1),'nif',round(c['nif_ms_per_step'],1),'TF',round(d['roofline']['achieved'],1),'clk',d['clocks']['sm_mhz'],d['clocks']['power_w_max'])
"; }
{ run B200PT_LIB=$A; run B200PT_LIB=$B; run B200PT_LIB=$A; run B200PT_LIB=$B; } 2>&1 | tee gpurun_out/bench_ab.log

#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu2.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu2.log
tail -3 gpurun_out/pytest_gpu2.log
A=$PWD/ipu_path_trace_b200/libb200pt.so; B=$PWD/ipu_path_trace_b200/libb200pt_b.so
for i in 1 2; do for lib in $A $B; do echo -n "$(basename $lib) "; B200PT_LIB=$lib timeout 120 python scratch/nif_bench.py 4194304 20; done; done 2>&1 | tee gpurun_out/ab2.log
run() { echo "== $*"; env "$@" timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); c=d['config']; print(round(d['value']/1e6,1),'Msamples/s ms',round(d['ms_per_step'],1),'trace',round(c['trace_ms_per_step'],1),'nif',round(c['nif_ms_per_step'],1),'TF',round(d['roofline']['achieved'],1),'clk',d['clocks']['sm_mhz'],d['clocks']['power_w_max'])
"; }
{
run B200PT_LIB=$A B200PT_OVERLAP=0
run B200PT_LIB=$A B200PT_OVERLAP=1
run B200PT_LIB=$B B200PT_OVERLAP=0
run B200PT_LIB=$B B200PT_OVERLAP=1
run B200PT_LIB=$B B200PT_OVERLAP=1 B200PT_NIF_MODE=tp10
run B200PT_LIB=$A B200PT_OVERLAP=1 B200PT_NIF_MODE=tp10
} 2>&1 | tee gpurun_out/bench_ab2.log

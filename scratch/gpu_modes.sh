#!/bin/bash
for i in 1 2; do for mode in mc10 ts10 tp10 ss1; do echo -n "$mode "; B200PT_NIF_MODE=$mode timeout 120 python scratch/nif_bench.py 4194304 20; done; done 2>&1
run() { echo "== $*"; env "$@" timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); c=d['config']; print(round(d['value']/1e6,1),'Msamples/s ms',round(d['ms_per_step'],1),'trace',round(c['trace_ms_per_step'],1),'nif',round(c['nif_ms_per_step'],1),'TF',round(d['roofline']['achieved'],1),'clk',d['clocks']['sm_mhz'],d['clocks']['power_w_max'], 'launches', d['gpu_launches'])
"; }
run B200PT_NIF_MODE=mc10
run B200PT_NIF_MODE=tp10
run B200PT_NIF_MODE=ts10
run B200PT_NIF_MODE=mc10 B200PT_SPL=4
run B200PT_NIF_MODE=mc10 B200PT_SPL=1
run B200PT_NIF_MODE=mc10 B200PT_OVERLAP=0

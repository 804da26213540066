#!/bin/bash
# NIF-only sustained rate vs in-step rate; overlap on/off; 8 vs 16 epilogue warps
A=$PWD/ipu_path_trace_b200/libb200pt_e8.so; B=$PWD/ipu_path_trace_b200/libb200pt.so
one() { python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value']/1e9,4), round(d['e2e']['value']/1e9,4), round(d['roofline']['achieved'],1), d['clocks']['sm_mhz'], round(d['config']['nif_ms_per_step'],2), round(d['config']['trace_ms_per_step'],2), round(d['ms_per_step'],2))"; }
for lib in $A $B; do
  echo -n "$(basename $lib) C4 sustained 300 reps: "; B200PT_LIB=$lib timeout 200 python scratch/nif_bench.py 4194304 300
  for ov in 1 0; do echo -n "$(basename $lib) C2 overlap=$ov: "; B200PT_OVERLAP=$ov B200PT_LIB=$lib timeout 300 python bench.py --steps 4 --warmup 3 --no-cpu-baseline | one; done
done 2>&1

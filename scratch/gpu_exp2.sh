#!/bin/bash
P=$PWD/ipu_path_trace_b200
one() { python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value']/1e9,4), round(d['e2e']['value']/1e9,4), round(d['roofline']['achieved'],1), d['clocks']['sm_mhz'], round(d['config']['nif_ms_per_step'],2), round(d['config']['trace_ms_per_step'],2), round(d['ms_per_step'],2))"; }
timeout 600 python -m pytest tests/test_gpu_nif.py -m gpu -x -q 2>&1 | tail -2
echo "== prof 16 front"; B200PT_LIB=$P/libb200pt_prof16f.so python scratch/nif_prof.py 4194304
for i in 1 2; do for lib in "$@"; do
  echo -n "$lib C4 burst: "; B200PT_LIB=$P/$lib timeout 200 python scratch/nif_bench.py 4194304 20
  echo -n "$lib C4 sustained: "; B200PT_LIB=$P/$lib timeout 200 python scratch/nif_bench.py 4194304 300
done; done
for lib in "$@"; do for ov in 1 0; do echo -n "$lib C2 overlap=$ov: "; B200PT_OVERLAP=$ov B200PT_LIB=$P/$lib timeout 300 python bench.py --steps 4 --warmup 3 --no-cpu-baseline | one; done; done 2>&1

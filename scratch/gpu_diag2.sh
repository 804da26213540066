#!/bin/bash
# 2-GPU diagnosis of the 8-GPU findings: CLI start-up / first-save costs, 4K save steps, e2e gap at a 1/8 share per GPU
mkdir -p gpurun_out/diag2
A=/tmp/assets.extra; mkdir -p $A; cp tests/golden/nif_metadata.txt $A/; python -c "
import ipu_path_trace_b200 as pt; from ipu_path_trace_b200.nif import write_nif_bin; write_nif_bin('$A/nif.bin', pt.synthetic_model())"
host/b200_path_trace -o /tmp/c2.png --assets $A -w 1104 -h 1000 -s 30000 --samples-per-step 300 --ipus 2 --save-interval 10 2> gpurun_out/diag2/cli_c2_n2.log; grep -E "acquisition|Initialisation|Saved|skipped|finished|Samples/sec:" gpurun_out/diag2/cli_c2_n2.log | cut -c1-200 | head -30
awk '/Completed render step/ {print $9}' gpurun_out/diag2/cli_c2_n2.log | sort -g | tail -5 | tr '\n' ' '; echo " <- slowest C2 steps (s)"
host/b200_path_trace -o /tmp/c5.png --assets $A -w 3840 -h 2160 -s 2560 --samples-per-step 64 --ipus 2 --save-interval 10 2> gpurun_out/diag2/cli_c5_n2.log; grep -E "acquisition|Initialisation|Saved|skipped|finished|Samples/sec:" gpurun_out/diag2/cli_c5_n2.log | cut -c1-200
awk '/Completed render step/ {print $7, $9}' gpurun_out/diag2/cli_c5_n2.log | tr '\n' ';' | cut -c1-1500; echo
for sh in 8 1; do echo -n "e2e gap, 1/$sh share on one GPU: "; B200PT_BENCH_SHARDS=$sh python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['e2e']['ms_per_step'])"; done

#!/bin/bash
mkdir -p gpurun_out/n8
B200PT_BENCH_LOGDIR=gpurun_out/n8 timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r2_bench_n8.json 2> gpurun_out/r2_bench_n8.err; echo "n8 exit $?"
python - <<'PY'
import json
d = json.loads([l for l in open("gpurun_out/r2_bench_n8.json") if l.startswith("{")][-1])
print(8, d["value"] / 1e9, d["e2e"]["value"] / 1e9, d["roofline"]["frac"], d["ms_per_step"], d["e2e"]["ms_per_step"], d["clocks"])
for k in ("cli_c2", "cli_c5"):
    print(k, json.dumps(d["config"].get(k)))
PY
grep -E "acquisition|Initialisation|skipped|finished|Samples/sec:" gpurun_out/n8/cli_c5_n8.log | cut -c1-160
awk '/Completed render step/ {print $9}' gpurun_out/n8/cli_c5_n8.log | tr '\n' ' ' | cut -c1-1200; echo
OMP_NUM_THREADS=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29542 bench.py --impl reference --gpus 8 --steps 2 --warmup 1 2>/dev/null | cut -c1-700

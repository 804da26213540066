#!/bin/bash
# Round 3, multi-GPU visit: the multi-GPU parity tests and bench at N = all GPUs of the lease with config.multi_gpu_check.
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
python -m pytest tests/test_gpu_multi.py -x -q -m gpu > gpurun_out/r3_gputests_multi_n$NG.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r3_gputests_multi_n$NG.log
B200PT_BENCH_LOGDIR=gpurun_out timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $NG --steps 10 --warmup 3 --cli-spp ${CLI_SPP:-30000} > gpurun_out/r3_bench_n$NG.json 2> gpurun_out/r3_bench_n$NG.err; echo "bench rc=$?"
python - <<PY
import json
d = json.loads(open("gpurun_out/r3_bench_n$NG.json").read().strip().splitlines()[-1])
print("value", d["value"], "e2e", d["e2e"]["value"], "ms", d["ms_per_step"]); print("multi_gpu_check", d["config"].get("multi_gpu_check"))
for k in ("cli_c2", "cli_c5"):
    if k in d["config"]: print(k, {a: b for a, b in d["config"][k].items() if a not in ("clock", "cmd", "workload")})
PY

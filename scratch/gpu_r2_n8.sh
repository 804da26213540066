#!/bin/bash
# 8-GPU evidence: multi-GPU parity tests, bench N=8 (with the C++ host's full-length C2 run and C5), bench N=4
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -q 2>&1 | tail -3
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/r2_bench_n8.json 2> gpurun_out/r2_bench_n8.err; echo "n8 exit $?"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus 4 --steps 10 --warmup 3 --no-extras > gpurun_out/r2_bench_n4.json 2> gpurun_out/r2_bench_n4.err; echo "n4 exit $?"
python - <<'PY'
import json
for n in (8, 4):
    try:
        d = json.load(open(f"gpurun_out/r2_bench_n{n}.json"))
        print(n, d["value"] / 1e9, d["e2e"]["value"] / 1e9, d["roofline"]["frac"], d["clocks"])
        for k in ("cli_c2", "cli_c5"):
            if k in d["config"]: print(k, json.dumps(d["config"][k]))
    except Exception as e:
        print(n, "failed", e)
PY
tail -5 gpurun_out/r2_bench_n8.err

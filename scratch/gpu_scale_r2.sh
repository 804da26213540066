#!/bin/bash
# same-box scaling table: bench.py at N = 1, 2, 4, 8 on one 8-GPU box (no extras, no CPU baseline)
python bench.py --gpus 1 --steps 10 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r2_scale_n1.json 2>/dev/null
for n in 2 4 8; do python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29550+n)) bench.py --gpus $n --steps 10 --warmup 3 --no-extras > gpurun_out/r2_scale_n$n.json 2>/dev/null; done
python - <<'PY'
import json
base = None
for n in (1, 2, 4, 8):
    d = json.loads([l for l in open(f"gpurun_out/r2_scale_n{n}.json") if l.startswith("{")][-1])
    base = base or d["value"]
    print(n, round(d["value"] / 1e9, 4), "e2e", round(d["e2e"]["value"] / 1e9, 4), "eff", round(d["value"] / base / n, 4), "e2e/value", round(d["e2e"]["value"] / d["value"], 4), "frac", round(d["roofline"]["frac"], 4), d["clocks"]["sm_mhz"])
PY

#!/bin/bash
# 2-GPU visit: C++ host with --ipus 2 (NCCL film gather) vs --ipus 1, C2 sample rates, bench.py under torchrun.
mkdir -p gpurun_out /tmp/m
python - <<'PY'
import json, os, sys
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from test_host_cpu import METADATA
import ipu_path_trace_b200 as pt
from ipu_path_trace_b200.nif import write_nif_bin
os.makedirs('/tmp/m/assets.extra', exist_ok=True)
open('/tmp/m/assets.extra/nif_metadata.txt', 'w').write(json.dumps(METADATA))
write_nif_bin('/tmp/m/assets.extra/nif.bin', pt.synthetic_model())
PY
CLI=host/b200_path_trace; A=/tmp/m/assets.extra
for n in 1 2; do $CLI -o /tmp/m/small$n.png --assets $A -w 96 -h 64 -s 16 --samples-per-step 4 --ipus $n 2>&1 | tail -3; done
OPENCV_IO_ENABLE_OPENEXR=1 python - <<'PY' 2>&1 | tee gpurun_out/multi_cli_equal.log
import cv2, numpy as np
a = cv2.imread('/tmp/m/small1.exr', cv2.IMREAD_UNCHANGED); b = cv2.imread('/tmp/m/small2.exr', cv2.IMREAD_UNCHANGED)
print("1 vs 2 GPUs: max rel diff", float(np.max(np.abs(a - b) / np.maximum(np.abs(a), 1e-6))), "png equal", np.array_equal(cv2.imread('/tmp/m/small1.png'), cv2.imread('/tmp/m/small2.png')))
PY
for n in 1 2; do echo "== CLI C2 --ipus $n"; $CLI -o /tmp/m/c2_$n.png --assets $A -w 1104 -h 1000 -s 3000 --samples-per-step 300 --save-interval 5 --ipus $n 2>&1 | grep -E "Samples/sec|Render finished|Escaped|error|\[E\]" ; done 2>&1 | tee gpurun_out/multi_cli_c2.log
python bench.py --gpus 1 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_g1.json 2> gpurun_out/bench_g1.err; tail -c 600 gpurun_out/bench_g1.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_g2.json 2> gpurun_out/bench_g2.err; echo "exit $?"; tail -c 900 gpurun_out/bench_g2.json; tail -3 gpurun_out/bench_g2.err

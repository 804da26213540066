#!/bin/bash
# 8-GPU visit: bench.py at N = 4 and 8 under torchrun, and the C++ host with --ipus 8 (NCCL film gather) at C2 and C5.
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
for n in 8 4; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2953$n bench.py --gpus $n --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_g$n.json 2> gpurun_out/bench_g$n.err; echo "N=$n exit $?"; python -c "
import json,sys
d=json.loads([l for l in open('gpurun_out/bench_g$n.json') if l.startswith('{')][0]); print(d['n_gpus'], round(d['value']/1e9,3), 'G samples/s  e2e', round(d['e2e']['value']/1e9,3), 'ms/step', round(d['ms_per_step'],1), 'gather ms', d['config']['film_gather_ms_per_save_interval'], 'TF', round(d['roofline']['achieved'],1))"
done
CLI=host/b200_path_trace; A=/tmp/m/assets.extra
echo "== CLI C2 --ipus 8"; timeout 300 $CLI -o /tmp/m/c2_8.png --assets $A -w 1104 -h 1000 -s 6000 --samples-per-step 300 --save-interval 10 --ipus 8 2>&1 | grep -E "step (3|7|11|15|19|20)/|Samples/sec|Render finished|\[E\]" | tee gpurun_out/cli_c2_8.log
echo "== CLI C5 3840x2160 --ipus 8"; timeout 300 $CLI -o /tmp/m/c5_8.png --assets $A -w 3840 -h 2160 -s 640 --samples-per-step 64 --save-interval 2 --ipus 8 2>&1 | grep -E "step (3|5|7|9|10)/|Samples/sec|Render finished|Saved images at step 10|\[E\]" | tee gpurun_out/cli_c5_8.log
ls -la /tmp/m/*.png /tmp/m/*.exr | tee -a gpurun_out/cli_c5_8.log

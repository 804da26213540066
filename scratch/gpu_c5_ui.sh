#!/bin/bash
# C5: 3840x2160 progressive render on 8 GPUs with a remote-UI client attached (raw HDR image per save interval).
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
N=${1:-8}
host/b200_path_trace -o /tmp/m/c5_ui.png --assets /tmp/m/assets.extra -w 3840 -h 2160 -s 640000 --samples-per-step 64 --save-interval 20 --ipus $N \
   --ui-port 29600 --ui-accept-timeout 180 > gpurun_out/c5_ui_server.log 2>&1 &
SERVER=$!
timeout 600 python scratch/c5_ui_client.py 29600 2>&1 | tee gpurun_out/c5_ui_client.log
wait $SERVER; echo "server exit $?" | tee -a gpurun_out/c5_ui_client.log
grep -E "User interface|Remote UI|stopped by|Render finished|MiB raw image|Loading|\[E\]|\[W\]" gpurun_out/c5_ui_server.log | tail -20
grep -E "Completed render step" gpurun_out/c5_ui_server.log | awk 'NR%25==1' | tail -12

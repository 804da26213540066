#!/bin/bash
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
host/b200_path_trace -o /tmp/m/n.png --assets /tmp/m/assets.extra -w 1104 -h 1000 -s 6000 --samples-per-step 300 --save-interval 20 2>&1 | tail -3
OPENCV_IO_ENABLE_OPENEXR=1 python - <<'PY' 2>&1 | tee gpurun_out/nan_probe.log
import cv2, numpy as np, sys
sys.path.insert(0,'.')
a = cv2.imread('/tmp/m/n.exr', cv2.IMREAD_UNCHANGED)
bad = ~np.isfinite(a).all(axis=2)
print("non-finite pixels:", int(bad.sum()), "of", bad.size, "max finite", float(a[~bad].max()), "mean", float(a[~bad].mean()))
ys, xs = np.nonzero(bad); print(list(zip(ys[:10].tolist(), xs[:10].tolist())), a[bad][:5])
# NIF alone: 200 M random directions through the tensor-core kernel and the SIMT kernel
import ipu_path_trace_b200 as pt
t = pt.PathTracer(64, 64); model = pt.synthetic_model(); t.init_nif_weights(model)
rng = np.random.default_rng(0)
tot = 0; nbad = 0; worst = 0.0
for rep in range(20):
    uv = rng.random((4_000_000, 2), dtype=np.float32)
    out = t.nif_infer(uv[:, 0], uv[:, 1])
    b = ~np.isfinite(out).all(axis=1); nbad += int(b.sum()); tot += len(uv); worst = max(worst, float(out[~b].max()))
    if b.any() and nbad < 50: print("non-finite NIF output at uv", uv[b][:3], out[b][:3])
print("NIF evaluations", tot, "non-finite", nbad, "max finite", worst)
PY

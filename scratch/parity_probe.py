"""How exact is the main path against the oracle now? Prints agreement fractions (scratch, not a test)."""
import sys, numpy as np
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import ipu_path_trace_b200 as pt
from oracle import oracle as O
orc = O.Oracle() if hasattr(O, 'Oracle') else None
golden = np.load('tests/golden/ref_golden.npz')
model = pt.synthetic_model()
W, H = (int(x) for x in golden["img_wh"])
t = pt.PathTracer(W, H); t.init_nif_weights(model); t.init_render_settings()
rec = np.ascontiguousarray(golden["img_records"]).view(O.CONTRIB).reshape(W * H, -1)
esc_w, dir_w, thr_w, len_w = orc.path_summary(rec)
esc, d, thr, length = t.trace_fixed(golden["img_rays"], golden["img_uniforms"], 30)
agree = (esc == esc_w) & (length == len_w)
m = agree & (esc == 1)
print("trace_fixed vs reference records: paths", agree.size, "flag/length agree", agree.mean(), "dir bit-equal", (d[m].view(np.uint32) == dir_w[m].astype(np.float32).view(np.uint32)).all(axis=1).mean(),
      "thr max rel", float(np.max(np.abs(thr[m] - thr_w[m]) / np.maximum(np.abs(thr_w[m]), 1e-12))))
t.close()
W, H = 640, 480
s = O.Settings(width=W, height=H, seed=77, env_map_rotation_degrees=33.0)
t = pt.PathTracer(W, H); t.init_nif_weights(model); t.init_render_settings(seed=77, env_map_rotation_degrees=33.0)
items = O.make_worklist(W, H); t.setup(items)
tot = 0; bad = 0; uvbad = 0; uvn = 0
for sample in range(4):
    esc, uv, thr, length, cam = t.trace_wavefront(sample)
    e2, uv2, t2, l2, cam2 = orc.trace_philox(items, sample, s)
    a = (esc == e2) & (length == l2); tot += a.size; bad += int((~a).sum())
    m = a & (esc == 1); uvn += int(m.sum()) * 2; uvbad += int((uv[m].view(np.uint32) != uv2[m].astype(np.float32).view(np.uint32)).sum())
print("wavefront vs oracle: paths", tot, "flag/length mismatches", bad, "uv words not bit-equal", uvbad, "of", uvn)

"""north_star gate three at the BASELINE size, once, with more samples than the test-suite can afford: a 1104x1000 image at
SPP samples per pixel, GPU (tcgen05 path) against the CPU oracle (mode 0) on the same Philox streams, and against the oracle
on a DIFFERENT seed through 16x16-pixel block means (independent streams)."""
import sys, time; sys.path.insert(0, '.')
import numpy as np
import ipu_path_trace_b200 as pt
from oracle import oracle as O
W, H = 1104, 1000
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 16
model = pt.synthetic_model()
nif = O.NifLayers(model.embedding_dimension, model.max, model.mean, model.log_tone_map, model.layers)
orc = O.Oracle(); print("host threads", orc.use_all_host_threads())
items = pt.create_work_list_for_image(W, H)
rgb = lambda x: np.stack([x["r"], x["g"], x["b"]], 1).astype(np.float64)
def gpu(seed):
    t = pt.PathTracer(W, H, nif_impl=pt.NIF_TCGEN05); t.init_nif_weights(model); t.init_render_settings(seed=seed, samples_per_step=spp)
    t.setup(items); t.path_trace(); out, ctr = t.read_results(); t.close(); return out, ctr
t0 = time.time(); g1, c1 = gpu(1); print(f"GPU {W}x{H}x{spp}: {time.time()-t0:.2f} s, escaped {c1.escaped}")
t0 = time.time(); o1, esc, _, _ = orc.render(items, 0, spp, O.Settings(width=W, height=H, seed=1), nif, mode=0); print(f"oracle seed 1: {time.time()-t0:.1f} s")
a, b = rgb(g1), rgb(o1)
ok = np.isfinite(b).all(1)
print("same streams: escaped equal", c1.escaped == esc, "pathLength equal", np.array_equal(g1["pathLength"], o1["pathLength"]),
      "rRMSE %.3e" % (np.sqrt(np.mean((a[ok]-b[ok])**2))/np.mean(b[ok])), "pixels within 2e-2: %.6f" % np.isclose(a[ok], b[ok], rtol=2e-2, atol=1e-3*spp/16).all(1).mean(),
      "oracle NaN pixels", int((~ok).sum()))
t0 = time.time(); o2, _, _, _ = orc.render(items, 0, spp, O.Settings(width=W, height=H, seed=2), nif, mode=0); print(f"oracle seed 2: {time.time()-t0:.1f} s")
c = rgb(o2)
def blocks(x, n=16):
    img = np.zeros((H, W, 3)); img[items["v"], items["u"]] = x
    hh, ww = H // n * n, W // n * n
    return img[:hh, :ww].reshape(hh // n, n, ww // n, n, 3).mean((1, 3))
for n in (8, 16, 40):
    ga, ob, oc = blocks(a, n), blocks(b, n), blocks(np.nan_to_num(c), n)
    rr = lambda x, y: np.sqrt(np.mean((x - y) ** 2)) / np.mean(y)
    print(f"independent streams, {n}x{n} block means: GPU(seed 1) vs oracle(seed 2) rRMSE {rr(ga, oc):.4f}; oracle(1) vs oracle(2) {rr(ob, oc):.4f}; GPU(1) vs oracle(1) {rr(ga, ob):.2e}")
print("whole-image mean radiance: GPU(1) / oracle(2) = %.5f" % (a.mean() / np.nanmean(c)))

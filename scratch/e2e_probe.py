"""Where do the milliseconds of the host-buffer (e2e) step go at a 1/8 share per GPU? Times every call of the pipelined loop."""
import sys, time; sys.path.insert(0, '.')
import numpy as np, torch
import ipu_path_trace_b200 as pt
W, H = 1104, 1000
shards = int(sys.argv[1]) if len(sys.argv) > 1 else 8
wl = pt.shard_tiles(W, H, shards)[0]
t = pt.PathTracer(W, H, max_worklist=len(wl), nif_impl=pt.NIF_TCGEN05)
t.init_nif_weights(pt.synthetic_model()); t.init_render_settings(seed=1, samples_per_step=300)
pin_in, pin_out = pt.PinnedRecords(len(wl)), pt.PinnedRecords(len(wl))
pin_in.array[:] = wl
t.setup(pin_in.array)
def sync(): torch.cuda.synchronize()
acc = {}
def tick(name, t0):
    acc.setdefault(name, []).append((time.perf_counter() - t0) * 1e3)
for it in range(13):
    sync(); t0 = time.perf_counter()
    t.path_trace(sync=False); tick("path_trace enqueue", t0)
    t1 = time.perf_counter(); t.setup_next(pin_in.array); tick("setup_next enqueue", t1)
    t2 = time.perf_counter(); out, c = t.read_results(pin_out.array); tick("read_results (waits)", t2)
    tick("e2e step", t0)
    t3 = time.perf_counter(); t.read_results(pin_out.array); tick("read_results again (pack + D2H + sync only)", t3)
    acc.setdefault("device total_ms", []).append(c.total_ms)
for it in range(13):
    sync(); t0 = time.perf_counter()
    t.path_trace(sync=False); t.film_accumulate(); t.sync(); tick("resident step", t0)
    acc.setdefault("device total_ms (resident)", []).append(t.counters().total_ms)
for k, v in acc.items(): print(f"{k:50s} median {np.median(v[3:]):8.3f} ms")

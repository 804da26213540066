"""C5 with a remote-UI client attached: connects to host/b200_path_trace --ui-port, logs progress / sample rates / raw image
arrivals, changes the field of view once (render restarts), then stops the render. Scratch driver for scratch/gpu_c5_ui.sh."""
import sys, time
import numpy as np
sys.path.insert(0, 'tests')
from ui_client import UiClient
port = int(sys.argv[1])
t0 = time.time()
c = UiClient(port, connect_timeout=180)
print(f"[{time.time()-t0:6.2f}s] connected")
def wait_images(n, timeout=120):
    ok = c.wait(lambda c: len(c.images) >= n, timeout=timeout)
    with c.lock:
        img = c.images[-1] if c.images else None
        rate = c.rates[-1] if c.rates else None
        prog = c.progress[-1] if c.progress else None
    print(f"[{time.time()-t0:6.2f}s] images {len(c.images)} ok={ok} shape={None if img is None else img.shape} mean={None if img is None else float(img.mean()):.4f} "
          f"progress={prog} last rate (paths/s, rays/s)={rate} steps seen={len(c.progress)}")
wait_images(1); wait_images(2)
c.send_float("exposure", 0.5)
c.send_float("fov", 60.0)
n = len(c.images)
print(f"[{time.time()-t0:6.2f}s] fov -> 60 sent")
wait_images(n + 2)
with c.lock:
    drops = sum(1 for a, b in zip(c.progress, c.progress[1:]) if b < a)
    rates = np.array(c.rates)
print(f"restarts seen by the client: {drops}; median path rate {np.median(rates[:,0])/1e9:.2f} G samples/s over {len(rates)} steps (max {rates[:,0].max()/1e9:.2f})")
c.send_bool("stop")
c.closed.wait(60)
print(f"[{time.time()-t0:6.2f}s] stopped, link closed={c.closed.is_set()}")

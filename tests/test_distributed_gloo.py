"""Multi-rank host logic on CPU: world_size-2 gloo. The CUDA tracer is replaced by a stand-in that fills
its film deterministically from the pixels it was given, so sharding, padding and the per-save-interval
gather (one reduce of disjoint-support films) are exercised without a GPU."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
W, H = 100, 52


def pixel_value(u, v, step):
    return np.stack([u * 1.0 + step, v * 2.0, (u + v) * 0.5], -1).astype(np.float32)      # B,G,R


class FakeTracer:
    def __init__(self, n, device):
        self.film = np.zeros((H, W, 3), np.float32)
        self.steps = 0

    def init_nif_weights(self, model): pass
    def init_render_settings(self, **kw): pass
    def setup(self, worklist): self.wl = worklist
    def path_trace(self, sync=True): self.pending = True
    def sync(self): pass

    def film_accumulate(self):
        self.film[self.wl["v"], self.wl["u"]] += pixel_value(self.wl["u"], self.wl["v"], self.steps)
        self.steps += 1

    def film_read(self, scale=1.0):
        return self.film * np.float32(scale)


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from ipu_path_trace_b200.distributed import DistributedRenderer
    r = DistributedRenderer(W, H, model=None, samples_per_step=4, make_tracer=FakeTracer,
                            device_tensor=lambda n: torch.zeros(n, dtype=torch.float32))
    for _ in range(3):
        r.step()
    img = r.gather()
    n_mine = len(r.shard.worklist)
    q.put((rank, n_mine, None if img is None else img.copy()))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_render_and_gather(world):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    res.sort(key=lambda x: x[0])
    assert sum(n for _, n, _ in res) == W * H                        # every pixel rendered exactly once
    assert all(img is None for rk, _, img in res if rk != 0)
    img = res[0][2]
    vv, uu = np.divmod(np.arange(W * H), W)
    want = sum(pixel_value(uu, vv, s) for s in range(3)).reshape(H, W, 3) / 3.0
    np.testing.assert_allclose(img, want, rtol=1e-6)


def test_shards_partition_the_image():
    sys.path.insert(0, ROOT)
    import ipu_path_trace_b200 as pt
    for (w, h, g) in [(1104, 1000, 8), (320, 240, 3), (17, 5, 2), (64, 64, 1)]:
        shards = pt.shard_tiles(w, h, g)
        seen = np.zeros((h, w), np.int32)
        for s in shards:
            seen[s["v"], s["u"]] += 1
        assert np.all(seen == 1)
        sizes = [len(s) for s in shards]
        # balanced to within a few tiles (edge tiles are partial, so not exactly equal)
        assert max(sizes) - min(sizes) <= max(3 * 256, 0.015 * w * h / g)

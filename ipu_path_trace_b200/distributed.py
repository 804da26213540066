"""Multi-GPU rendering: one process per GPU, pixel tiles dealt to ranks, one film gather per save interval.

The reference spreads the shuffled pixel work list over all tiles of all IPUs and gives every IPU its own
NIF replica "so that there is no inter-ipu exchange of ray data" (src/PathTracerApp.cpp:205-252,
src/shard_utils.cpp:28-38); results come back through one host stream and are merged by
AccumulatedImage::accumulate (src/AccumulatedImage.cpp:59-74). Here each rank renders the pixel tiles
`load_balancer.shard_tiles` deals to it into its own fp32 film; nothing is exchanged inside the sample
loop. Every save interval the films, which have disjoint pixel support, are summed onto rank 0 with a
single `torch.distributed.reduce` (NCCL over NVLink on GPUs; gloo in the CPU tests): a gather written as
a reduce, exact because every pixel has exactly one non-zero contributor.

`torch.distributed` is plumbing only: the film lives in the CUDA library; it is copied into a torch
tensor on the same device (b200pt_film_read_device) for the collective.
"""
from __future__ import annotations

import os
from typing import Callable, Optional

import numpy as np

from .load_balancer import shard_tiles


def env_rank_world() -> tuple[int, int, int]:
    """(rank, world_size, local_rank) from the torchrun environment (1-process defaults)."""
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


class ShardedFilm:
    """Rank-local view of a width x height render dealt over `world` ranks."""

    def __init__(self, width: int, height: int, rank: int, world: int, tile: int = 16, seed: int = 142):
        if not (0 <= rank < world):
            raise ValueError("rank out of range")
        self.width, self.height, self.rank, self.world = width, height, rank, world
        self.worklist = shard_tiles(width, height, world, tile=tile, seed=seed)[rank]

    def owned_mask(self) -> np.ndarray:
        m = np.zeros((self.height, self.width), bool)
        m[self.worklist["v"], self.worklist["u"]] = True
        return m


def gather_film(local_film, dst: int = 0, group=None):
    """Sums the per-rank films (disjoint pixel support) onto rank `dst`. `local_film` is a torch tensor
    (CUDA for NCCL, CPU for gloo) holding this rank's film with zeros outside its pixels; it is reduced
    in place. Returns the tensor on `dst`, None elsewhere. One collective per save interval."""
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return local_film
    dist.reduce(local_film, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return local_film if dist.get_rank(group) == dst else None


class DistributedRenderer:
    """Drives one PathTracer per rank through the reference's render loop (PathTracerApp.cpp:648-776):
    setup once, then per step path_trace + film accumulate; `save(step)` gathers the film on rank 0.

    `make_tracer(worklist_len, device)` must return an object with the PathTracer programs; the default
    builds the CUDA PathTracer. Tests inject a stand-in to exercise sharding and the gather on CPU."""

    def __init__(self, width: int, height: int, model, samples_per_step: int, seed: int = 1,
                 make_tracer: Optional[Callable] = None, device_tensor: Optional[Callable] = None, **tracer_kw):
        self.rank, self.world, self.local = env_rank_world()
        self.width, self.height = width, height
        self.shard = ShardedFilm(width, height, self.rank, self.world)
        if make_tracer is None:
            from .tracer import PathTracer
            self.tracer = PathTracer(width, height, device=self.local, max_worklist=max(1, len(self.shard.worklist)), **tracer_kw)
        else:
            self.tracer = make_tracer(len(self.shard.worklist), self.local)
        self._device_tensor = device_tensor
        self.tracer.init_nif_weights(model)
        self.tracer.init_render_settings(seed=seed, samples_per_step=samples_per_step)
        self.tracer.setup(self.shard.worklist)
        self.steps = 0
        self._film = None

    def step(self, sync: bool = True):
        self.tracer.path_trace(sync=False)
        self.tracer.film_accumulate()
        if sync:
            self.tracer.sync()
        self.steps += 1

    def gather(self, scale: Optional[float] = None):
        """Film / steps (AccumulatedImage::saveImages divides by the step count, AccumulatedImage.cpp:48-56)
        gathered on rank 0 as an (H, W, 3) B,G,R float32 numpy array; None on other ranks."""
        import torch
        scale = (1.0 / max(1, self.steps)) if scale is None else scale
        if self._film is None:
            if self._device_tensor is not None:
                self._film = self._device_tensor(self.height * self.width * 3)
            else:
                self._film = torch.zeros(self.height * self.width * 3, dtype=torch.float32, device=f"cuda:{self.local}")
                # the library writes this tensor from its own (non-blocking) stream: torch's zero-fill must have landed
                torch.cuda.synchronize(self._film.device)
        if self._film.is_cuda:
            self.tracer.film_read_device(self._film.data_ptr(), scale)      # synchronises the library's stream
        else:
            self._film.copy_(torch.from_numpy(self.tracer.film_read(scale).reshape(-1)))
        out = gather_film(self._film)
        if self._film.is_cuda:
            # NCCL's reduce is asynchronous on every rank: the next gather must not overwrite the send buffer under it
            torch.cuda.synchronize(self._film.device)
        if out is None:
            return None
        return out.detach().cpu().numpy().reshape(self.height, self.width, 3)

"""Work lists and GPU sharding: the reference's LoadBalancer / shard_utils re-read for GPUs.

The reference deals equal chunks of a shuffled per-pixel work list to IPU tiles
(src/LoadBalancer.cpp:38-86,118-139) and gives every IPU a full NIF replica so that no ray data
crosses chips (src/PathTracerApp.cpp:205-252, src/shard_utils.cpp:28-38). Here the unit of
distribution is a square pixel tile: tiles are shuffled with the reference's seed (142) and dealt
round-robin to the GPUs, which keeps each warp's pixels coherent while spreading sky and
sphere regions evenly. Every GPU renders only its own pixels; the only exchange is one gather
of film pixels per save interval.
"""
from __future__ import annotations

import numpy as np

TRACE_RECORD = np.dtype(
    [("u", "<u2"), ("v", "<u2"), ("r", "<f4"), ("g", "<f4"), ("b", "<f4"), ("sampleCount", "<u2"), ("pathLength", "<u2")]
)
assert TRACE_RECORD.itemsize == 20     # src/codelets/TraceRecord.hpp:7-19
PAD_COORD = 65535                      # LoadBalancer.cpp:66-71


def create_work_list_for_image(width: int, height: int) -> np.ndarray:
    """Row-major list of every pixel (LoadBalancer.cpp:38-52)."""
    items = np.zeros(width * height, dtype=TRACE_RECORD)
    vv, uu = np.divmod(np.arange(width * height), width)
    items["u"] = uu
    items["v"] = vv
    return items


def pad_work_list(items: np.ndarray, multiple: int) -> np.ndarray:
    """Pads with null work (u = v = 65535) up to a multiple (LoadBalancer.cpp:62-71)."""
    n = len(items)
    padded = (n + multiple - 1) // multiple * multiple
    if padded == n:
        return items
    out = np.zeros(padded, dtype=TRACE_RECORD)
    out[:n] = items
    out["u"][n:] = PAD_COORD
    out["v"][n:] = PAD_COORD
    return out


def shard_tiles(width: int, height: int, num_gpus: int, tile: int = 16, seed: int = 142) -> list[np.ndarray]:
    """Deals tile x tile pixel tiles to `num_gpus` work lists after a seeded shuffle.
    Returns one TraceRecord array per GPU; together they cover every pixel exactly once."""
    if num_gpus < 1:
        raise ValueError("num_gpus must be >= 1")
    tx = (width + tile - 1) // tile
    ty = (height + tile - 1) // tile
    order = np.arange(tx * ty)
    np.random.default_rng(seed).shuffle(order)
    shards = []
    for g in range(num_gpus):
        mine = order[g::num_gpus]
        t_y, t_x = np.divmod(mine, tx)
        dy, dx = np.divmod(np.arange(tile * tile), tile)
        uu = (t_x[:, None] * tile + dx[None, :]).reshape(-1)
        vv = (t_y[:, None] * tile + dy[None, :]).reshape(-1)
        keep = (uu < width) & (vv < height)
        items = np.zeros(int(keep.sum()), dtype=TRACE_RECORD)
        items["u"] = uu[keep]
        items["v"] = vv[keep]
        shards.append(items)
    return shards


def clear_accumulators(items: np.ndarray) -> int:
    """LoadBalancer::clearInactiveAccumulators: zero the accumulators, return the path-length sum ("rays")."""
    total = int(items["pathLength"].astype(np.int64).sum())
    for k in ("r", "g", "b"):
        items[k] = 0
    items["sampleCount"] = 0
    items["pathLength"] = 0
    return total


def allocate_work_by_path_length(items: np.ndarray, num_bins: int) -> np.ndarray:
    """LoadBalancer::allocateWorkByPathLength (LoadBalancer.cpp:141-192): sort by last step's path
    length and hand every bin one cheap and one expensive item in turn."""
    order = np.argsort(items["pathLength"], kind="stable")
    srt = items[order]
    n = len(srt)
    lo, hi = 0, n - 1
    bins: list[list[int]] = [[] for _ in range(num_bins)]
    while lo <= hi:
        for b in bins:
            if lo > hi:
                break
            b.append(lo)
            lo += 1
            if lo > hi:
                break
            b.append(hi)
            hi -= 1
    flat = np.array([i for b in bins for i in b], dtype=np.int64)
    return srt[flat]

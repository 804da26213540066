"""Generates tests/golden/cornell.npz from the UNMODIFIED reference CPU tracer of the Cornell-box scene
(oracle/_ref/libref_cornell.so = external/light/src/trace.hpp compiled by oracle/Makefile). Run in the build
container, where /root/reference exists:  python tests/golden/make_cornell_golden.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402

ALT = dict(refractive_index=1.33, roulette_depth=2, stop_prob=0.35, epsilon=1e-6, intersection_epsilon=1e-4)
RENDER = dict(width=40, height=28, tile_width=16, tile_height=16, samples=8, seed=7, aa_noise_scale=1.0 / 700)


def main():
    O.build()
    ref = O.CornellReference()
    rng = np.random.default_rng(2024)
    n = 4096
    cam = rng.uniform(-1.0, 1.0, (n, 2)).astype(np.float32)
    state = rng.integers(1, 2 ** 63, (n, 2), dtype=np.uint64)
    out = {"cam_xy": cam, "rng_state": state}
    rgb, depth, st = ref.trace_paths(cam, state)
    out.update(rgb_default=rgb, depth_default=depth, state_default=st)
    rgb, depth, st = ref.trace_paths(cam, state, **ALT)
    out.update(rgb_alt=rgb, depth_alt=depth, state_alt=st)
    img, stats = ref.render(**RENDER)
    out.update(render_rgb=img, render_stats=np.array([stats["total_ray_casts"], stats["max_path_length"]], np.uint64))
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "cornell.npz"), **out)
    lit = (out["rgb_default"].sum(axis=1) > 0).mean()
    print(f"paths: {n}, lit fraction {lit:.3f}, mean depth {out['depth_default'].mean():.2f}; render mean {img.mean():.2f}, stats {stats}")


if __name__ == "__main__":
    main()

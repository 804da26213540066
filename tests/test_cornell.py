"""Second built-in scene: the Cornell box of the reference's stand-alone CPU renderer
(external/light/src/trace.cpp, trace.hpp). The oracle is that renderer compiled UNMODIFIED (oracle/_ref/libref_cornell.so)
and the golden vectors generated from it (tests/golden/cornell.npz, tests/golden/make_cornell_golden.py). The CUDA path must
return the reference's bits for the same xoroshiro128** streams: colours, ray-cast counts and the advanced generator state."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
ALT = dict(refractive_index=1.33, roulette_depth=2, stop_prob=0.35, epsilon=1e-6, intersection_epsilon=1e-4)
RENDER = dict(width=40, height=28, tile_width=16, tile_height=16, samples=8, seed=7, aa_noise_scale=1.0 / 700)


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(ROOT, "tests", "golden", "cornell.npz"))


@pytest.fixture(scope="module")
def cref(O):
    if not O.CornellReference.available():
        pytest.skip("oracle/_ref/libref_cornell.so not built (needs /root/reference)")
    return O.CornellReference()


@pytest.fixture(scope="module")
def corc(O, orc):
    return O.CornellOracle()          # orc has built liboracle.so


# ---- CPU: the fixtures are the reference's own outputs --------------------------------------------------------------

def test_restatement_reproduces_the_golden_vectors(gold, corc):
    """oracle/cornell_oracle.cpp against the compiled reference's outputs: every path, the image and the statistics, bit for bit."""
    for which, kw in (("default", {}), ("alt", ALT)):
        rgb, depth, st = corc.trace_paths(gold["cam_xy"], gold["rng_state"], **kw)
        assert np.array_equal(rgb.view(np.uint32), gold["rgb_" + which].view(np.uint32))
        assert np.array_equal(depth, gold["depth_" + which]) and np.array_equal(st, gold["state_" + which])
    img, stats = corc.render(**RENDER)
    assert np.array_equal(img.view(np.uint32), gold["render_rgb"].view(np.uint32))
    assert [stats["total_ray_casts"], stats["max_path_length"]] == gold["render_stats"].tolist()


def test_restatement_equals_the_compiled_reference_on_fresh_inputs(corc, cref):
    rng = np.random.default_rng(5)
    n = 50000
    cam = rng.uniform(-1.5, 1.5, (n, 2)).astype(np.float32)
    state = rng.integers(1, 2 ** 63, (n, 2), dtype=np.uint64)
    for kw in ({}, dict(refractive_index=1.1, roulette_depth=1, stop_prob=0.5), dict(roulette_depth=25, stop_prob=0.0)):
        a, b = corc.trace_paths(cam, state, **kw), cref.trace_paths(cam, state, **kw)
        assert np.array_equal(a[0].view(np.uint32), b[0].view(np.uint32)) and np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])
    kw = dict(width=33, height=21, tile_width=7, tile_height=5, samples=3, seed=99, aa_noise_scale=0.02)
    (ia, sa), (ib, sb) = corc.render(**kw), cref.render(**kw)
    assert np.array_equal(ia.view(np.uint32), ib.view(np.uint32)) and sa == sb


def test_golden_vectors_are_the_compiled_reference(gold, cref):
    rgb, depth, st = cref.trace_paths(gold["cam_xy"], gold["rng_state"])
    assert np.array_equal(rgb.view(np.uint32), gold["rgb_default"].view(np.uint32))
    assert np.array_equal(depth, gold["depth_default"]) and np.array_equal(st, gold["state_default"])
    rgb, depth, st = cref.trace_paths(gold["cam_xy"], gold["rng_state"], **ALT)
    assert np.array_equal(rgb.view(np.uint32), gold["rgb_alt"].view(np.uint32))
    assert np.array_equal(depth, gold["depth_alt"]) and np.array_equal(st, gold["state_alt"])
    img, stats = cref.render(**RENDER)
    assert np.array_equal(img.view(np.uint32), gold["render_rgb"].view(np.uint32))
    assert [stats["total_ray_casts"], stats["max_path_length"]] == gold["render_stats"].tolist()


def test_golden_vectors_cover_the_scene(gold):
    """Emitter hits through every material, roulette stops, depth-limit paths and unlit paths are all present."""
    rgb, depth = gold["rgb_default"], gold["depth_default"]
    lit = rgb.sum(axis=1) > 0
    assert 0.05 < lit.mean() < 0.5 and (depth == 20).any() and (depth[lit] == 0).any() and (depth[lit] >= 5).any()
    assert (depth[~lit] == 3).any()                       # stopped by the first roulette draw
    assert gold["render_rgb"].shape == (28, 40, 3) and gold["render_stats"][1] <= 20


def test_cornell_needs_the_gpu_library(lib_built):
    """No CPU path: without a device the entry points fail with the library's message; bad arguments are rejected first."""
    import torch
    from ipu_path_trace_b200 import _lib, cornell
    lib = _lib.load()
    cfg = cornell.CornellOptions(width=8, height=8).to_c()
    cfg.struct_size = 4
    rgb = np.zeros((8, 8, 3), np.float32)
    assert lib.b200pt_cornell_render(C.byref(cfg), rgb.ctypes.data, None) != 0
    assert b"struct_size" in lib.b200pt_cornell_last_error()
    with pytest.raises(RuntimeError, match="must be > 0"):
        cornell.render(cornell.CornellOptions(width=0))
    with pytest.raises(RuntimeError, match="stop_prob"):
        cornell.render(cornell.CornellOptions(stop_prob=1.0))
    if not torch.cuda.is_available():
        with pytest.raises(RuntimeError, match="no CUDA device"):
            cornell.render(cornell.CornellOptions(width=8, height=8, samples=1))


# ---- GPU: parity through the C ABI ----------------------------------------------------------------------------------

def _opts(**kw):
    from ipu_path_trace_b200 import cornell
    return cornell.CornellOptions(**kw)


@pytest.mark.gpu
@pytest.mark.parametrize("which", ["default", "alt"])
def test_paths_match_the_reference_bit_for_bit(lib_built, gold, which):
    from ipu_path_trace_b200 import cornell
    rgb, depth, st = cornell.trace_paths(_opts(**(ALT if which == "alt" else {})), gold["cam_xy"], gold["rng_state"])
    assert np.array_equal(depth, gold["depth_" + which])
    assert np.array_equal(st, gold["state_" + which])
    assert np.array_equal(rgb.view(np.uint32), gold["rgb_" + which].view(np.uint32))


@pytest.mark.gpu
def test_tile_stream_render_is_the_reference_image(lib_built, gold):
    """TILE_STREAMS reproduces trace.cpp's image for a seed: jumped per-tile streams, ragged edge tiles, p += colour / spp."""
    from ipu_path_trace_b200 import cornell
    img, stats = cornell.render(_opts(rng_mode=cornell.TILE_STREAMS, **RENDER))
    assert [stats["total_ray_casts"], stats["max_path_length"]] == gold["render_stats"].tolist()
    assert np.array_equal(img.view(np.uint32), gold["render_rgb"].view(np.uint32))


@pytest.mark.gpu
def test_against_the_oracle_on_fresh_inputs(lib_built, corc):
    """Larger seeded inputs than the fixture, against the CPU restatement (itself pinned to the compiled reference above)."""
    cref = corc
    from ipu_path_trace_b200 import cornell
    rng = np.random.default_rng(99)
    n = 200000
    cam = rng.uniform(-1.2, 1.2, (n, 2)).astype(np.float32)
    state = rng.integers(1, 2 ** 63, (n, 2), dtype=np.uint64)
    want = cref.trace_paths(cam, state)
    got = cornell.trace_paths(_opts(), cam, state)
    for g, w in zip(got, want):
        assert np.array_equal(g.view(np.uint32) if g.dtype == np.float32 else g, w.view(np.uint32) if w.dtype == np.float32 else w)
    kw = dict(width=70, height=33, tile_width=8, tile_height=16, samples=5, seed=12345, aa_noise_scale=0.01)
    wimg, wstats = cref.render(**kw)
    img, stats = cornell.render(_opts(rng_mode=cornell.TILE_STREAMS, **kw))
    assert np.array_equal(img.view(np.uint32), wimg.view(np.uint32))
    assert stats["total_ray_casts"] == wstats["total_ray_casts"] and stats["max_path_length"] == wstats["max_path_length"]


@pytest.mark.gpu
def test_pixel_stream_render_converges_to_the_same_image(lib_built):
    """The GPU schedule draws different random numbers: same estimator, so block means agree at high sample counts,
    ray casts per path agree, and the image does not depend on the slice count chosen for occupancy."""
    from ipu_path_trace_b200 import cornell
    w, h = 64, 64
    a, sa = cornell.render(_opts(width=w, height=h, samples=4096, seed=1))
    b, sb = cornell.render(_opts(width=w, height=h, samples=4096, seed=2))
    t, st = cornell.render(_opts(width=w, height=h, samples=256, seed=3, rng_mode=cornell.TILE_STREAMS))
    blocks = lambda x: x.reshape(4, h // 4, 4, w // 4, 3).mean(axis=(1, 3))
    rel = lambda x, y: float(np.abs(blocks(x) - blocks(y)).mean() / blocks(y).mean())
    assert rel(a, b) < 0.05, rel(a, b)
    assert rel(a, t) < 0.12, rel(a, t)
    per_path = lambda s, spp: s["total_ray_casts"] / (w * h * spp)
    assert abs(per_path(sa, 4096) - per_path(st, 256)) / per_path(st, 256) < 0.01
    assert sa["max_path_length"] == 20 and sa["device_ms"] > 0
    assert np.isfinite(a).all() and a.min() >= 0


@pytest.mark.gpu
def test_cli_b200_trace_writes_the_reference_image(lib_built, gold, tmp_path):
    """host/b200_trace = external/light's `trace` command line on the GPU: with --rng tile the EXR holds the reference's
    radiance bit for bit and the 8-bit image is min(radiance, 255) (trace.cpp:82-98); options file and statistics as :46-54,296-310."""
    import subprocess
    import cv2
    os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
    subprocess.run(["make", "-C", os.path.join(ROOT, "host"), "b200_trace"], check=True, capture_output=True)
    out = str(tmp_path / "cornell.png")
    r = subprocess.run([os.path.join(ROOT, "host", "b200_trace"), "-o", out, "-w", "40", "-h", "28", "-s", "8", "--seed", "7", "--rng", "tile", "--no-gui"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    exr = cv2.imread(out + ".exr", cv2.IMREAD_UNCHANGED)
    assert np.array_equal(np.ascontiguousarray(exr[:, :, ::-1]).view(np.uint32), gold["render_rgb"].view(np.uint32))
    png = cv2.imread(out, cv2.IMREAD_UNCHANGED)
    want = np.rint(np.minimum(gold["render_rgb"][:, :, ::-1], 255.0)).astype(np.uint8)
    assert np.array_equal(png, want)
    total, longest = (int(x) for x in gold["render_stats"])
    assert f"Total Rays: {total} " in r.stdout and f"Max path length: {longest}" in r.stdout and "Render time:" in r.stdout
    opts = dict(l.split(" ", 1) for l in open(out + ".command_line_options").read().splitlines())
    assert opts["samples"] == "8" and opts["stop-prob"] == "0.2" and opts["tile-width"] == "16" and opts["rng"] == "tile"
    assert open(out + ".hist.txt").read().split() == [str(longest)]
    # errors: missing output name, bad mode
    bad = subprocess.run([os.path.join(ROOT, "host", "b200_trace"), "-w", "8"], capture_output=True, text=True)
    assert bad.returncode != 0 and "outfile" in bad.stderr
    bad = subprocess.run([os.path.join(ROOT, "host", "b200_trace"), "-o", out, "--rng", "other"], capture_output=True, text=True)
    assert bad.returncode != 0 and "--rng" in bad.stderr


def _splitmix64(z):
    z = (z + np.uint64(0x9e3779b97f4a7c15))
    z = (z ^ (z >> np.uint64(30))) * np.uint64(0xbf58476d1ce4e5b9)
    z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94d049bb133111eb)
    return z ^ (z >> np.uint64(31))


def _xo_next(s0, s1):
    rotl = lambda x, k: (x << np.uint64(k)) | (x >> np.uint64(64 - k))
    r = rotl(s0 * np.uint64(5), 7) * np.uint64(9)
    b = s1 ^ s0
    return r, rotl(s0, 24) ^ b ^ (b << np.uint64(16)), rotl(b, 37)


@pytest.mark.gpu
def test_pixel_stream_schedule_is_the_per_path_function_sample_by_sample(lib_built):
    """PIXEL_STREAMS runs a bounce-level state machine (a lane starts its next sample as soon as its path ends). Its image must be
    what the parity-tested per-path entry point gives when the same per-thread stream is replayed on the host: splitmix64 seeding,
    two jitter draws (y first), the path, next sample - bit for bit when one thread owns a pixel (slices = 1)."""
    from ipu_path_trace_b200 import cornell
    w, h, spp, seed, aa = 1024, 600, 3, 11, 1.0 / 700            # 614 400 pixels >= two waves of threads -> slices = 1
    img, stats = cornell.render(_opts(width=w, height=h, samples=spp, seed=seed, aa_noise_scale=aa))
    with np.errstate(over="ignore"):
        t = np.arange(w * h, dtype=np.uint64)
        s0 = _splitmix64(np.uint64(seed) ^ (np.uint64(0xd1342543de82ef95) * (t + np.uint64(1))))
        s1 = _splitmix64(s0)
        libm = C.CDLL("libm.so.6")
        libm.tanf.restype = C.c_float
        libm.tanf.argtypes = [C.c_float]
        f = np.float32
        tan_theta = f(libm.tanf(f(f(3.14159265358979323846) / f(2)) / f(2)))
        col = (t % np.uint64(w)).astype(np.float32)
        row = (t // np.uint64(w)).astype(np.float32)
        x = col / f(w) - f(.5)
        y = row / f(h) - f(.5)
        base_x = f(2) * x * (f(w) / f(h)) * tan_theta
        base_y = f(-2) * y * tan_theta
        total = np.zeros((w * h, 3), np.float32)
        casts = 0
        for _ in range(spp):
            r, s0, s1 = _xo_next(s0, s1)
            ny = (2.0 * (((np.uint64(0x3FF) << np.uint64(52)) | (r >> np.uint64(12))).view(np.float64) - 1.0) - 1.0).astype(np.float32)
            r, s0, s1 = _xo_next(s0, s1)
            nx = (2.0 * (((np.uint64(0x3FF) << np.uint64(52)) | (r >> np.uint64(12))).view(np.float64) - 1.0) - 1.0).astype(np.float32)
            cam = np.stack([base_x + nx * f(aa), base_y + ny * f(aa)], axis=1)
            rgb, depth, st = cornell.trace_paths(_opts(), cam, np.stack([s0, s1], axis=1))
            s0, s1 = st[:, 0].copy(), st[:, 1].copy()
            total = total + rgb / f(spp)
            casts += int(depth.sum())
    assert stats["total_ray_casts"] == casts
    assert np.array_equal(img.reshape(-1, 3).view(np.uint32), total.view(np.uint32))

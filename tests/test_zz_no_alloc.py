"""b200pt_cornell_render_noalloc, the one entry point of the library that allocates nothing.

Who needs it: code that is not allowed to ask the CUDA driver for memory at the moment it renders. Think of a plug-in that
was given one output buffer and a fixed slice of the GPU by its host application, a frame loop that cannot take the
device-wide stall cudaMalloc brings, or a process sharing the GPU with an allocator that already holds everything that is
left. For them the tile-stream Cornell render keeps its working memory (per-tile generator states, two counters) in an array
that is part of the library image, and writes the image into a device buffer the caller passes in.

This file is named so that pytest collects it last: `pytest -x` stops at the first failure, and a problem in this extra
entry point must not hide the parity tests. The GPU test has to hold both in a fresh process (it then performs the library's
first-use set-up itself) and after every other test has used the same library and device.
"""
import ctypes as C
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _opts(**kw):
    from ipu_path_trace_b200 import cornell
    return cornell.CornellOptions(rng_mode=cornell.TILE_STREAMS, **kw)


def test_limit_and_argument_errors_need_no_gpu(lib_built):
    """Rejections happen before any device work, in the order: configuration, rng_mode, tile limit, buffer."""
    import torch
    from ipu_path_trace_b200 import _lib, cornell
    lib = _lib.load()
    assert cornell.NOALLOC_MAX_TILES == 16384
    cfg = _opts(width=8, height=8).to_c()
    cfg.struct_size = 4
    assert lib.b200pt_cornell_render_noalloc(C.byref(cfg), None, None) == -1
    assert b"struct_size" in lib.b200pt_cornell_last_error()
    if not torch.cuda.is_available():
        with pytest.raises(RuntimeError, match="no CUDA device"):
            cornell.render_noalloc(_opts(width=8, height=8, samples=1), 0)


@pytest.mark.gpu
def test_noalloc_render_equals_the_allocating_render_twice_in_a_row(lib_built):
    """Same bits as b200pt_cornell_render (tile streams) on two consecutive calls with DIFFERENT shapes and seeds, the second
    smaller than the first: the shared array then still holds the first call's states and counters, and the caller's buffer
    still holds the first image, so anything the entry point forgot to reset would show. Then the first configuration again."""
    import torch
    from ipu_path_trace_b200 import cornell
    first = dict(width=70, height=33, tile_width=8, tile_height=16, samples=5, seed=12345, aa_noise_scale=0.01)
    second = dict(width=40, height=28, tile_width=16, tile_height=16, samples=8, seed=7, aa_noise_scale=1.0 / 700)
    buf = torch.full((70 * 33 * 3,), float("nan"), dtype=torch.float32, device="cuda:0")
    for kw in (first, second, first):
        want_img, want_stats = cornell.render(_opts(**kw))
        n = kw["width"] * kw["height"] * 3
        stats = cornell.render_noalloc(_opts(**kw), buf.data_ptr())
        got = buf[:n].cpu().numpy().reshape(kw["height"], kw["width"], 3)
        assert stats["total_ray_casts"] == want_stats["total_ray_casts"] and stats["max_path_length"] == want_stats["max_path_length"]
        assert stats["device_ms"] == 0.0                    # documented: no events, hence no timing
        assert np.array_equal(got.view(np.uint32), want_img.view(np.uint32))     # bit-exact, the tolerance of tests/test_cornell.py


@pytest.mark.gpu
def test_noalloc_render_makes_no_device_allocation_and_rejects_what_does_not_fit(lib_built):
    """Free device memory is the same before and after a (non-first) call; too many tiles, the other rng_mode and a host
    pointer are refused with messages that say what to do instead."""
    import torch
    from ipu_path_trace_b200 import cornell
    kw = dict(width=64, height=48, tile_width=16, tile_height=16, samples=2, seed=3)
    buf = torch.zeros(640 * 480 * 3, dtype=torch.float32, device="cuda:0")
    cornell.render_noalloc(_opts(**kw), buf.data_ptr())                  # first use on this device may load code
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info(0)
    cornell.render_noalloc(_opts(**kw), buf.data_ptr())
    free1, _ = torch.cuda.mem_get_info(0)
    assert free1 == free0
    with pytest.raises(RuntimeError, match="16385 tiles exceed the 16384"):
        cornell.render_noalloc(_opts(width=16385, height=1, tile_width=1, tile_height=1, samples=1), buf.data_ptr())
    cornell.render_noalloc(_opts(width=640, height=400, tile_width=4, tile_height=4, samples=1), buf.data_ptr())       # 16000 tiles: fits
    with pytest.raises(RuntimeError, match="TILE_STREAMS only"):
        cornell.render_noalloc(cornell.CornellOptions(width=8, height=8, samples=1), buf.data_ptr())
    host = np.zeros((8, 8, 3), np.float32)
    with pytest.raises(RuntimeError, match="device memory"):
        cornell.render_noalloc(_opts(width=8, height=8, samples=1), host.ctypes.data)
    with pytest.raises(RuntimeError, match="null image buffer"):
        cornell.render_noalloc(_opts(width=8, height=8, samples=1), 0)

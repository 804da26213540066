"""The Cornell-box scene of the reference's stand-alone CPU renderer (external/light/src/trace.cpp) on the GPU:
Python mirror of b200pt_cornell_render / b200pt_cornell_trace_paths (include/b200pt.h). Option names and defaults
are those of trace.cpp:19-35. There is no CPU path."""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import _lib

PIXEL_STREAMS, TILE_STREAMS = 0, 1


@dataclass
class CornellOptions:
    width: int = 256
    height: int = 256
    tile_width: int = 16
    tile_height: int = 16
    samples: int = 32
    refractive_index: float = 1.5
    roulette_depth: int = 3
    stop_prob: float = 0.2
    aa_noise_scale: float = 1.0 / 700
    seed: int = 1
    epsilon: float = float(np.finfo(np.float32).eps)
    intersection_epsilon: float = 1e-5
    rng_mode: int = PIXEL_STREAMS
    device: int = 0

    def to_c(self) -> _lib.CornellConfig:
        return _lib.CornellConfig(C.sizeof(_lib.CornellConfig), self.device, self.width, self.height, self.tile_width, self.tile_height,
                                  self.samples, self.roulette_depth, self.seed, self.aa_noise_scale, self.refractive_index, self.stop_prob,
                                  self.epsilon, self.intersection_epsilon, self.rng_mode)


def _check(lib, rc: int, what: str) -> None:
    if rc != 0:
        raise RuntimeError(f"{what} failed ({rc}): {lib.b200pt_cornell_last_error().decode()}")


def render(opts: CornellOptions):
    """-> (height x width x 3 float32 R,G,B mean radiance, stats dict)."""
    lib = _lib.load()
    rgb = np.zeros((opts.height, opts.width, 3), np.float32)
    stats = _lib.CornellStats()
    cfg = opts.to_c()
    _check(lib, lib.b200pt_cornell_render(C.byref(cfg), rgb.ctypes.data, C.byref(stats)), "b200pt_cornell_render")
    return rgb, {"total_ray_casts": int(stats.total_ray_casts), "max_path_length": int(stats.max_path_length), "device_ms": float(stats.device_ms)}


NOALLOC_MAX_TILES = 16384          # B200PT_CORNELL_NOALLOC_MAX_TILES


def render_noalloc(opts: CornellOptions, rgb_device_ptr: int):
    """b200pt_cornell_render_noalloc: the TILE_STREAMS render for callers that may not allocate GPU memory while rendering
    (a plug-in on a fixed memory share, a frame loop that avoids cudaMalloc's device-wide sync). The image is written to the
    caller's DEVICE buffer at `rgb_device_ptr` (height x width x 3 float32 on opts.device); the kernel's working memory is an
    array built into the library, good for NOALLOC_MAX_TILES tiles. Calls on one device must not overlap. -> stats dict."""
    lib = _lib.load()
    stats = _lib.CornellStats()
    cfg = opts.to_c()
    _check(lib, lib.b200pt_cornell_render_noalloc(C.byref(cfg), C.c_void_p(rgb_device_ptr), C.byref(stats)), "b200pt_cornell_render_noalloc")
    return {"total_ray_casts": int(stats.total_ray_casts), "max_path_length": int(stats.max_path_length), "device_ms": float(stats.device_ms)}


def trace_paths(opts: CornellOptions, cam_xy: np.ndarray, rng_state: np.ndarray):
    """cam_xy (n,2) float32, rng_state (n,2) uint64 -> (rgb (n,3), depth (n,), advanced state (n,2))."""
    lib = _lib.load()
    cam = np.ascontiguousarray(cam_xy, np.float32)
    state = np.array(rng_state, np.uint64, copy=True, order="C")
    n = cam.shape[0]
    rgb = np.zeros((n, 3), np.float32)
    depth = np.zeros(n, np.uint32)
    cfg = opts.to_c()
    _check(lib, lib.b200pt_cornell_trace_paths(C.byref(cfg), n, cam.ctypes.data, state.ctypes.data, rgb.ctypes.data, depth.ctypes.data),
           "b200pt_cornell_trace_paths")
    return rgb, depth, state

"""PathTracer: the host-side mirror of the reference's device programs over the C ABI.

Method names are the reference's program names (src/PathTracerApp.cpp:479-483):
init_nif_weights, init_render_settings, setup, path_trace, read_results - so that the render loop
of PathTracerApp::execute (:648-776) reads the same here. Everything runs in libb200pt.so on the
GPU; this module holds no arithmetic of its own and has no fallback.
"""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass

import numpy as np

from . import _lib
from .load_balancer import TRACE_RECORD
from .nif import NifModel

AA_NORMAL, AA_UNIFORM, AA_TRUNCATED_NORMAL = 0, 1, 2
AA_NOISE_TYPES = {"normal": AA_NORMAL, "uniform": AA_UNIFORM, "truncated-normal": AA_TRUNCATED_NORMAL}
NIF_AUTO, NIF_TCGEN05, NIF_SIMT = 0, 1, 2

KAT_PIXEL_TO_RAY, KAT_SPHERE, KAT_DISC, KAT_DIFFUSE, KAT_REFLECT, KAT_REFRACT, KAT_ROULETTE, KAT_ESCAPED_UV, \
    KAT_PHILOX, KAT_AA_NOISE, KAT_NIF_FEATURES = range(11)


class B200PTError(RuntimeError):
    pass


@dataclass
class Counters:
    paths: int
    escaped: int
    path_segments: int
    trace_ms: float
    nif_ms: float
    total_ms: float
    kernel_launches: int
    nif_impl_used: int


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class PinnedRecords:
    """A TraceRecord array in page-locked host memory (b200pt_host_alloc): the host side of setup / read_results at the
    full host-link rate. `.array` is a numpy view; the memory is released by close() or with the object."""

    def __init__(self, n: int):
        self.lib = _lib.load()
        p = C.c_void_p()
        rc = self.lib.b200pt_host_alloc(C.byref(p), C.c_size_t(max(1, n) * TRACE_RECORD.itemsize))
        if rc != 0 or not p.value:
            raise B200PTError(f"b200pt_host_alloc failed ({rc})")
        self.ptr = p
        buf = (C.c_ubyte * (n * TRACE_RECORD.itemsize)).from_address(p.value)
        self.array = np.frombuffer(buf, dtype=TRACE_RECORD, count=n)

    def close(self):
        if self.ptr is not None:
            self.array = None
            self.lib.b200pt_host_free(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class PathTracer:
    """One context = one GPU (`--ipus N` of the reference becomes N contexts, one per GPU)."""

    def __init__(self, width: int, height: int, device: int = 0, max_path_length: int = 10, roulette_depth: int = 3,
                 stop_prob: float = 0.3, refractive_index: float = 1.5, aa_noise_type: int = AA_NORMAL,
                 max_worklist: int | None = None, samples_per_launch: int = 0, nif_impl: int = NIF_AUTO):
        self.lib = _lib.load()
        self.width, self.height = width, height
        cfg = _lib.Config(C.sizeof(_lib.Config), device, width, height, max_path_length, roulette_depth, stop_prob,
                          refractive_index, aa_noise_type, max_worklist or width * height, samples_per_launch, nif_impl)
        h = C.c_void_p()
        rc = self.lib.b200pt_create(C.byref(cfg), C.byref(h))
        if rc != 0:
            raise B200PTError(f"b200pt_create failed ({rc}): {self.lib.b200pt_last_error(None).decode()}")
        self.h = h
        self.n_items = 0

    def close(self):
        if getattr(self, "h", None):
            self.lib.b200pt_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc: int, what: str):
        if rc != 0:
            raise B200PTError(f"{what} failed ({rc}): {self.lib.b200pt_last_error(self.h).decode()}")

    # ---- the five programs -----------------------------------------------------------------
    def init_nif_weights(self, model: NifModel):
        keep = []
        arr = (_lib.NifLayer * len(model.layers))()
        for i, l in enumerate(model.layers):
            k = np.ascontiguousarray(l["kernel"], dtype=np.float16)
            b = None if l.get("bias") is None else np.ascontiguousarray(l["bias"], dtype=np.float16)
            keep += [k, b]
            arr[i] = _lib.NifLayer(k.shape[0], k.shape[1], int(bool(l["relu"])), 0, k.ctypes.data, None if b is None else b.ctypes.data)
        desc = _lib.NifDesc(len(model.layers), model.embedding_dimension, model.max, (C.c_float * 3)(*model.mean),
                            int(model.log_tone_map), arr)
        self._check(self.lib.b200pt_upload_nif(self.h, C.byref(desc)), "init_nif_weights")

    def init_render_settings(self, seed: int = 1, aa_noise_scale: float = 0.3, fov_degrees: float = 90.0,
                             env_map_rotation_degrees: float = 0.0, samples_per_step: int = 1):
        # degrees -> radians exactly as PathTracerApp.cpp:574,584 do it in fp32
        fov = float(np.float32(fov_degrees) * np.float32(np.pi / 180.0))
        az = float(np.float32((env_map_rotation_degrees / 360.0) * (2.0 * math.pi)))
        self._check(self.lib.b200pt_set_render_settings(self.h, seed, aa_noise_scale, fov, az, samples_per_step),
                    "init_render_settings")

    def set_scene(self, prims):
        """Scene as data (b200pt_set_scene): list of dicts with shape ('sphere'|'disc'), bsdf ('diffuse'|'specular'|
        'refractive'), centre, normal (disc), radius, colour, emission; None / [] restores the reference's built-in scene."""
        shapes = {"sphere": 0, "disc": 1}
        bsdfs = {"diffuse": 0, "specular": 1, "refractive": 2}
        prims = list(prims or [])
        arr = (_lib.Primitive * max(1, len(prims)))()
        for i, p in enumerate(prims):
            arr[i].shape = shapes[p["shape"]]; arr[i].material = bsdfs[p.get("bsdf", "diffuse")]
            arr[i].centre = (C.c_float * 3)(*[float(x) for x in p["centre"]])
            arr[i].normal = (C.c_float * 3)(*[float(x) for x in p.get("normal", (0, 1, 0))])
            arr[i].radius = float(p["radius"])
            arr[i].colour = (C.c_float * 3)(*[float(x) for x in p.get("colour", (1, 1, 1))])
            arr[i].emission = (C.c_float * 3)(*[float(x) for x in p.get("emission", (0, 0, 0))])
        self._check(self.lib.b200pt_set_scene(self.h, arr, len(prims)), "set_scene")

    def setup(self, worklist: np.ndarray):
        wl = np.ascontiguousarray(worklist, dtype=TRACE_RECORD)
        self._check(self.lib.b200pt_setup(self.h, _ptr(wl), len(wl)), "setup")
        self.n_items = len(wl)

    def setup_next(self, worklist: np.ndarray):
        """Stage the NEXT step's work list while a path_trace may be in flight (the reference's double-buffered work list,
        PathTracerApp.cpp:703-709); it becomes active at the next path_trace. `worklist` must stay alive and unmodified until
        then - a reference is kept here."""
        wl = np.ascontiguousarray(worklist, dtype=TRACE_RECORD)
        self._staged = wl
        self._check(self.lib.b200pt_setup_next(self.h, _ptr(wl), len(wl)), "setup_next")
        self._staged_items = len(wl)

    def path_trace(self, sync: bool = True):
        if getattr(self, "_staged", None) is not None:
            self.n_items = self._staged_items
        self._check(self.lib.b200pt_path_trace(self.h), "path_trace")
        self._staged = None
        if sync:
            self.sync()

    def sync(self):
        self._check(self.lib.b200pt_sync(self.h), "sync")

    def read_results(self, out: np.ndarray | None = None):
        if out is None:
            out = np.zeros(self.n_items, dtype=TRACE_RECORD)
        ctr = _lib.Counters()
        self._check(self.lib.b200pt_read_results(self.h, _ptr(out), len(out), C.byref(ctr)), "read_results")
        return out, self._counters(ctr)

    # ---- control / film ----------------------------------------------------------------------
    @staticmethod
    def _counters(c) -> Counters:
        return Counters(c.paths, c.escaped, c.path_segments, c.trace_ms, c.nif_ms, c.total_ms, c.kernel_launches, c.nif_impl_used)

    def counters(self) -> Counters:
        ctr = _lib.Counters()
        self._check(self.lib.b200pt_get_counters(self.h, C.byref(ctr)), "get_counters")
        return self._counters(ctr)

    def set_sample_index(self, next_sample: int):
        self._check(self.lib.b200pt_set_sample_index(self.h, next_sample), "set_sample_index")

    def nif_flops_per_eval(self) -> int:
        return int(self.lib.b200pt_nif_flops_per_eval(self.h))

    def film_accumulate(self):
        self._check(self.lib.b200pt_film_accumulate(self.h), "film_accumulate")

    def film_clear(self):
        self._check(self.lib.b200pt_film_clear(self.h), "film_clear")

    def film_read(self, scale: float = 1.0) -> np.ndarray:
        out = np.zeros((self.height, self.width, 3), np.float32)
        self._check(self.lib.b200pt_film_read(self.h, _ptr(out), scale), "film_read")
        return out

    def film_read_device(self, device_ptr: int, scale: float = 1.0):
        self._check(self.lib.b200pt_film_read_device(self.h, C.c_void_p(device_ptr), scale), "film_read_device")

    def film_tonemap(self, scale: float, exposure: float = 0.0, gamma: float = 2.2) -> np.ndarray:
        out = np.zeros((self.height, self.width, 3), np.uint8)
        self._check(self.lib.b200pt_film_tonemap(self.h, _ptr(out), scale, exposure, gamma), "film_tonemap")
        return out

    # ---- stage-level entry points ------------------------------------------------------------
    def nif_infer(self, u, v, impl: int = NIF_AUTO) -> np.ndarray:
        u = np.ascontiguousarray(u, np.float32); v = np.ascontiguousarray(v, np.float32)
        out = np.zeros((u.size, 3), np.float32)
        self._check(self.lib.b200pt_nif_infer(self.h, u.size, _ptr(u), _ptr(v), _ptr(out), impl), "nif_infer")
        return out

    def nif_infer_device(self, n: int, u_ptr: int, v_ptr: int, out_ptr: int, impl: int = NIF_AUTO, repeats: int = 1) -> float:
        ms = C.c_double(0)
        self._check(self.lib.b200pt_nif_infer_device(self.h, n, C.c_void_p(u_ptr), C.c_void_p(v_ptr), C.c_void_p(out_ptr),
                                                     impl, repeats, C.byref(ms)), "nif_infer_device")
        return ms.value

    def trace_wavefront(self, sample: int, n: int | None = None):
        n = self.n_items if n is None else n
        esc = np.zeros(n, np.int32); uv = np.zeros((n, 2), np.float32); thr = np.zeros((n, 3), np.float32)
        length = np.zeros(n, np.int32); cam = np.zeros((n, 2), np.uint16)
        self._check(self.lib.b200pt_trace_wavefront(self.h, sample, n, _ptr(esc), _ptr(uv), _ptr(thr), _ptr(length), _ptr(cam)),
                    "trace_wavefront")
        return esc, uv, thr, length, cam

    def trace_fixed(self, cam_rays_bits: np.ndarray, uniform_bits: np.ndarray, per_path: int):
        cam = np.ascontiguousarray(cam_rays_bits, np.uint16); uni = np.ascontiguousarray(uniform_bits, np.uint16)
        n = cam.size // 2
        assert uni.size >= n * per_path
        esc = np.zeros(n, np.int32); d = np.zeros((n, 3), np.float32); thr = np.zeros((n, 3), np.float32)
        length = np.zeros(n, np.int32)
        self._check(self.lib.b200pt_trace_fixed(self.h, n, _ptr(cam), _ptr(uni), per_path, _ptr(esc), _ptr(d), _ptr(thr), _ptr(length)),
                    "trace_fixed")
        return esc, d, thr, length

    def kat(self, op: int, rows: np.ndarray, out_cols: int) -> np.ndarray:
        rows = np.ascontiguousarray(rows, np.float32)
        out = np.zeros((rows.shape[0], out_cols), np.float32)
        self._check(self.lib.b200pt_kat(self.h, op, rows.shape[0], _ptr(rows), rows.shape[1], _ptr(out), out_cols), "kat")
        return out

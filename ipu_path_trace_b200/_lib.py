"""ctypes binding of libb200pt.so (include/b200pt.h). There is no fallback: if the CUDA
library is missing or cannot be loaded, importing a symbol from here raises."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200PT_LIB") or os.path.join(HERE, "libb200pt.so")

# Every symbol include/b200pt.h declares (tests check the library exports each one).
EXPORTS = [
    "b200pt_create", "b200pt_destroy", "b200pt_last_error", "b200pt_version",
    "b200pt_upload_nif", "b200pt_set_render_settings", "b200pt_setup", "b200pt_setup_next", "b200pt_host_alloc", "b200pt_host_free", "b200pt_set_scene",
    "b200pt_path_trace", "b200pt_sync",
    "b200pt_read_results", "b200pt_film_accumulate", "b200pt_film_clear", "b200pt_film_read",
    "b200pt_film_read_device", "b200pt_film_tonemap", "b200pt_set_sample_index", "b200pt_get_counters",
    "b200pt_nif_flops_per_eval", "b200pt_nif_infer", "b200pt_nif_infer_device", "b200pt_trace_wavefront",
    "b200pt_trace_fixed", "b200pt_kat",
    "b200pt_cornell_render", "b200pt_cornell_trace_paths", "b200pt_cornell_last_error", "b200pt_cornell_render_noalloc",
]


class Primitive(C.Structure):
    _fields_ = [("shape", C.c_int32), ("material", C.c_int32), ("centre", C.c_float * 3), ("normal", C.c_float * 3),
                ("radius", C.c_float), ("colour", C.c_float * 3), ("emission", C.c_float * 3)]


class Config(C.Structure):
    _fields_ = [
        ("struct_size", C.c_uint32), ("device", C.c_int32), ("width", C.c_uint32), ("height", C.c_uint32),
        ("max_path_length", C.c_uint32), ("roulette_depth", C.c_uint32), ("stop_prob", C.c_float),
        ("refractive_index", C.c_float), ("aa_noise_type", C.c_int32), ("max_worklist", C.c_uint32),
        ("samples_per_launch", C.c_uint32), ("nif_impl", C.c_uint32),
    ]


class NifLayer(C.Structure):
    _fields_ = [("in_", C.c_int32), ("out", C.c_int32), ("relu", C.c_int32), ("dtype", C.c_int32),
                ("kernel", C.c_void_p), ("bias", C.c_void_p)]


class NifDesc(C.Structure):
    _fields_ = [("num_layers", C.c_int32), ("embedding_dimension", C.c_int32), ("max", C.c_float),
                ("mean", C.c_float * 3), ("log_tone_map", C.c_int32), ("layers", C.POINTER(NifLayer))]


class Counters(C.Structure):
    _fields_ = [("paths", C.c_uint64), ("escaped", C.c_uint64), ("path_segments", C.c_uint64),
                ("trace_ms", C.c_double), ("nif_ms", C.c_double), ("total_ms", C.c_double),
                ("kernel_launches", C.c_uint64), ("nif_impl_used", C.c_int32)]


class CornellConfig(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("device", C.c_int32), ("width", C.c_uint32), ("height", C.c_uint32),
                ("tile_width", C.c_uint32), ("tile_height", C.c_uint32), ("samples", C.c_uint32), ("roulette_depth", C.c_uint32),
                ("seed", C.c_uint64), ("aa_noise_scale", C.c_float), ("refractive_index", C.c_float), ("stop_prob", C.c_float),
                ("epsilon", C.c_float), ("intersection_epsilon", C.c_float), ("rng_mode", C.c_int32)]


class CornellStats(C.Structure):
    _fields_ = [("total_ray_casts", C.c_uint64), ("max_path_length", C.c_uint64), ("device_ms", C.c_double)]


def build(verbose: bool = False) -> str:
    """Compile libb200pt.so in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    cmd = ["make", "-C", os.path.join(HERE, "csrc"), "-j8", "CCBIN=/usr/bin/g++"]
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if verbose or res.returncode != 0:
        print(res.stdout)
    if res.returncode != 0:
        raise RuntimeError("building libb200pt.so failed")
    return LIB_PATH


_lib = None


def load() -> C.CDLL:
    """Load the CUDA library; raises if it has not been built (no CPU path exists)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with ipu_path_trace_b200.build() / __graft_entry__.build(). "
            "This package has no CPU or PyTorch fallback.")
    lib = C.CDLL(LIB_PATH)
    vp, u32, i32, f32, sz = C.c_void_p, C.c_uint32, C.c_int32, C.c_float, C.c_size_t
    lib.b200pt_create.argtypes = [C.POINTER(Config), C.POINTER(vp)]
    lib.b200pt_destroy.argtypes = [vp]
    lib.b200pt_destroy.restype = None
    lib.b200pt_last_error.argtypes = [vp]
    lib.b200pt_last_error.restype = C.c_char_p
    lib.b200pt_version.restype = C.c_char_p
    lib.b200pt_upload_nif.argtypes = [vp, C.POINTER(NifDesc)]
    lib.b200pt_set_render_settings.argtypes = [vp, C.c_uint64, f32, f32, f32, u32]
    lib.b200pt_setup.argtypes = [vp, vp, sz]
    lib.b200pt_set_scene.argtypes = [vp, C.POINTER(Primitive), u32]
    lib.b200pt_setup_next.argtypes = [vp, vp, sz]
    lib.b200pt_host_alloc.argtypes = [C.POINTER(vp), sz]
    lib.b200pt_host_free.argtypes = [vp]
    lib.b200pt_path_trace.argtypes = [vp]
    lib.b200pt_sync.argtypes = [vp]
    lib.b200pt_read_results.argtypes = [vp, vp, sz, C.POINTER(Counters)]
    lib.b200pt_film_accumulate.argtypes = [vp]
    lib.b200pt_film_clear.argtypes = [vp]
    lib.b200pt_film_read.argtypes = [vp, vp, f32]
    lib.b200pt_film_read_device.argtypes = [vp, vp, f32]
    lib.b200pt_film_tonemap.argtypes = [vp, vp, f32, f32, f32]
    lib.b200pt_set_sample_index.argtypes = [vp, u32]
    lib.b200pt_get_counters.argtypes = [vp, C.POINTER(Counters)]
    lib.b200pt_nif_flops_per_eval.argtypes = [vp]
    lib.b200pt_nif_flops_per_eval.restype = C.c_int64
    lib.b200pt_nif_infer.argtypes = [vp, sz, vp, vp, vp, i32]
    lib.b200pt_nif_infer_device.argtypes = [vp, sz, vp, vp, vp, i32, i32, C.POINTER(C.c_double)]
    lib.b200pt_trace_wavefront.argtypes = [vp, u32, sz, vp, vp, vp, vp, vp]
    lib.b200pt_trace_fixed.argtypes = [vp, sz, vp, vp, u32, vp, vp, vp, vp]
    lib.b200pt_kat.argtypes = [vp, i32, sz, vp, i32, vp, i32]
    lib.b200pt_cornell_render.argtypes = [C.POINTER(CornellConfig), vp, C.POINTER(CornellStats)]
    lib.b200pt_cornell_trace_paths.argtypes = [C.POINTER(CornellConfig), sz, vp, vp, vp, vp]
    lib.b200pt_cornell_last_error.restype = C.c_char_p
    lib.b200pt_cornell_render_noalloc.argtypes = [C.POINTER(CornellConfig), vp, C.POINTER(CornellStats)]
    _lib = lib
    return lib

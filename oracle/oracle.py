"""TEST INFRASTRUCTURE ONLY — ctypes front-end for the CPU oracle.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this module. The product package
(``ipu_path_trace_b200``) never does; it fails loudly without its CUDA library.

Two libraries:
  * ``oracle/liboracle.so``            our CPU restatement (pt_oracle.cpp, nif_oracle.cpp, render_oracle.cpp)
  * ``oracle/_ref/libref_codelets.so`` the UNMODIFIED reference codelets compiled from
    /root/reference through the Poplar shim (ref_driver.cpp); present when it was built
    in the container that has the reference tree. It pins the restatement.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass, field

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

TRACE_RECORD = np.dtype(
    [("u", "<u2"), ("v", "<u2"), ("r", "<f4"), ("g", "<f4"), ("b", "<f4"), ("sampleCount", "<u2"), ("pathLength", "<u2")]
)
CONTRIB = np.dtype([("clr", "<f4", (3,)), ("weight", "<f4"), ("type", "u1"), ("pad", "u1", (3,))])
assert TRACE_RECORD.itemsize == 20 and CONTRIB.itemsize == 20

C_DIFFUSE, C_EMIT, C_SPECULAR, C_REFRACT, C_ESCAPED, C_DEBUG, C_END, C_SKIP = range(8)
AA_NORMAL, AA_UNIFORM, AA_TRUNCATED_NORMAL = 0, 1, 2


def build(ref: bool = True) -> None:
    """Compile the oracle (and, where /root/reference exists, oracle/_ref)."""
    subprocess.run(["make", "-s", "-C", HERE, "all"], check=True)
    if ref and os.path.isdir("/root/reference"):
        subprocess.run(["make", "-s", "-C", HERE, "ref"], check=True)


def half_round(x):
    """float32 -> nearest binary16 -> float32 (numpy's conversion is IEEE RNE)."""
    return np.asarray(x, dtype=np.float32).astype(np.float16).astype(np.float32)


def half_bits(x) -> np.ndarray:
    return np.asarray(x, dtype=np.float32).astype(np.float16).view(np.uint16)


class _Settings(C.Structure):
    _fields_ = [
        ("width", C.c_uint32), ("height", C.c_uint32),
        ("fov", C.c_float), ("aaScale", C.c_float), ("azimuth", C.c_float),
        ("refractiveIndex", C.c_float), ("stopProb", C.c_float),
        ("rouletteDepth", C.c_uint32), ("maxPathLength", C.c_uint32),
        ("seed", C.c_uint64), ("aaNoiseType", C.c_int32),
    ]


@dataclass
class Settings:
    """Render settings with the reference's defaults (PathTracerApp.cpp:797-829).

    Values the reference streams to the device as ``half`` are rounded through binary16
    here, once, so that the oracle and the CUDA library see identical numbers."""
    width: int = 256
    height: int = 256
    fov_degrees: float = 90.0
    aa_noise_scale: float = 0.3
    env_map_rotation_degrees: float = 0.0
    refractive_index: float = 1.5
    stop_prob: float = 0.3
    roulette_depth: int = 3
    max_path_length: int = 10
    seed: int = 1
    aa_noise_type: int = AA_NORMAL

    def to_c(self) -> _Settings:
        fov = np.float32(self.fov_degrees) * np.float32(np.pi / 180.0)   # PathTracerApp.cpp:574
        az = np.float32((self.env_map_rotation_degrees / 360.0) * (2.0 * np.pi))  # :584
        return _Settings(
            self.width, self.height,
            float(half_round(fov)), float(half_round(self.aa_noise_scale)), float(az),
            float(half_round(self.refractive_index)), float(half_round(self.stop_prob)),
            self.roulette_depth, self.max_path_length, self.seed, self.aa_noise_type,
        )


class _NifLayer(C.Structure):
    _fields_ = [("in_", C.c_int32), ("out", C.c_int32), ("relu", C.c_int32), ("hasBias", C.c_int32),
                ("kernel", C.c_void_p), ("bias", C.c_void_p)]


class _Nif(C.Structure):
    _fields_ = [("numLayers", C.c_int32), ("embeddingDim", C.c_int32), ("maxv", C.c_float),
                ("mean", C.c_float * 3), ("logToneMap", C.c_int32), ("layers", C.POINTER(_NifLayer))]


@dataclass
class NifLayers:
    """Host-side NIF description: list of (kernel[in,out] fp16, bias[out] fp16|None, relu)."""
    embedding_dim: int
    maxv: float
    mean: tuple            # eps already folded in (NifMetaData.cpp:48-53)
    log_tone_map: bool
    layers: list = field(default_factory=list)   # dicts: kernel, bias, relu

    def to_c(self):
        keep = []
        arr = (_NifLayer * len(self.layers))()
        for i, l in enumerate(self.layers):
            k = np.ascontiguousarray(l["kernel"], dtype=np.float16)
            keep.append(k)
            b = None
            if l.get("bias") is not None:
                b = np.ascontiguousarray(l["bias"], dtype=np.float16)
                keep.append(b)
            arr[i] = _NifLayer(k.shape[0], k.shape[1], int(bool(l["relu"])), int(b is not None),
                               k.ctypes.data, b.ctypes.data if b is not None else None)
        nif = _Nif(len(self.layers), self.embedding_dim, float(self.maxv),
                   (C.c_float * 3)(*[float(m) for m in self.mean]), int(self.log_tone_map), arr)
        keep.append(arr)
        return nif, keep


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


class Oracle:
    """The CPU restatement (liboracle.so)."""

    def __init__(self):
        path = os.path.join(HERE, "liboracle.so")
        if not os.path.exists(path):
            build(ref=False)
        self.lib = C.CDLL(path)
        self.lib.orc_render.restype = C.c_longlong
        self.lib.orc_nif_flops_per_eval.restype = C.c_longlong
        assert self.lib.orc_sizeof_settings() == C.sizeof(_Settings)
        assert self.lib.orc_sizeof_nif() == C.sizeof(_Nif)
        assert self.lib.orc_sizeof_nif_layer() == C.sizeof(_NifLayer)

    def num_threads(self) -> int:
        return int(self.lib.orc_num_threads())

    def use_all_host_threads(self) -> int:
        """Lift an inherited OMP_NUM_THREADS=1 (torch.distributed.run sets it): the CPU arm runs on every
        host thread this process may use. Returns the thread count now in force."""
        n = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
        self.lib.orc_set_num_threads(int(n))
        return self.num_threads()

    def camera_rays(self, items: np.ndarray, noise_bits: np.ndarray, s: Settings) -> np.ndarray:
        n = len(items)
        items = np.ascontiguousarray(items, dtype=TRACE_RECORD)
        noise_bits = np.ascontiguousarray(noise_bits, dtype=np.uint16)
        assert noise_bits.size >= 2 * n
        rays = np.zeros(2 * n, dtype=np.uint16)
        cs = s.to_c()
        self.lib.orc_camera_rays(n, _ptr(items), _ptr(noise_bits), C.byref(cs), _ptr(rays))
        return rays

    def trace_records(self, ray_bits: np.ndarray, uniform_bits: np.ndarray, per_path: int, s: Settings) -> np.ndarray:
        n = ray_bits.size // 2
        ray_bits = np.ascontiguousarray(ray_bits, dtype=np.uint16)
        uniform_bits = np.ascontiguousarray(uniform_bits, dtype=np.uint16)
        assert uniform_bits.size >= n * per_path
        rec = np.zeros((n, s.max_path_length), dtype=CONTRIB)
        cs = s.to_c()
        self.lib.orc_trace_records(n, _ptr(ray_bits), _ptr(uniform_bits), per_path, C.byref(cs), _ptr(rec))
        return rec

    def escaped_uv(self, records: np.ndarray, azimuth: float):
        n, L = records.shape
        records = np.ascontiguousarray(records)
        u = np.zeros(n, np.float32); v = np.zeros(n, np.float32)
        self.lib.orc_escaped_uv(n, _ptr(records), L, C.c_float(azimuth), _ptr(u), _ptr(v))
        return u, v

    def light_and_accumulate(self, records: np.ndarray, bgr: np.ndarray, items: np.ndarray):
        n, L = records.shape
        records = np.ascontiguousarray(records).copy()
        bgr = np.ascontiguousarray(bgr, dtype=np.float32)
        items = np.ascontiguousarray(items, dtype=TRACE_RECORD).copy()
        self.lib.orc_light_and_accumulate(n, _ptr(records), L, _ptr(bgr), _ptr(items))
        return items

    def path_summary(self, records: np.ndarray):
        n, L = records.shape
        records = np.ascontiguousarray(records)
        esc = np.zeros(n, np.int32); length = np.zeros(n, np.int32)
        d = np.zeros((n, 3), np.float32); t = np.zeros((n, 3), np.float32)
        self.lib.orc_path_summary(n, _ptr(records), L, _ptr(esc), _ptr(d), _ptr(t), _ptr(length))
        return esc, d, t, length

    def trace_philox(self, items: np.ndarray, sample: int, s: Settings):
        assert s.max_path_length <= 64
        n = len(items)
        items = np.ascontiguousarray(items, dtype=TRACE_RECORD)
        esc = np.zeros(n, np.int32); length = np.zeros(n, np.int32)
        uv = np.zeros((n, 2), np.float32); t = np.zeros((n, 3), np.float32)
        cam = np.zeros((n, 2), np.uint16)
        cs = s.to_c()
        self.lib.orc_trace_philox(n, _ptr(items), C.c_uint32(sample), C.byref(cs), _ptr(esc), _ptr(uv), _ptr(t), _ptr(length), _ptr(cam))
        return esc, uv, t, length, cam

    def set_scene(self, prims) -> None:
        """Scene as data (the reference's comment at codelets.cpp:90-92): list of dicts with shape ('sphere'|'disc'),
        bsdf ('diffuse'|'specular'|'refractive'), centre, normal (disc), radius, colour, emission. None / [] restores the
        built-in scene."""
        rows = scene_rows(prims)
        self.lib.orc_set_scene(len(rows), _ptr(rows) if len(rows) else None)

    def philox(self, c, k) -> np.ndarray:
        out = np.zeros(4, np.uint32)
        self.lib.orc_philox(*[C.c_uint32(int(x)) for x in c], *[C.c_uint32(int(x)) for x in k], _ptr(out))
        return out

    def aa_noise(self, words: np.ndarray, noise_type: int) -> np.ndarray:
        words = np.ascontiguousarray(words, dtype=np.uint32)
        n = words.size // 2
        out = np.zeros(2 * n, np.uint16)
        self.lib.orc_aa_noise(n, _ptr(words), noise_type, _ptr(out))
        return out

    def float_to_half(self, x: np.ndarray) -> np.ndarray:
        x = np.ascontiguousarray(x, dtype=np.float32)
        out = np.zeros(x.size, np.uint16)
        self.lib.orc_float_to_half(x.size, _ptr(x), _ptr(out))
        return out.reshape(x.shape)

    def half_to_float(self, x: np.ndarray) -> np.ndarray:
        x = np.ascontiguousarray(x, dtype=np.uint16)
        out = np.zeros(x.size, np.float32)
        self.lib.orc_half_to_float(x.size, _ptr(x), _ptr(out))
        return out.reshape(x.shape)

    def nif_encode(self, u, v, embedding_dim: int) -> np.ndarray:
        u = np.ascontiguousarray(u, dtype=np.float32); v = np.ascontiguousarray(v, dtype=np.float32)
        n = u.size
        feat = np.zeros((n, 4 * embedding_dim), np.uint16)
        self.lib.orc_nif_encode(n, _ptr(u), _ptr(v), embedding_dim, _ptr(feat))
        return feat.view(np.float16)

    def nif_forward(self, nif: NifLayers, u, v, mode: int = 0) -> np.ndarray:
        u = np.ascontiguousarray(u, dtype=np.float32); v = np.ascontiguousarray(v, dtype=np.float32)
        n = u.size
        out = np.zeros((n, 3), np.float32)
        cn, keep = nif.to_c()
        rc = self.lib.orc_nif_forward(C.byref(cn), n, _ptr(u), _ptr(v), mode, _ptr(out))
        assert rc == 0
        return out

    def nif_decode(self, x3, maxv: float, mean3, log_tone_map: bool) -> np.ndarray:
        x3 = np.ascontiguousarray(x3, dtype=np.float32)
        out = np.zeros_like(x3)
        mean = np.ascontiguousarray(mean3, dtype=np.float32)
        self.lib.orc_nif_decode(x3.shape[0], _ptr(x3), C.c_float(maxv), _ptr(mean), int(bool(log_tone_map)), _ptr(out))
        return out

    def nif_flops_per_eval(self, nif: NifLayers) -> int:
        cn, keep = nif.to_c()
        return int(self.lib.orc_nif_flops_per_eval(C.byref(cn)))

    def render(self, items: np.ndarray, sample_begin: int, sample_count: int, s: Settings, nif: NifLayers, mode: int = 0):
        """Runs `sample_count` iterations in place on a copy of `items`.
        Returns (items, escaped_paths, seconds_trace, seconds_nif)."""
        assert s.max_path_length <= 64
        items = np.ascontiguousarray(items, dtype=TRACE_RECORD).copy()
        cs = s.to_c()
        cn, keep = nif.to_c()
        t_tr = C.c_double(0); t_nif = C.c_double(0)
        esc = self.lib.orc_render(len(items), _ptr(items), C.c_uint32(sample_begin), C.c_uint32(sample_count),
                                  C.byref(cs), C.byref(cn), mode, C.byref(t_tr), C.byref(t_nif))
        return items, int(esc), t_tr.value, t_nif.value


class Reference:
    """The unmodified reference codelets (oracle/_ref/libref_codelets.so)."""

    @staticmethod
    def available() -> bool:
        return os.path.exists(os.path.join(HERE, "_ref", "libref_codelets.so"))

    def __init__(self):
        self.lib = C.CDLL(os.path.join(HERE, "_ref", "libref_codelets.so"))
        assert self.lib.ref_sizeof_trace_record() == 20 and self.lib.ref_sizeof_contribution() == 20

    @staticmethod
    def worker_noise_layout(noise_pairs: np.ndarray) -> np.ndarray:
        """Lay out per-item noise pairs (n,2) the way GenerateCameraRays' 6 workers read them
        (codelets.cpp:53-58,69-76): item j=w+6m reads noise[w+12m] and noise[w+12m+6]."""
        n = noise_pairs.shape[0]
        m = (n + 5) // 6
        buf = np.zeros(12 * m, dtype=np.uint16)
        j = np.arange(n)
        buf[(j % 6) + 12 * (j // 6)] = noise_pairs[:, 0]
        buf[(j % 6) + 12 * (j // 6) + 6] = noise_pairs[:, 1]
        return buf

    def camera_rays(self, items: np.ndarray, noise_pairs_bits: np.ndarray, s: Settings) -> np.ndarray:
        n = len(items)
        items = np.ascontiguousarray(items, dtype=TRACE_RECORD)
        noise = self.worker_noise_layout(np.ascontiguousarray(noise_pairs_bits, dtype=np.uint16).reshape(n, 2))
        rays = np.zeros(2 * n, np.uint16)
        cs = s.to_c()
        self.lib.ref_generate_camera_rays(n, _ptr(items), _ptr(noise), noise.size, _ptr(rays), cs.width, cs.height,
                                          C.c_uint16(int(half_bits(cs.aaScale))), C.c_uint16(int(half_bits(cs.fov))))
        return rays

    def ray_trace(self, ray_bits: np.ndarray, uniform_bits: np.ndarray, s: Settings, per_path: int = 0) -> np.ndarray:
        n = ray_bits.size // 2
        ray_bits = np.ascontiguousarray(ray_bits, dtype=np.uint16)
        uniform_bits = np.ascontiguousarray(uniform_bits, dtype=np.uint16)
        rec = np.zeros((n, s.max_path_length), dtype=CONTRIB)
        cs = s.to_c()
        self.lib.ref_ray_trace(n, _ptr(ray_bits), _ptr(uniform_bits), uniform_bits.size, int(per_path > 0), per_path,
                               _ptr(rec), s.max_path_length, C.c_uint16(int(half_bits(cs.refractiveIndex))),
                               C.c_uint16(int(half_bits(cs.stopProb))), C.c_ushort(s.roulette_depth))
        return rec

    def preprocess_escaped(self, records: np.ndarray, azimuth: float):
        n, L = records.shape
        records = np.ascontiguousarray(records).copy()
        u = np.zeros(n, np.float32); v = np.zeros(n, np.float32)
        self.lib.ref_preprocess_escaped(n, _ptr(records), L, C.c_float(azimuth), _ptr(u), _ptr(v))
        return u, v, records

    def postprocess_and_accumulate(self, records: np.ndarray, bgr: np.ndarray, items: np.ndarray):
        n, L = records.shape
        records = np.ascontiguousarray(records).copy()
        bgr = np.ascontiguousarray(bgr, dtype=np.float32).copy()
        items = np.ascontiguousarray(items, dtype=TRACE_RECORD).copy()
        self.lib.ref_postprocess_escaped(n, _ptr(records), L, _ptr(bgr))
        self.lib.ref_accumulate(n, _ptr(records), L, _ptr(items))
        return items

    # function-level KATs
    def pixel_to_ray(self, x, y, w, h, fov):
        x = np.ascontiguousarray(x, np.float32); y = np.ascontiguousarray(y, np.float32)
        out = np.zeros((x.size, 3), np.float32)
        self.lib.ref_pixel_to_ray(x.size, _ptr(x), _ptr(y), C.c_uint(w), C.c_uint(h), C.c_float(fov), _ptr(out))
        return out

    def sphere_intersect(self, origin, direction, centre, radius):
        o = np.ascontiguousarray(origin, np.float32); d = np.ascontiguousarray(direction, np.float32)
        c = np.ascontiguousarray(centre, np.float32); r = np.ascontiguousarray(radius, np.float32)
        t = np.zeros(r.size, np.float32)
        self.lib.ref_sphere_intersect(r.size, _ptr(o), _ptr(d), _ptr(c), _ptr(r), _ptr(t))
        return t

    def disc_intersect(self, origin, direction, normal, centre, radius):
        o = np.ascontiguousarray(origin, np.float32); d = np.ascontiguousarray(direction, np.float32)
        nn = np.ascontiguousarray(normal, np.float32)
        c = np.ascontiguousarray(centre, np.float32); r = np.ascontiguousarray(radius, np.float32)
        t = np.zeros(r.size, np.float32)
        self.lib.ref_disc_intersect(r.size, _ptr(o), _ptr(d), _ptr(nn), _ptr(c), _ptr(r), _ptr(t))
        return t

    def diffuse(self, normal, u1, u2, rr):
        nn = np.ascontiguousarray(normal, np.float32)
        u1 = np.ascontiguousarray(u1, np.float32); u2 = np.ascontiguousarray(u2, np.float32)
        d = np.zeros((u1.size, 3), np.float32); w = np.zeros(u1.size, np.float32)
        self.lib.ref_diffuse(u1.size, _ptr(nn), _ptr(u1), _ptr(u2), C.c_float(rr), _ptr(d), _ptr(w))
        return d, w

    def reflect(self, direction, normal):
        d = np.ascontiguousarray(direction, np.float32); nn = np.ascontiguousarray(normal, np.float32)
        out = np.zeros_like(d)
        self.lib.ref_reflect(d.shape[0], _ptr(d), _ptr(nn), _ptr(out))
        return out

    def refract(self, direction, normal, ri, rnd):
        d = np.ascontiguousarray(direction, np.float32); nn = np.ascontiguousarray(normal, np.float32)
        rnd = np.ascontiguousarray(rnd, np.float32)
        out = np.zeros_like(d); flag = np.zeros(d.shape[0], np.int32)
        self.lib.ref_refract(d.shape[0], _ptr(d), _ptr(nn), C.c_float(ri), _ptr(rnd), _ptr(out), _ptr(flag))
        return out, flag

    def roulette(self, rnd, stop_prob):
        rnd = np.ascontiguousarray(rnd, np.float32)
        stop = np.zeros(rnd.size, np.int32); w = np.zeros(rnd.size, np.float32)
        self.lib.ref_roulette(rnd.size, _ptr(rnd), C.c_float(stop_prob), _ptr(stop), _ptr(w))
        return stop, w


def scene_rows(prims) -> np.ndarray:
    """-> (n, 15) float32: shape, bsdf, centre[3], normal[3], radius, colour[3], emission[3]."""
    shapes = {"sphere": 0, "disc": 1}
    bsdfs = {"diffuse": 0, "specular": 1, "refractive": 2}
    rows = np.zeros((len(prims or []), 15), np.float32)
    for i, p in enumerate(prims or []):
        rows[i, 0] = shapes[p["shape"]]; rows[i, 1] = bsdfs[p.get("bsdf", "diffuse")]
        rows[i, 2:5] = p["centre"]; rows[i, 5:8] = p.get("normal", (0, 1, 0)); rows[i, 8] = p["radius"]
        rows[i, 9:12] = p.get("colour", (1, 1, 1)); rows[i, 12:15] = p.get("emission", (0, 0, 0))
    return rows


def reference_scene(with_light: bool = False) -> list:
    """The scene of codelets.cpp:110-144 as data; with_light adds the emitter the reference has commented out
    (:117 Sphere(Vec(12, 8, -5.3), .3), :142 Object(zero colour, lightW = 10000, diffuse))."""
    g = 2.0
    s = [dict(shape="sphere", centre=(-1.8575, -0.98714, -3.6), radius=0.6, colour=(1 * g, .89 * g, .55 * g), bsdf="diffuse"),
         dict(shape="sphere", centre=(0.74795, -0.55, -4.3816), radius=1.05, colour=(1, 1, 1), bsdf="specular"),
         dict(shape="sphere", centre=(1.9929, -1.08666, -3.23), radius=0.5, colour=(.75, .75, .75), bsdf="refractive"),
         dict(shape="sphere", centre=(-0.19931, -1.183, -2.75), radius=0.4, colour=(.8 * g, .06 * g, .391 * g), bsdf="diffuse"),
         dict(shape="sphere", centre=(-0.19931, -1.183, -2.75), radius=0.4001, colour=(1, 1, 1), bsdf="refractive"),
         dict(shape="disc", centre=(0, -1.6, -5.22), normal=(0, 1, 0), radius=3.5, colour=(.98 * g, .76 * g, .66 * g), bsdf="diffuse")]
    if with_light:
        s.insert(5, dict(shape="sphere", centre=(12, 8, -5.3), radius=0.3, colour=(0, 0, 0), emission=(10000, 10000, 10000), bsdf="diffuse"))
    return s


def make_worklist(width: int, height: int) -> np.ndarray:
    """Row-major list of every pixel (LoadBalancer.cpp:38-52)."""
    items = np.zeros(width * height, dtype=TRACE_RECORD)
    vv, uu = np.divmod(np.arange(width * height), width)
    items["u"] = uu
    items["v"] = vv
    return items


class NifTwinsReference:
    """The reference's host-side twins of the NIF's two ends - NifModel::generateInputSamples / makeCoefficients /
    makeGridCoordsUV and NifModel::decodeSamples (src/neural_networks/NifModel.cpp:403-443,466-509) - cut out of the
    reference source at build time and compiled unmodified (oracle/_ref/libref_niftwins.so). TEST INFRASTRUCTURE ONLY."""

    PATH = os.path.join(HERE, "_ref", "libref_niftwins.so")

    @staticmethod
    def available() -> bool:
        return os.path.exists(NifTwinsReference.PATH)

    def __init__(self):
        self.lib = C.CDLL(self.PATH)

    def encode_grid(self, height: int, width: int, embedding_dim: int):
        n = height * width
        f = np.zeros((n, 4 * embedding_dim), np.float32)
        u = np.zeros(n, np.float32); v = np.zeros(n, np.float32)
        assert self.lib.refnif_encode_grid(height, width, embedding_dim, _ptr(f), _ptr(u), _ptr(v)) == 0
        return f, u, v

    def decode(self, x3, maxv: float, mean3, eps: float, log_tone_map: bool) -> np.ndarray:
        x3 = np.ascontiguousarray(x3, dtype=np.float32)
        out = np.zeros_like(x3)
        mean = np.ascontiguousarray(mean3, dtype=np.float32)
        assert self.lib.refnif_decode(x3.shape[0], _ptr(x3), C.c_float(maxv), _ptr(mean), C.c_float(eps), int(bool(log_tone_map)), _ptr(out)) == 0
        return out


class CornellReference:
    """The reference's CPU tracer of the Cornell-box scene, compiled unmodified (oracle/_ref/libref_cornell.so,
    external/light/src/trace.hpp). TEST INFRASTRUCTURE ONLY."""

    PATH = os.path.join(HERE, "_ref", "libref_cornell.so")

    @staticmethod
    def available() -> bool:
        return os.path.exists(CornellReference.PATH)

    def __init__(self):
        self.lib = C.CDLL(self.PATH)

    def trace_paths(self, cam_xy, rng_state, refractive_index=1.5, roulette_depth=3, stop_prob=0.2,
                    epsilon=float(np.finfo(np.float32).eps), intersection_epsilon=1e-5):
        cam = np.ascontiguousarray(cam_xy, np.float32)
        state = np.array(rng_state, np.uint64, copy=True, order="C")
        n = cam.shape[0]
        rgb = np.zeros((n, 3), np.float32)
        depth = np.zeros(n, np.uint32)
        self.lib.ref_cornell_trace_paths(C.c_size_t(n), _ptr(cam), _ptr(state), C.c_float(refractive_index), C.c_uint32(roulette_depth),
                                         C.c_float(stop_prob), C.c_float(epsilon), C.c_float(intersection_epsilon), _ptr(rgb), _ptr(depth))
        return rgb, depth, state

    def render(self, width, height, tile_width, tile_height, samples, seed, aa_noise_scale=1.0 / 700, refractive_index=1.5,
               roulette_depth=3, stop_prob=0.2, epsilon=float(np.finfo(np.float32).eps), intersection_epsilon=1e-5):
        rgb = np.zeros((height, width, 3), np.float32)
        stats = np.zeros(2, np.uint64)
        self.lib.ref_cornell_render(C.c_uint32(width), C.c_uint32(height), C.c_uint32(tile_width), C.c_uint32(tile_height), C.c_uint32(samples),
                                    C.c_uint64(seed), C.c_float(aa_noise_scale), C.c_float(refractive_index), C.c_uint32(roulette_depth),
                                    C.c_float(stop_prob), C.c_float(epsilon), C.c_float(intersection_epsilon), _ptr(rgb), _ptr(stats))
        return rgb, {"total_ray_casts": int(stats[0]), "max_path_length": int(stats[1])}


class CornellOracle(CornellReference):
    """CPU restatement of the same renderer (oracle/cornell_oracle.cpp in liboracle.so): available everywhere, pinned bit for bit
    against CornellReference and the golden vectors by tests/test_cornell.py. TEST INFRASTRUCTURE ONLY."""

    PATH = os.path.join(HERE, "liboracle.so")

    @staticmethod
    def available() -> bool:
        return os.path.exists(CornellOracle.PATH)

    def __init__(self):
        lib = C.CDLL(self.PATH)

        class _Renamed:          # same call signatures as the compiled reference's driver, orc_ prefix
            ref_cornell_trace_paths = lib.orc_cornell_trace_paths
            ref_cornell_render = lib.orc_cornell_render
        self.lib = _Renamed

"""Host side of the neural image field (NIF): metadata, weights, synthetic models.

Mirrors what the reference's loader produces for ``init_nif_weights``:
  * NifMetaData  src/neural_networks/NifMetaData.cpp:11-71 (JSON fields, eps folded into mean :48-53,
                 hidden size scraped from train_command :56-64)
  * NifModel::Data  src/neural_networks/NifModel.cpp:40-85 (Dense layers in file order, Keras
                 [in,out] kernels, activation "linear" -> none :73-76)
The shipped weights blob (assets.extra/converted.hdf5) is missing from the reference snapshot
(.MISSING_LARGE_BLOBS:1) and no HDF5 library exists in this image, so weights travel in a trivial
container, ``nif.bin`` (see write_nif_bin); tools/h5_to_nifbin.py converts a Keras HDF5 file where
h5py is available. Tests and benchmarks use seeded synthetic weights in the documented topology.
"""
from __future__ import annotations

import json
import os
import struct
from dataclasses import dataclass, field

import numpy as np

MAGIC = b"NIFBIN01"


@dataclass
class NifMetaData:
    embedding_dimension: int
    name: str
    image_shape: tuple
    eps: float
    log_tone_map: bool
    max: float
    mean: tuple            # eps already folded in when log_tone_map (NifMetaData.cpp:48-53)
    hidden_size: int

    @staticmethod
    def load(path: str) -> "NifMetaData":
        with open(path, "r") as f:
            pt = json.load(f)
        if not pt:
            raise RuntimeError(f"Empty property tree after parsing file: '{path}'")
        try:
            enc = pt["encode_params"]
            eps = np.float32(enc["eps"])
            log_tm = bool(enc["log_tone_map"])
            mean = [np.float32(m) for m in enc["mean"]]
            if log_tm:
                mean = [np.float32(m - eps) for m in mean]
            hidden = 0
            cmd = list(pt.get("train_command", []))
            for i, tok in enumerate(cmd):
                if tok == "--layer-size" and i + 1 < len(cmd):
                    hidden = int(cmd[i + 1])
            return NifMetaData(int(pt["embedding_dimension"]), str(pt["name"]),
                               tuple(int(x) for x in pt["original_image_shape"]), float(eps), log_tm,
                               float(np.float32(enc["max"])), tuple(float(m) for m in mean), hidden)
        except KeyError as e:
            raise RuntimeError(f"Error reading property: {e} from file: '{path}'")


@dataclass
class NifModel:
    """Dense stack + decode parameters, the payload of init_nif_weights."""
    embedding_dimension: int
    max: float
    mean: tuple
    log_tone_map: bool
    layers: list = field(default_factory=list)   # dicts: kernel [in,out] fp16, bias [out] fp16 | None, relu bool

    @property
    def num_features(self) -> int:
        return 4 * self.embedding_dimension

    def flops_per_eval(self) -> int:
        """NifModel.cpp:122-134: sum of 2*in*out (+ out when biased)."""
        return sum(2 * l["kernel"].shape[0] * l["kernel"].shape[1] + (l["kernel"].shape[1] if l.get("bias") is not None else 0)
                   for l in self.layers)

    def param_count(self) -> int:
        return sum(l["kernel"].size + (l["bias"].size if l.get("bias") is not None else 0) for l in self.layers)


def synthetic_model(meta: NifMetaData | None = None, hidden: int = 320, hidden_layers: int = 6, concat_at: int = 3,
                    embedding_dimension: int = 12, seed: int = 42, outputs: int = 3) -> NifModel:
    """Seeded stand-in for the missing weights, in the topology the shipped model was trained with
    (train_command: --layer-count 6 --layer-size 320 --embedding-dimension 12 --fp16):
    4E -> hidden x hidden_layers (layer `concat_at` takes concat(activations, features)) -> 3.
    He-uniform kernels, biases U(-0.1, 0.1); the last layer is scaled so that the decoded
    log-radiance stays within a few units of the mean."""
    if meta is not None:
        embedding_dimension = meta.embedding_dimension
        if meta.hidden_size:
            hidden = meta.hidden_size
        mx, mean, log_tm = meta.max, meta.mean, meta.log_tone_map
    else:
        mx, mean, log_tm = 3.4299468994140625, (-2.3514461517333984 - 1e-8, -2.2660605907440186 - 1e-8, -1.9648972749710083 - 1e-8), True
    rng = np.random.default_rng(seed)
    nfeat = 4 * embedding_dimension
    layers = []
    width = nfeat
    for l in range(hidden_layers):
        fan_in = width + (nfeat if (l == concat_at and l > 0) else 0)
        lim = np.sqrt(6.0 / fan_in)
        layers.append({"kernel": rng.uniform(-lim, lim, size=(fan_in, hidden)).astype(np.float16),
                       "bias": rng.uniform(-0.1, 0.1, size=hidden).astype(np.float16), "relu": True})
        width = hidden
    lim = 0.5 * np.sqrt(6.0 / width)
    layers.append({"kernel": rng.uniform(-lim, lim, size=(width, outputs)).astype(np.float16),
                   "bias": rng.uniform(-0.1, 0.1, size=outputs).astype(np.float16), "relu": False})
    return NifModel(embedding_dimension, float(mx), tuple(float(m) for m in mean), bool(log_tm), layers)


def write_nif_bin(path: str, model: NifModel) -> None:
    """nif.bin: MAGIC, u32 layers, u32 E, f32 max, 3 x f32 mean, u32 log_tone_map, then per layer
    u32 in, u32 out, u32 relu, u32 has_bias, u32 dtype (0 = f16), kernel bytes [in,out], bias bytes."""
    with open(path, "wb") as f:
        f.write(MAGIC)
        f.write(struct.pack("<IIffffI", len(model.layers), model.embedding_dimension, model.max, *model.mean, int(model.log_tone_map)))
        for l in model.layers:
            k = np.ascontiguousarray(l["kernel"], dtype=np.float16)
            b = l.get("bias")
            f.write(struct.pack("<IIIII", k.shape[0], k.shape[1], int(bool(l["relu"])), int(b is not None), 0))
            f.write(k.tobytes())
            if b is not None:
                f.write(np.ascontiguousarray(b, dtype=np.float16).tobytes())


def read_nif_bin(path: str) -> NifModel:
    with open(path, "rb") as f:
        data = f.read()
    if data[:8] != MAGIC:
        raise RuntimeError(f"{path}: not a nif.bin file")
    off = 8
    n, E, mx, m0, m1, m2, ltm = struct.unpack_from("<IIffffI", data, off)
    off += 28
    layers = []
    for _ in range(n):
        i, o, relu, has_b, dt = struct.unpack_from("<IIIII", data, off)
        off += 20
        np_dt = np.float16 if dt == 0 else np.float32
        sz = np.dtype(np_dt).itemsize
        k = np.frombuffer(data, dtype=np_dt, count=i * o, offset=off).reshape(i, o).astype(np.float16)
        off += i * o * sz
        b = None
        if has_b:
            b = np.frombuffer(data, dtype=np_dt, count=o, offset=off).astype(np.float16)
            off += o * sz
        layers.append({"kernel": k, "bias": b, "relu": bool(relu)})
    return NifModel(E, mx, (m0, m1, m2), bool(ltm), layers)


def load_assets(assets_dir: str, allow_synthetic: bool = False) -> tuple[NifModel, NifMetaData, str]:
    """--assets <dir>: reads assets.extra/nif_metadata.txt (PathTracerApp.cpp:106-123) and the weights:
    assets.extra/nif.bin if present. The Keras HDF5 (converted.hdf5) needs h5py, absent here.
    Returns (model, metadata, source) where source says where the weights came from."""
    # the reference's --assets is the assets.extra directory itself (PathTracerApp.cpp:106-110); the saved
    # Keras model directory that contains it is accepted too
    if not os.path.exists(os.path.join(assets_dir, "nif_metadata.txt")) and os.path.isdir(os.path.join(assets_dir, "assets.extra")):
        assets_dir = os.path.join(assets_dir, "assets.extra")
    meta_path = os.path.join(assets_dir, "nif_metadata.txt")
    meta = NifMetaData.load(meta_path)
    bin_path = os.path.join(assets_dir, "nif.bin")
    if os.path.exists(bin_path):
        m = read_nif_bin(bin_path)
        return NifModel(meta.embedding_dimension, meta.max, meta.mean, meta.log_tone_map, m.layers), meta, bin_path
    if allow_synthetic:
        return synthetic_model(meta), meta, "synthetic(seed=42)"
    raise RuntimeError(f"no weights under {assets_dir} (nif.bin missing; converted.hdf5 is read by the C++ host, host/Hdf5File.cpp)")

"""Reads a Keras NIF model (`converted.hdf5` + `nif_metadata.txt`) into the Python NifModel THROUGH the C++ host's HDF5
subset reader (host/Hdf5File.cpp via libb200pt_host.so), walking `model_config` like src/keras/Hdf5Model.cpp:8-133.
Used by the real-weights parity tests, which run the moment the reference's weights blob is supplied. TEST HELPER."""
import ctypes as C
import json
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "host")

# Where the shipped model would be: env override, the reference tree, or a copy dropped next to the golden vectors.
CANDIDATE_ASSETS = [
    os.environ.get("B200PT_NIF_ASSETS", ""),
    "/root/reference/nif_models/urban_alley_01_4k_fp16_yuv/assets.extra",
    os.path.join(ROOT, "tests", "golden", "urban_alley_01_4k_fp16_yuv", "assets.extra"),
]


def shipped_assets():
    for d in CANDIDATE_ASSETS:
        if d and os.path.exists(os.path.join(d, "converted.hdf5")) and os.path.exists(os.path.join(d, "nif_metadata.txt")):
            return d
    return None


def host_lib():
    subprocess.run(["make", "-C", HOST, "libb200pt_host.so"], check=True, capture_output=True)
    lib = C.CDLL(os.path.join(HOST, "libb200pt_host.so"))
    lib.b200host_last_error.restype = C.c_char_p
    return lib


def _dataset(lib, h5, path):
    shape = (C.c_size_t * 4)()
    rank, es = C.c_int(4), C.c_size_t()
    buf = np.zeros(4 << 20, np.uint8)
    rc = lib.b200host_h5_dataset(h5.encode(), path.encode(), buf.ctypes.data_as(C.c_void_p), C.c_size_t(buf.size), shape, C.byref(rank), C.byref(es))
    if rc:
        raise RuntimeError(lib.b200host_last_error().decode())
    dims = tuple(shape[i] for i in range(rank.value))
    dt = {2: np.float16, 4: np.float32}[es.value]
    n = int(np.prod(dims)) * es.value
    return buf[:n].view(dt).reshape(dims).copy()


def load_keras_nif(assets_dir, lib=None):
    """-> (NifModel, NifMetaData, list of 'in x out act' strings)."""
    from ipu_path_trace_b200.nif import NifMetaData, NifModel
    lib = lib or host_lib()
    meta = NifMetaData.load(os.path.join(assets_dir, "nif_metadata.txt"))
    h5 = os.path.join(assets_dir, "converted.hdf5")
    out = C.create_string_buffer(1 << 18)
    if lib.b200host_h5_attribute(h5.encode(), b"/", b"model_config", out, len(out)):
        raise RuntimeError(lib.b200host_last_error().decode())
    cfg = json.loads(out.value.decode())
    assert cfg["class_name"] == "Functional"                      # Hdf5Model.cpp:16-19
    layers, topo = [], []
    for l in cfg["config"]["layers"]:
        if l["class_name"] != "Dense":                            # InputLayer / Concatenate are implicit (Hdf5Model.cpp:40-43)
            assert l["class_name"] in ("InputLayer", "Concatenate"), l["class_name"]
            continue
        c = l["config"]
        name = c["name"]
        kernel = _dataset(lib, h5, f"/model_weights/{name}/{name}/kernel:0")
        bias = _dataset(lib, h5, f"/model_weights/{name}/{name}/bias:0") if c["use_bias"] else None
        assert kernel.shape[1] == c["units"]
        layers.append({"kernel": kernel.astype(np.float16), "bias": None if bias is None else bias.astype(np.float16),
                       "relu": c["activation"] == "relu"})
        topo.append(f"{kernel.shape[0]}x{kernel.shape[1]} {c['activation']}{' +bias' if c['use_bias'] else ''} {kernel.dtype}")
    return NifModel(meta.embedding_dimension, meta.max, meta.mean, meta.log_tone_map, layers), meta, topo

"""Generates tests/golden/nif_twins.npz from the reference's host-side NIF twins compiled unmodified
(oracle/_ref/libref_niftwins.so: NifModel::generateInputSamples / makeCoefficients / makeGridCoordsUV / decodeSamples,
src/neural_networks/NifModel.cpp:403-443,466-509). Run in the container that has /root/reference:
    python tests/golden/make_nif_twins_golden.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402

O.build(ref=True)
ref = O.NifTwinsReference()
out = {}
for tag, (h, w, e) in {"a": (8, 16, 12), "b": (5, 7, 4), "c": (32, 64, 12)}.items():
    f, u, v = ref.encode_grid(h, w, e)
    out[f"{tag}_shape"] = np.array([h, w, e])
    out[f"{tag}_features"] = f
    out[f"{tag}_u"] = u
    out[f"{tag}_v"] = v
rng = np.random.default_rng(17)
x = rng.uniform(-1.5, 1.5, (4096, 3)).astype(np.float32)
mean = np.float32([-2.3514461517333984, -2.2660605907440186, -1.9648972749710083])      # the shipped nif_metadata.txt, raw
out["dec_x"] = x
out["dec_mean"] = mean
out["dec_max"] = np.float32(3.4299468994140625)
out["dec_eps"] = np.float32(1e-8)
out["dec_log"] = ref.decode(x, float(out["dec_max"]), mean, 1e-8, True)
out["dec_lin"] = ref.decode(x, float(out["dec_max"]), mean, 1e-8, False)
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "nif_twins.npz"), **out)
print("wrote nif_twins.npz:", {k: v.shape for k, v in out.items()})

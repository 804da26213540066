"""Generates tests/golden/ref_golden.npz from the UNMODIFIED reference codelets.

Run in the build container only (needs /root/reference to build oracle/_ref):
    python tests/golden/make_golden.py
The reference ships no golden vectors for this path, so these are outputs of the
reference's own code (src/codelets/codelets.cpp + external/light/src/*.hpp compiled through
oracle/shim) on seeded inputs. Inputs are stored next to the outputs, so neither the oracle
nor the CUDA path needs the reference tree to be checked against them.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402


def unit(v):
    return (v / np.linalg.norm(v, axis=-1, keepdims=True)).astype(np.float32)


def main():
    O.build(ref=True)
    ref = O.Reference()
    rng = np.random.default_rng(20261104)
    g = {}
    n = 512
    # pixelToRay light.hpp:194-206
    g["p2r_x"] = rng.uniform(0, 1104, n).astype(np.float32)
    g["p2r_y"] = rng.uniform(0, 1000, n).astype(np.float32)
    g["p2r_whf"] = np.array([1104, 1000, np.float32(1.5703125)], np.float32)
    g["p2r_out"] = ref.pixel_to_ray(g["p2r_x"], g["p2r_y"], 1104, 1000, float(g["p2r_whf"][2]))
    # Sphere::intersect light.hpp:131-147 (rays aimed near the sphere so hits and misses both occur)
    o = rng.uniform(-1, 1, (n, 3)).astype(np.float32)
    c = rng.uniform(-3, 3, (n, 3)).astype(np.float32)
    r = rng.uniform(0.3, 1.5, n).astype(np.float32)
    d = unit((c - o) + rng.normal(0, 0.6, (n, 3)) * r[:, None])
    g["sph_o"], g["sph_d"], g["sph_c"], g["sph_r"] = o, d, c, r
    g["sph_t"] = ref.sphere_intersect(o, d, c, r)
    # Disc::intersect light.hpp:106-120
    nn = unit(rng.normal(0, 1, (n, 3)))
    dc = rng.uniform(-3, 3, (n, 3)).astype(np.float32)
    dr = rng.uniform(0.5, 3.5, n).astype(np.float32)
    dd = unit((dc - o) + rng.normal(0, 1.0, (n, 3)) * dr[:, None])
    g["disc_o"], g["disc_d"], g["disc_n"], g["disc_c"], g["disc_r"] = o, dd, nn, dc, dr
    g["disc_t"] = ref.disc_intersect(o, dd, nn, dc, dr)
    # diffuse sdf.hpp:8-31
    u1 = O.half_round(rng.random(n)); u2 = O.half_round(rng.random(n))
    g["dif_n"], g["dif_u1"], g["dif_u2"] = nn, u1, u2
    g["dif_rr"] = np.float32(1.0 / (1.0 - float(O.half_round(0.3))))
    g["dif_d"], g["dif_w"] = ref.diffuse(nn, u1, u2, float(g["dif_rr"]))
    # reflect sdf.hpp:36-40 / refract sdf.hpp:45-67
    rd = unit(rng.normal(0, 1, (n, 3)))
    g["refl_d"], g["refl_n"] = rd, nn
    g["refl_out"] = ref.reflect(rd, nn)
    rnd = O.half_round(rng.random(n))
    g["refr_d"], g["refr_n"], g["refr_rnd"] = rd, nn, rnd
    g["refr_ri"] = np.float32(1.5)
    g["refr_out"], g["refr_flag"] = ref.refract(rd, nn, 1.5, rnd)
    # rouletteWeight light.hpp:25-29
    g["rou_rnd"] = O.half_round(rng.random(n))
    g["rou_p"] = O.half_round(0.3)
    g["rou_stop"], g["rou_w"] = ref.roulette(g["rou_rnd"], float(g["rou_p"]))
    # GenerateCameraRays + RayTraceKernel + PreProcess + PostProcess + Accumulate on a 64x48 image
    W, H = 64, 48
    s = O.Settings(width=W, height=H)
    items = O.make_worklist(W, H)
    noise = O.half_bits(rng.standard_normal((len(items), 2)).astype(np.float32))
    uni = O.half_bits(rng.random((len(items), 30)).astype(np.float32))
    g["img_wh"] = np.array([W, H], np.int32)
    g["img_noise"], g["img_uniforms"] = noise, uni
    rays = ref.camera_rays(items, noise, s)
    g["img_rays"] = rays
    rec = ref.ray_trace(rays, uni.reshape(-1), s, per_path=30)
    g["img_records"] = rec.view(np.uint8).reshape(len(items), -1)
    g["img_azimuth"] = np.float32(0.7)
    u, v, _ = ref.preprocess_escaped(rec, 0.7)
    g["img_u"], g["img_v"] = u, v
    bgr = rng.uniform(0.05, 3.0, (len(items), 3)).astype(np.float32)
    g["img_bgr"] = bgr
    out = ref.postprocess_and_accumulate(rec, bgr, items)
    g["img_accum"] = out.view(np.uint8).reshape(len(items), -1)
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_golden.npz")
    np.savez_compressed(path, **g)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()

"""End-to-end: the C++ host binary (host/b200_path_trace, the reference's command line) renders through the
C ABI on the GPU; its PNG/EXR are checked against the CPU oracle's image at matched samples per pixel
(north_star: converged images within 1 % relative RMSE) and against the Python host mirror."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "host")
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def cli(lib_built):
    subprocess.run(["make", "-C", HOST, "b200_path_trace"], check=True, capture_output=True)
    return os.path.join(HOST, "b200_path_trace")


def _assets(tmp_path, model, h5=False):
    from test_host_cpu import METADATA
    d = tmp_path / "assets.extra"
    d.mkdir(exist_ok=True)
    (d / "nif_metadata.txt").write_text(json.dumps(METADATA))
    if h5:
        from h5_writer import write_keras_nif
        write_keras_nif(str(d / "converted.hdf5"), model)
    else:
        from ipu_path_trace_b200.nif import write_nif_bin
        write_nif_bin(str(d / "nif.bin"), model)
    return str(d)


def _oracle_film(O, orc, lib_built, model, w, h, spp):
    items = lib_built.create_work_list_for_image(w, h)
    nif = O.NifLayers(model.embedding_dimension, model.max, model.mean, model.log_tone_map, model.layers)
    want, _, _, _ = orc.render(items, 0, spp, O.Settings(width=w, height=h), nif, mode=0)
    film = np.zeros((h, w, 3), np.float64)
    film[want["v"], want["u"], 0] = want["b"] / spp
    film[want["v"], want["u"], 1] = want["g"] / spp
    film[want["v"], want["u"], 2] = want["r"] / spp
    return film


@pytest.mark.parametrize("mode", ["device-film", "host-film", "hdf5+balance"])
def test_cli_renders_the_oracle_image(cli, lib_built, nif_model, O, orc, tmp_path, mode):
    import cv2
    os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
    w, h, spp, per_step = 96, 64, 16, 4
    assets = _assets(tmp_path, nif_model, h5=mode.startswith("hdf5"))
    out = str(tmp_path / "render.png")
    cmd = [cli, "-o", out, "--assets", assets, "-w", str(w), "-h", str(h), "-s", str(spp), "--samples-per-step", str(per_step),
           "--save-interval", "2", "--seed", "1", "--exposure", "0.5"]
    if mode == "host-film":
        cmd.append("--host-film")
    if mode == "hdf5+balance":
        cmd.append("--enable-load-balancing")
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "Render finished" in r.stderr and "Samples/sec" in r.stderr and "Saved images at step 4" in r.stderr
    exr = cv2.imread(out[:-4] + ".exr", cv2.IMREAD_UNCHANGED).astype(np.float64)
    png = cv2.imread(out, cv2.IMREAD_UNCHANGED)
    assert exr.shape == (h, w, 3) and png.shape == (h, w, 3)
    # EXR = sum over steps of per-step means / steps = mean radiance per pixel (AccumulatedImage.cpp:48-74)
    want = _oracle_film(O, orc, lib_built, nif_model, w, h, spp)
    rrmse = np.sqrt(np.mean((exr - want) ** 2)) / np.mean(want)
    assert rrmse < 0.01, rrmse
    # PNG = tone map of the EXR (:23-46)
    ldr = np.clip(np.rint(np.power(exr * 2 ** 0.5, 1 / 2.2) * 255), 0, 255)
    assert np.abs(png.astype(int) - ldr).max() <= 1


def test_cli_error_contract(cli, tmp_path, nif_model):
    assets = _assets(tmp_path, nif_model)
    r = subprocess.run([cli, "-o", str(tmp_path / "x.png"), "--assets", str(tmp_path / "nowhere"), "-w", "32", "-h", "32"], capture_output=True, text=True)
    assert r.returncode != 0 and "Could not load NIF model" in r.stderr
    r = subprocess.run([cli, "-o", str(tmp_path / "x.png"), "--assets", assets, "-w", "32", "-h", "32", "--aa-noise-type", "pink"], capture_output=True, text=True)
    assert r.returncode != 0 and "Invalid AA noise type: pink" in r.stderr
    r = subprocess.run([cli, "-o", str(tmp_path / "x.png"), "--assets", assets, "-w", "32", "-h", "32", "--roulette-depth", "0"], capture_output=True, text=True)
    assert r.returncode != 0 and "roulette_depth" in r.stderr

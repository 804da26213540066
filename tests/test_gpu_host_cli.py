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


def _free_port():
    import socket
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _rrmse(a, b):
    return float(np.sqrt(np.mean((a - b) ** 2)) / np.mean(b))


def test_cli_remote_ui_session(cli, nif_model, tmp_path):
    """--ui-port: a client of the reference's wire protocol steers the render (PathTracerApp.cpp:618-675,726-759):
    it receives progress, sample rates and the raw image; a new field of view restarts the render with the new
    camera; stop ends the process cleanly."""
    import cv2
    from ui_client import UiClient
    os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
    w, h = 96, 64
    assets = _assets(tmp_path, nif_model)
    base = [cli, "--assets", assets, "-w", str(w), "-h", str(h), "--samples-per-step", "128", "--seed", "3"]
    refs = {}
    for fov in (90, 50):                                   # head-less references of the two cameras
        out = str(tmp_path / f"ref{fov}.png")
        r = subprocess.run(base + ["-o", out, "-s", "32768", "--save-interval", "256", "--fov", str(fov), "--env-map-rotation", "30" if fov == 50 else "0"],
                           capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr[-2000:]
        refs[fov] = cv2.imread(out[:-4] + ".exr", cv2.IMREAD_UNCHANGED).astype(np.float64)[:, :, ::-1]      # -> R,G,B
    assert _rrmse(refs[50], refs[90]) > 0.2

    port = _free_port()
    out = str(tmp_path / "ui.png")
    p = subprocess.Popen(base + ["-o", out, "-s", str(128 * 1000000), "--save-interval", "16", "--ui-port", str(port), "--ui-accept-timeout", "120"],
                         stderr=subprocess.PIPE, text=True)
    try:
        c = UiClient(port, connect_timeout=120)
        assert c.wait(lambda c: len(c.images) >= 1 and len(c.rates) >= 16, timeout=120)
        first = c.images[0].astype(np.float64)
        assert first.shape == (h, w, 3)
        assert _rrmse(first, refs[90]) < 0.08, _rrmse(first, refs[90])
        assert c.rates[-1][0] > 0 and c.rates[-1][1] > c.rates[-1][0]          # rays/s > paths/s
        assert 0 < c.progress[-1] < 1e-3
        c.send_float("exposure", 1.0)                                         # no restart
        c.send_float("fov", 50.0)
        c.send_float("env_rotation", 30.0)
        assert c.wait(lambda c: any(_rrmse(i.astype(np.float64), refs[50]) < 0.08 for i in c.images[-2:]), timeout=120), \
            [_rrmse(i.astype(np.float64), refs[50]) for i in c.images[-3:]]
        with c.lock:
            drops = sum(1 for a, b in zip(c.progress, c.progress[1:]) if b < a)
        assert 1 <= drops <= 2
        c.send_bool("stop")
        rc = p.wait(timeout=60)
        err = p.stderr.read()
        assert rc == 0, err[-2000:]
        assert "Rendering stopped by remote UI" in err and "Render finished" in err
        c.close()
    finally:
        if p.poll() is None:
            p.kill()


def test_cli_remote_ui_runs_to_completion_and_saves(cli, nif_model, tmp_path):
    """With a client attached the files are written on the final step and the last raw image is that EXR."""
    import cv2
    from ui_client import UiClient
    os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
    w, h = 64, 48
    assets = _assets(tmp_path, nif_model)
    port = _free_port()
    out = str(tmp_path / "ui.png")
    p = subprocess.Popen([cli, "--assets", assets, "-w", str(w), "-h", str(h), "-s", "96", "--samples-per-step", "8", "--save-interval", "5", "-o", out,
                          "--ui-port", str(port), "--ui-accept-timeout", "120", "--host-film"], stderr=subprocess.PIPE, text=True)
    try:
        c = UiClient(port, connect_timeout=120)
        rc = p.wait(timeout=120)
        err = p.stderr.read()
        assert rc == 0, err[-2000:]
        assert c.closed.wait(10)
        assert "Saved images at step 12" in err and "Saved images at step 5" not in err
        assert len(c.images) >= 1 and abs(c.progress[-1] - 1.0) < 1e-6
        exr = cv2.imread(out[:-4] + ".exr", cv2.IMREAD_UNCHANGED)
        np.testing.assert_allclose(c.images[-1][:, :, ::-1], exr, rtol=2e-3, atol=1e-6)
        c.close()
    finally:
        if p.poll() is None:
            p.kill()


def test_cli_load_balancing_really_redeals(cli, lib_built, nif_model, O, orc, tmp_path):
    """--enable-load-balancing (LoadBalancer.cpp:141-192, PathTracerApp.cpp:707-735; ADVICE r1: it used to be a no-op).
    The two work lists alternate: the list uploaded at steps 1-3 is the initial deal (step 1's task does not re-deal,
    step 2's task re-deals the OTHER list), from step 4 on every upload is a list re-dealt by path length, so its order
    digest differs from the initial one. The image stays the oracle's (RNG is keyed on the pixel, not the slot)."""
    import re
    import cv2
    os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
    import torch
    gpus = min(2, torch.cuda.device_count())
    w, h, spp, per_step = 96, 64, 24, 4
    assets = _assets(tmp_path, nif_model)
    out = str(tmp_path / "lb.png")
    r = subprocess.run([cli, "-o", out, "--assets", assets, "-w", str(w), "-h", str(h), "-s", str(spp), "--samples-per-step", str(per_step),
                        "--save-interval", "6", "--seed", "1", "--enable-load-balancing", "--ipus", str(gpus), "--log-level", "debug"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    digests = dict((int(a), b) for a, b in re.findall(r"Work list order digest at step (\d+): (\d+)", r.stderr))
    assert sorted(digests) == [1, 2, 3, 4, 5, 6], r.stderr[-3000:]
    assert digests[1] == digests[2] == digests[3]
    assert digests[4] != digests[1] and digests[5] != digests[1] and digests[6] != digests[4]
    assert r.stderr.count("Load balancing started") == 5          # every step but the first
    exr = cv2.imread(out[:-4] + ".exr", cv2.IMREAD_UNCHANGED).astype(np.float64)
    want = _oracle_film(O, orc, lib_built, nif_model, w, h, spp)
    assert _rrmse(exr, want) < 0.01


@pytest.mark.parametrize("film", ["--host-film", "device"])
def test_cli_survives_a_client_that_drops_the_socket_mid_render(cli, nif_model, tmp_path, film):
    """the async host task may be inside startSendingRawImage when the link dies. A client connects, takes the
    first raw image and then closes the socket without `detach` or `stop`, in the middle of the save intervals; the render
    must carry on head-less to the end, write its files and exit 0 (PathTracerApp.cpp:543-546: "If the Remote-UI detaches
    just continue rendering")."""
    import cv2
    from ui_client import UiClient
    os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
    w, h = 128, 96
    assets = _assets(tmp_path, nif_model)
    port = _free_port()
    out = str(tmp_path / "drop.png")
    # 4096 steps of 64 samples (~5 s): the render must still be running when the client has its first image and hangs up
    cmd = [cli, "--assets", assets, "-w", str(w), "-h", str(h), "-s", "262144", "--samples-per-step", "64", "--save-interval", "32", "-o", out,
           "--ui-port", str(port), "--ui-accept-timeout", "120"] + ([film] if film.startswith("--") else [])
    log = open(tmp_path / "stderr.log", "w")                    # 4096 log lines: a pipe nobody drains would fill up and block the render
    p = subprocess.Popen(cmd, stderr=log, text=True)
    try:
        c = UiClient(port, connect_timeout=120)
        assert c.wait(lambda c: len(c.images) >= 1 and len(c.progress) >= 4, timeout=120)
        c.close()                                               # hang up, no good-bye
        rc = p.wait(timeout=180)
        log.close()
        err = open(tmp_path / "stderr.log").read()
        assert rc == 0, err[-2000:]
        assert "Remote UI connection lost" in err or "Remote UI disconnected" in err
        assert "Render finished" in err and "Saved images at step 4096" in err
        exr = cv2.imread(out[:-4] + ".exr", cv2.IMREAD_UNCHANGED)
        assert exr is not None and exr.shape == (h, w, 3) and np.isfinite(exr).all() and exr.mean() > 0
    finally:
        if p.poll() is None:
            p.kill()

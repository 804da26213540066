"""CPU tests of the C++ host (host/): option parser, work lists, film, PNG/EXR writers, NIF asset loaders
and the HDF5 subset reader, through libb200pt_host.so (no GPU code inside). References:
src/main.cpp:8-66, src/PathTracerApp.cpp:19-27,794-830, src/LoadBalancer.cpp, src/AccumulatedImage.cpp,
src/neural_networks/NifMetaData.cpp, src/keras/Hdf5Model.cpp."""
import ctypes as C
import json
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "host")
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

# The shipped metadata of the reference's NIF model, verbatim (nif_models/urban_alley_01_4k_fp16_yuv/assets.extra/
# nif_metadata.txt; a 1.2 KB data fixture, the weights blob next to it is missing from the snapshot).
SHIPPED_METADATA = os.path.join(ROOT, "tests", "golden", "nif_metadata.txt")
METADATA = json.load(open(SHIPPED_METADATA))


@pytest.fixture(scope="session")
def host():
    subprocess.run(["make", "-C", HOST, "libb200pt_host.so"], check=True, capture_output=True)
    lib = C.CDLL(os.path.join(HOST, "libb200pt_host.so"))
    lib.b200host_last_error.restype = C.c_char_p
    lib.b200host_create_jobs.restype = C.c_long
    lib.b200host_balance.restype = C.c_long
    lib.b200host_round_spp.restype = C.c_size_t
    lib.b200host_round_spp.argtypes = [C.c_size_t, C.c_size_t]
    return lib


def _opts(host, *argv):
    args = (C.c_char_p * (len(argv) + 1))(b"b200_path_trace", *[a.encode() for a in argv])
    out = C.create_string_buffer(1 << 14)
    rc = host.b200host_parse_options(len(argv) + 1, args, out, len(out))
    if rc:
        raise RuntimeError(host.b200host_last_error().decode())
    return dict(l.split("=", 1) for l in out.value.decode().splitlines())


def test_option_defaults_match_the_reference(host):
    o = _opts(host, "-o", "out.png", "--assets", "a")
    want = {"ipus": "1", "save-interval": "1", "width": "256", "height": "256", "samples": "512", "samples-per-step": "512",
            "interactive-samples": "8", "refractive-index": "1.5", "roulette-depth": "3", "stop-prob": "0.3", "aa-noise-scale": "0.3",
            "fov": "90", "exposure": "0", "gamma": "2.2", "env-map-rotation": "0", "seed": "1", "aa-noise-type": "normal",
            "max-path-length": "10", "max-nif-batch-size": "44160", "ui-port": "0", "log-level": "info", "enable-load-balancing": "0"}
    for k, v in want.items():
        assert o[k] == v, k
    assert o["outfile"] == "out.png" and o["assets"] == "a"


def test_option_forms_and_errors(host):
    o = _opts(host, "-w", "1104", "-h1000", "--samples=100000", "--samples-per-step", "300", "-o", "x.png", "--assets=/m",
              "--ipus", "8", "--enable-load-balancing", "-a", "0.5", "-n", "1.33", "--env-map-rotation", "-45")
    assert (o["width"], o["height"], o["samples"], o["samples-per-step"], o["ipus"]) == ("1104", "1000", "100000", "300", "8")
    assert o["enable-load-balancing"] == "1" and o["aa-noise-scale"] == "0.5" and o["refractive-index"] == "1.33"
    assert o["env-map-rotation"] == "-45"
    with pytest.raises(RuntimeError, match="required but missing"):
        _opts(host, "-o", "x.png")
    with pytest.raises(RuntimeError, match="unrecognised option"):
        _opts(host, "-o", "x.png", "--assets", "a", "--bogus")
    with pytest.raises(RuntimeError, match="is missing"):
        _opts(host, "--assets", "a", "-o")
    with pytest.raises(RuntimeError, match="both save-exe and load-exe"):
        _opts(host, "-o", "x", "--assets", "a", "--save-exe", "p", "--load-exe", "q")
    with pytest.raises(RuntimeError, match="Show help"):
        _opts(host, "--help")


def test_round_samples_per_pixel(host):
    assert host.b200host_round_spp(100000, 300) == 100200        # BASELINE config C2: 334 steps
    assert host.b200host_round_spp(512, 512) == 512
    assert host.b200host_round_spp(1, 300) == 300


REC = np.dtype([("u", "<u2"), ("v", "<u2"), ("r", "<f4"), ("g", "<f4"), ("b", "<f4"), ("sampleCount", "<u2"), ("pathLength", "<u2")])


@pytest.mark.parametrize("w,h,gpus,tile", [(1104, 1000, 1, 16), (1104, 1000, 8, 16), (320, 240, 3, 16), (33, 17, 4, 8), (5, 3, 8, 16), (1, 1, 2, 16)])
def test_tracing_jobs_cover_every_pixel_once(host, w, h, gpus, tile):
    per = host.b200host_create_jobs(w, h, gpus, tile, None, 0)
    assert per > 0
    items = np.zeros(per * gpus, dtype=REC)
    assert host.b200host_create_jobs(w, h, gpus, tile, items.ctypes.data_as(C.c_void_p), C.c_size_t(len(items))) == per
    valid = items[(items["u"] < w) & (items["v"] < h)]
    pad = items[(items["u"] >= w) | (items["v"] >= h)]
    assert len(valid) == w * h
    assert len(np.unique(valid["v"].astype(np.int64) * w + valid["u"])) == w * h
    assert np.all(pad["u"] == 65535) and np.all(pad["v"] == 65535)          # LoadBalancer.cpp:66-71
    if w * h >= gpus * tile * tile * 8:
        counts = [np.sum(items[g * per:(g + 1) * per]["u"] < w) for g in range(gpus)]
        assert max(counts) - min(counts) <= tile * tile * 2 + tile * max(w, h) // tile


def test_load_balancer_deals_cheap_with_expensive(host):
    rng = np.random.default_rng(3)
    jobs, per = 4, 250
    items = np.zeros(jobs * per, dtype=REC)
    items["u"] = np.arange(len(items)) % 50
    items["v"] = np.arange(len(items)) // 50
    items["pathLength"] = rng.integers(1, 2000, len(items))
    items["r"] = 1.0
    before = items.copy()
    rays = host.b200host_balance(items.ctypes.data_as(C.c_void_p), C.c_size_t(per), C.c_size_t(jobs), 1)
    assert rays == int(before["pathLength"].astype(np.int64).sum())
    # same multiset of work items, now dealt so that every job's cost is close to the mean
    key = lambda a: np.sort(a["v"].astype(np.int64) * 1000 + a["u"])
    assert np.array_equal(key(items), key(before))
    cost = items["pathLength"].astype(np.int64).reshape(jobs, per).sum(1)
    naive = before["pathLength"].astype(np.int64).reshape(jobs, per).sum(1)
    assert cost.max() - cost.min() <= 2000 * 2
    assert cost.max() - cost.min() <= max(1, naive.max() - naive.min())


def _film(host, items, w, h, step, exposure, gamma, name=b""):
    hdr = np.zeros((h, w, 3), np.float32)
    ldr = np.zeros((h, w, 3), np.uint8)
    rc = host.b200host_film_accumulate_save(items.ctypes.data_as(C.c_void_p), C.c_size_t(len(items)), w, h, step, C.c_float(exposure),
                                            C.c_float(gamma), name, hdr.ctypes.data_as(C.c_void_p), ldr.ctypes.data_as(C.c_void_p))
    assert rc == 0, host.b200host_last_error()
    return hdr, ldr


def test_film_accumulate_and_tonemap_follow_accumulated_image(host, tmp_path):
    import cv2
    os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
    rng = np.random.default_rng(5)
    w, h, spp, step = 37, 21, 7, 3
    items = np.zeros(w * h + 11, dtype=REC)
    items["u"][:w * h] = np.arange(w * h) % w
    items["v"][:w * h] = np.arange(w * h) // w
    items["u"][w * h:] = 65535
    items["v"][w * h:] = 65535
    for k in "rgb":
        items[k] = rng.gamma(1.5, 2.0, len(items)).astype(np.float32) * spp
    items["sampleCount"] = spp
    out = str(tmp_path / "img.png")
    hdr, ldr = _film(host, items, w, h, step, 0.5, 2.2, out.encode())
    # AccumulatedImage.cpp:59-74: hdr(v,u) += (b,g,r) / sampleCount, once per step
    want = np.zeros((h, w, 3), np.float32)
    v = items[:w * h]
    for _ in range(step):
        want[v["v"], v["u"], 0] += v["b"] / np.float32(spp)
        want[v["v"], v["u"], 1] += v["g"] / np.float32(spp)
        want[v["v"], v["u"], 2] += v["r"] / np.float32(spp)
    np.testing.assert_allclose(hdr, want, rtol=1e-6)
    # :23-46 with OpenCV's own arithmetic: pow, then convertTo(CV_8UC3, 255)
    scaled = hdr * np.float32(1.0 / step)
    ref_ldr = cv2.convertScaleAbs(np.power(scaled * np.float32(2.0 ** 0.5), np.float32(1 / 2.2)), alpha=255.0)
    assert np.abs(ldr.astype(int) - ref_ldr.astype(int)).max() <= 1
    assert np.mean(ldr == ref_ldr) > 0.99
    # the files: OpenCV reads back what AccumulatedImage::saveImages would have written with cv::imwrite
    png = cv2.imread(out, cv2.IMREAD_UNCHANGED)
    assert png.shape == (h, w, 3) and np.array_equal(png, ldr)
    exr = cv2.imread(out[:-4] + ".exr", cv2.IMREAD_UNCHANGED)
    assert exr.dtype == np.float32 and exr.shape == (h, w, 3)
    np.testing.assert_array_equal(exr, scaled)


@pytest.mark.parametrize("w,h", [(1, 1), (16, 16), (17, 33), (640, 37), (1104, 1000)])     # the last one takes the multi-band deflate
def test_png_exr_round_trip_and_interop(host, tmp_path, w, h):
    import cv2
    os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
    rng = np.random.default_rng(w * 1000 + h)
    img8 = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    imgf = (rng.standard_normal((h, w, 3)) * 10).astype(np.float32)
    imgf[0, 0] = [0.0, 1e-30, 6.5e4]
    if h >= 1000:      # upper half smooth (deflated blocks), lower half noise (blocks the writer's probe stores raw): both in one file
        imgf[: h // 2] = (np.arange(w, dtype=np.float32)[None, :, None] / w + np.arange(h // 2, dtype=np.float32)[:, None, None] / h)
    p, e = str(tmp_path / "a.png").encode(), str(tmp_path / "a.exr").encode()
    assert host.b200host_write_png(p, img8.ctypes.data_as(C.c_void_p), w, h) == 0
    assert host.b200host_write_exr(e, imgf.ctypes.data_as(C.c_void_p), w, h) == 0
    assert np.array_equal(cv2.imread(p.decode(), cv2.IMREAD_UNCHANGED).reshape(h, w, 3), img8)
    assert np.array_equal(cv2.imread(e.decode(), cv2.IMREAD_UNCHANGED).reshape(h, w, 3), imgf)
    if h >= 1000:
        assert 0.55 * imgf.nbytes < os.path.getsize(e) < 0.95 * imgf.nbytes       # the smooth half shrank, the noise half was stored
    # our readers on our files
    ww, hh = C.c_uint32(), C.c_uint32()
    back8 = np.zeros_like(img8)
    assert host.b200host_read_png(p, back8.ctypes.data_as(C.c_void_p), C.c_size_t(back8.size), C.byref(ww), C.byref(hh)) == 0
    assert (ww.value, hh.value) == (w, h) and np.array_equal(back8, img8)
    backf = np.zeros_like(imgf)
    assert host.b200host_read_exr(e, backf.ctypes.data_as(C.c_void_p), C.c_size_t(backf.size), C.byref(ww), C.byref(hh)) == 0
    assert (ww.value, hh.value) == (w, h) and np.array_equal(backf, imgf)
    # our readers on OpenCV's files (what the reference itself writes)
    p2, e2 = str(tmp_path / "b.png"), str(tmp_path / "b.exr")
    cv2.imwrite(p2, img8)
    cv2.imwrite(e2, imgf, [cv2.IMWRITE_EXR_TYPE, cv2.IMWRITE_EXR_TYPE_FLOAT])
    assert host.b200host_read_png(p2.encode(), back8.ctypes.data_as(C.c_void_p), C.c_size_t(back8.size), C.byref(ww), C.byref(hh)) == 0
    assert np.array_equal(back8, img8)
    rc = host.b200host_read_exr(e2.encode(), backf.ctypes.data_as(C.c_void_p), C.c_size_t(backf.size), C.byref(ww), C.byref(hh))
    if rc == 0:          # OpenCV may pick a compression this reader does not implement (PIZ); NONE/ZIP(S) must agree
        assert np.array_equal(backf, imgf)
    else:
        assert b"compression" in host.b200host_last_error()


def _assets(tmp_path, model=None, meta=METADATA, h5=None, **h5kw):
    d = tmp_path / "assets.extra"
    d.mkdir(exist_ok=True)
    (d / "nif_metadata.txt").write_text(json.dumps(meta, indent=2))
    if model is not None and h5 is None:
        from ipu_path_trace_b200.nif import write_nif_bin
        write_nif_bin(str(d / "nif.bin"), model)
    if h5:
        from h5_writer import write_keras_nif
        write_keras_nif(str(d / "converted.hdf5"), model, **h5kw)
    return str(d)


def _load(host, path, synthetic=0):
    out = C.create_string_buffer(4096)
    k0 = np.zeros(48 * 320, np.uint16)
    rc = host.b200host_load_assets(path.encode(), synthetic, out, len(out), k0.ctypes.data_as(C.c_void_p), C.c_size_t(k0.size))
    if rc:
        raise RuntimeError(host.b200host_last_error().decode())
    s = out.value.decode()
    shapes = s.split(";shapes=")[1].split(";")[0]
    fields = dict(kv.split("=", 1) for kv in s.replace(";shapes=" + shapes, "").split(";"))
    return fields, shapes, k0


def test_metadata_and_nif_bin_loader(host, tmp_path):
    from ipu_path_trace_b200.nif import synthetic_model
    m = synthetic_model()
    f, shapes, k0 = _load(host, _assets(tmp_path, m))
    assert f["layers"] == "7" and f["E"] == "12" and f["hidden"] == "320" and f["log"] == "1"
    assert int(f["flops"]) == m.flops_per_eval() == 1089283
    assert shapes == "48x320rbh|320x320rbh|320x320rbh|368x320rbh|320x320rbh|320x320rbh|320x3nbh|"
    assert np.array_equal(k0, m.layers[0]["kernel"].view(np.uint16).reshape(-1))
    # NifMetaData.cpp:48-53: eps folded into the mean when log tone mapping
    mean = [float(x) for x in f["mean"].split(",")]
    for got, raw in zip(mean, METADATA["encode_params"]["mean"]):
        assert np.float32(got) == np.float32(np.float32(raw) - np.float32(1e-8))
    assert np.float32(float(f["max"])) == np.float32(3.4299468994140625)


def test_shipped_metadata_file_parses_like_nifmetadata(host, tmp_path):
    """NifMetaData.cpp:11-71 on the reference's own nif_metadata.txt (byte-for-byte copy under tests/golden/), through the C++
    loader (host/NifAssets.cpp) and the Python mirror: embedding 12, eps folded into the mean (:48-53), max, the hidden
    size scraped from `--layer-size` in train_command (:55-64), name and image shape."""
    import shutil
    from ipu_path_trace_b200.nif import NifMetaData
    d = tmp_path / "assets.extra"
    d.mkdir()
    shutil.copy(SHIPPED_METADATA, d / "nif_metadata.txt")
    want_mean = [np.float32(np.float32(m) - np.float32(1e-8)) for m in (-2.3514461517333984, -2.2660605907440186, -1.9648972749710083)]
    md = NifMetaData.load(str(d / "nif_metadata.txt"))
    assert md.embedding_dimension == 12 and md.hidden_size == 320 and md.log_tone_map is True
    assert md.name == "/home/ubuntu/workspace/hdris/urban_alley_01_4k.exr" and md.image_shape == (2048, 4096, 3)
    assert np.float32(md.max) == np.float32(3.4299469) and np.float32(md.eps) == np.float32(1e-8)
    assert [np.float32(m) for m in md.mean] == want_mean
    f, shapes, _ = _load(host, str(d), synthetic=1)        # no weights next to it: synthetic ones in the topology the metadata names
    assert f["E"] == "12" and f["hidden"] == "320" and f["log"] == "1"
    assert np.float32(float(f["max"])) == np.float32(3.4299469)
    assert [np.float32(float(x)) for x in f["mean"].split(",")] == want_mean
    assert shapes.startswith("48x320") and shapes.count("|") == 7


def test_asset_errors(host, tmp_path):
    with pytest.raises(RuntimeError, match="Could not open file"):
        _load(host, str(tmp_path / "missing"))
    p = _assets(tmp_path)
    with pytest.raises(RuntimeError, match="No weights"):
        _load(host, p)
    f, shapes, _ = _load(host, p, synthetic=1)
    assert f["source"].startswith("synthetic") and shapes.count("|") == 7
    bad = dict(METADATA); del bad["encode_params"]
    (tmp_path / "b").mkdir()
    with pytest.raises(RuntimeError, match="Error reading property"):
        _load(host, _assets(tmp_path / "b", meta=bad), synthetic=1)
    (tmp_path / "c" ).mkdir()
    d = _assets(tmp_path / "c")
    open(os.path.join(d, "nif_metadata.txt"), "w").write("{ not json")
    with pytest.raises(RuntimeError, match="JSON parse error"):
        _load(host, d, synthetic=1)


@pytest.mark.parametrize("kw", [dict(), dict(vlen_attrs=True), dict(dtype=np.float32), dict(chunked=True), dict(chunked=True, dtype=np.float32, vlen_attrs=True)])
def test_keras_hdf5_weights_load_like_hdf5model(host, tmp_path, kw):
    """converted.hdf5 laid out as Keras/h5py write it (tests/h5_writer.py, from the format specification)
    is read by the host's HDF5 subset exactly as Hdf5Model.cpp:62-133 would hand it over."""
    from ipu_path_trace_b200.nif import synthetic_model
    m = synthetic_model(seed=7)
    path = _assets(tmp_path, m, h5=True, **kw)
    f, shapes, k0 = _load(host, path)
    half = kw.get("dtype", np.float16) == np.float16
    t = "h" if half else "f"
    assert shapes == f"48x320rb{t}|320x320rb{t}|320x320rb{t}|368x320rb{t}|320x320rb{t}|320x320rb{t}|320x3nb{t}|"
    assert f["source"].endswith("converted.hdf5") and int(f["flops"]) == 1089283
    if half:
        assert np.array_equal(k0, m.layers[0]["kernel"].view(np.uint16).reshape(-1))
    # dataset / attribute / listing entry points
    h5 = os.path.join(path, "converted.hdf5").encode()
    shape = (C.c_size_t * 4)()
    rank, es = C.c_int(4), C.c_size_t()
    buf = np.zeros(368 * 320 * 4, np.uint8)
    assert host.b200host_h5_dataset(h5, b"/model_weights/dense_3/dense_3/kernel:0", buf.ctypes.data_as(C.c_void_p), C.c_size_t(buf.size),
                                    shape, C.byref(rank), C.byref(es)) == 0, host.b200host_last_error()
    assert rank.value == 2 and (shape[0], shape[1]) == (368, 320) and es.value == (2 if half else 4)
    want = np.asarray(m.layers[3]["kernel"], dtype=np.float16 if half else np.float32)
    assert np.array_equal(buf[:want.nbytes].view(want.dtype).reshape(368, 320), want)
    out = C.create_string_buffer(1 << 15)
    assert host.b200host_h5_attribute(h5, b"/", b"keras_version", out, len(out)) == 0 and out.value == b"2.6.0"
    assert host.b200host_h5_attribute(h5, b"/", b"backend", out, len(out)) == 0 and out.value == b"tensorflow"
    assert host.b200host_h5_attribute(h5, b"/", b"model_config", out, len(out)) == 0
    cfg = json.loads(out.value.decode())
    assert cfg["class_name"] == "Functional" and sum(l["class_name"] == "Dense" for l in cfg["config"]["layers"]) == 7
    assert host.b200host_h5_list(h5, b"/model_weights", out, len(out)) == 0
    names = out.value.decode().split()
    assert names == sorted(names) and len(names) == 9 and "dense_6" in names
    assert host.b200host_h5_dataset(h5, b"/model_weights/nope", buf.ctypes.data_as(C.c_void_p), C.c_size_t(buf.size), shape, C.byref(rank), C.byref(es)) != 0
    assert b"doesn't exist" in host.b200host_last_error()


def test_hdf5_reader_rejects_garbage(host, tmp_path):
    p = tmp_path / "x.hdf5"
    p.write_bytes(b"not an hdf5 file" * 10)
    out = C.create_string_buffer(64)
    assert host.b200host_h5_attribute(str(p).encode(), b"/", b"a", out, len(out)) != 0
    assert b"not an HDF5 file" in host.b200host_last_error()
    from ipu_path_trace_b200.nif import synthetic_model
    from h5_writer import write_keras_nif
    good = tmp_path / "g.hdf5"
    write_keras_nif(str(good), synthetic_model(hidden=64, hidden_layers=2, concat_at=1, embedding_dimension=4))
    data = good.read_bytes()
    (tmp_path / "t.hdf5").write_bytes(data[:len(data) // 3])       # truncated file: an error, never a crash
    rc = host.b200host_h5_attribute(str(tmp_path / "t.hdf5").encode(), b"/", b"model_config", out, len(out))
    assert rc != 0 and b"HDF5" in host.b200host_last_error()


def test_cli_binary_fails_loudly_without_gpu(tmp_path):
    """The product has no CPU path: without a CUDA device the CLI exits non-zero with the library's message."""
    subprocess.run(["make", "-C", HOST, "b200_path_trace"], check=True, capture_output=True)
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    d = _assets(tmp_path)
    r = subprocess.run([os.path.join(HOST, "b200_path_trace"), "-o", str(tmp_path / "o.png"), "--assets", d, "--synthetic-weights", "-w", "32", "-h", "32"],
                       capture_output=True, text=True)
    assert r.returncode != 0
    assert "no CUDA device" in r.stderr and "no CPU path" in r.stderr
    assert not (tmp_path / "o.png").exists()


# ---- remote user interface (src/InterfaceServer.hpp, external/packetcomms) --------------------------------

def _free_port():
    import socket
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _ui_session(host, port, result, **kw):
    out = C.create_string_buffer(1024)
    rc = host.b200host_ui_session(port, kw.get("w", 12), kw.get("h", 7), kw.get("steps", 40), kw.get("save", 4), kw.get("step_ms", 20),
                                  kw.get("accept", 20), out, len(out))
    result["rc"] = rc
    result["err"] = host.b200host_last_error().decode() if rc else ""
    result["out"] = dict(kv.split("=", 1) for kv in out.value.decode().split(";")) if rc == 0 else {}


def test_ui_protocol_progress_image_restart_and_stop(host):
    """A client speaking the packetcomms wire format gets Hello, progress, sample rates and the raw image row by row
    (B,G,R -> R,G,B, divided by the step); changing fov / rotation / interactive samples / NIF restarts the render,
    exposure and gamma do not; stop ends it (InterfaceServer.hpp:113-176,270-331, PathTracerApp.cpp:531-564,651-675)."""
    import threading
    from ui_client import HELLO, UiClient
    port, res = _free_port(), {}
    t = threading.Thread(target=_ui_session, args=(host, port, res), kwargs=dict(steps=100000, save=3))
    t.start()
    c = UiClient(port)
    assert c.wait(lambda c: len(c.images) >= 1 and len(c.progress) >= 3)
    assert c.control[0] == HELLO
    img = c.images[0]
    assert img.shape == (7, 12, 3)
    # film = (fov, rot, isamples) * step in B,G,R; sent as R,G,B / step -> (isamples, rot, fov) = (8, 0, 1)
    np.testing.assert_array_equal(img, np.broadcast_to(np.float32([8, 0, 1]), img.shape))
    assert c.rates[0] == (1000.0, 2000.0)
    assert abs(c.progress[1] - 2 / 100000) < 1e-9

    # tone-mapping parameters do not restart
    c.send_float("exposure", 1.5)
    c.send_float("gamma", 1.8)
    n = len(c.progress)
    assert c.wait(lambda c: len(c.progress) >= n + 3)
    assert c.progress[-1] > c.progress[n - 1]
    # view parameters restart: progress drops back to step 1 and the image shows the new state
    c.send_float("fov", 180.0)
    c.send_float("env_rotation", 45.0)
    c.send_u32("interactive_samples", 3)
    c.send_string("load_nif", "/some/where/assets.extra")
    assert c.wait(lambda c: any(np.allclose(i[0, 0], [3, 45, np.float32(np.pi)]) for i in c.images))
    with c.lock:
        drops = sum(1 for a, b in zip(c.progress, c.progress[1:]) if b < a)
    assert 1 <= drops <= 4
    c.send_bool("stop")
    t.join(30)
    assert not t.is_alive() and res["rc"] == 0, res
    o = res["out"]
    assert o["stopped"] == "1" and o["detached"] == "0" and 1 <= int(o["restarts"]) <= 4
    assert float(o["exposure"]) == 1.5 and abs(float(o["gamma"]) - 1.8) < 1e-6
    assert abs(float(o["fov"]) - np.pi) < 1e-6 and float(o["rot"]) == 45.0 and o["isamples"] == "3"
    assert o["nif"] == "/some/where/assets.extra"
    assert c.closed.wait(10)
    c.close()


def test_ui_detach_and_lost_connection_continue_headless(host):
    import threading
    from ui_client import UiClient
    for how in ("detach", "hangup"):
        port, res = _free_port(), {}
        t = threading.Thread(target=_ui_session, args=(host, port, res), kwargs=dict(steps=25, save=5, step_ms=10))
        t.start()
        c = UiClient(port)
        assert c.wait(lambda c: len(c.progress) >= 2)
        if how == "detach":
            c.send_bool("detach")
            assert c.closed.wait(10)
        c.close()
        t.join(30)
        assert not t.is_alive() and res["rc"] == 0, res
        assert res["out"]["steps"] == "25" and res["out"]["stopped"] == "0"
        assert res["out"]["detached"] == ("1" if how == "detach" else "0")


def test_ui_heartbeat_bad_hello_and_accept_timeout(host):
    import socket
    import struct
    import threading
    from ui_client import HEARTBEAT, UiClient
    # no client: start() gives up after the timeout with an error (the reference waits for ever)
    res = {}
    _ui_session(host, _free_port(), res, accept=1)
    assert res["rc"] != 0 and "no client connected" in res["err"]
    # an idle link carries a heart beat every second (PacketMuxer.cpp:140-160)
    port, res = _free_port(), {}
    t = threading.Thread(target=_ui_session, args=(host, port, res), kwargs=dict(steps=2, save=100, step_ms=1400))
    t.start()
    c = UiClient(port)
    assert c.wait(lambda c: HEARTBEAT in c.control, timeout=10)
    t.join(30)
    c.close()
    assert res["rc"] == 0 and res["out"]["steps"] == "2"
    # a peer that does not start with Hello is dropped; the render carries on head-less
    port, res = _free_port(), {}
    t = threading.Thread(target=_ui_session, args=(host, port, res), kwargs=dict(steps=10, save=100, step_ms=20))
    t.start()
    s = None
    for _ in range(200):
        try:
            s = socket.create_connection(("127.0.0.1", port), timeout=1.0)
            break
        except OSError:
            import time
            time.sleep(0.05)
    s.sendall(struct.pack(">II", 9, 4) + struct.pack("<f", 30.0))      # fov before Hello
    t.join(30)
    s.close()
    assert res["rc"] == 0 and res["out"]["steps"] == "10" and res["out"]["restarts"] == "0"


def test_cornell_cli_options_and_no_cpu_path(tmp_path):
    """host/b200_trace keeps the option table of external/light/src/trace.cpp:19-35 and, like everything else here, has no CPU path."""
    subprocess.run(["make", "-C", HOST, "b200_trace"], check=True, capture_output=True)
    exe = os.path.join(HOST, "b200_trace")
    h = subprocess.run([exe, "--help"], capture_output=True, text=True)
    for opt in ("--outfile", "--width", "--height", "--tile-width", "--tile-height", "--samples", "--refractive-index", "--roulette-depth", "--stop-prob",
                "--aa-noise-scale", "--seed", "--epsilon", "--intersection-epsilon", "--no-gui"):
        assert opt in h.stdout, opt
    for short, default in (("-w [ --width ] arg (=256)", None), ("-s [ --samples ] arg (=32)", None), ("-n [ --refractive-index ] arg (=1.5)", None),
                           ("--stop-prob arg (=0.2)", None), ("--roulette-depth arg (=3)", None), ("--tile-width arg (=16)", None)):
        assert short in h.stdout, short
    r = subprocess.run([exe, "-w", "8"], capture_output=True, text=True)
    assert r.returncode != 0 and "outfile" in r.stderr
    import torch
    if not torch.cuda.is_available():
        out = str(tmp_path / "c.png")
        r = subprocess.run([exe, "-o", out, "-w", "16", "-h", "16", "-s", "2", "--no-gui"], capture_output=True, text=True)
        assert r.returncode != 0 and "no CUDA device" in r.stderr and "no CPU path" in r.stderr
        assert not os.path.exists(out) and not os.path.exists(out + ".exr")
        opts = dict(l.split(" ", 1) for l in open(out + ".command_line_options").read().splitlines())
        assert opts["width"] == "16" and opts["samples"] == "2" and opts["seed"] == "1"


def test_ui_protocol_with_the_references_own_client(host):
    """The same session as test_ui_protocol_progress_image_restart_and_stop, but the peer is the REFERENCE's packetcomms
    (PacketDemuxer / PacketMuxer / TcpSocket) + cereal compiled unmodified into oracle/_ref/libref_uiclient.so, with the packet
    table and the HdrHeader / HdrPacket / SampleRates serialisers of src/InterfaceServer.hpp:21-83: host/InterfaceServer.hpp
    and host/PacketComms.hpp interoperate with the reference's implementation of the wire format, in both directions
    (Hello handshake PacketDemuxer.cpp:110,230-250; framing PacketMuxer.cpp:114-138)."""
    import threading
    import ui_ref_client as R
    if not R.available():
        pytest.skip("oracle/_ref/libref_uiclient.so not built here (needs /root/reference)")
    from ui_client import PACKETS
    port, res = _free_port(), {}
    t = threading.Thread(target=_ui_session, args=(host, port, res), kwargs=dict(steps=100000, save=3))
    t.start()
    c = R.RefUiClient(port)
    assert c.packet_types() == PACKETS                      # the hand-copied table equals the reference's
    assert c.wait(lambda c: c._counts()[2] >= 1 and c._counts()[0] >= 3)
    assert c.ok()                                           # the reference's demuxer accepted our Hello
    img = c.image(0)
    assert img.shape == (7, 12, 3)
    np.testing.assert_array_equal(img, np.broadcast_to(np.float32([8, 0, 1]), img.shape))
    assert c.rates[0] == (1000.0, 2000.0)
    assert abs(c.progress[1] - 2 / 100000) < 1e-9
    c.send_float("exposure", 1.5)
    c.send_float("gamma", 1.8)
    n = len(c.progress)
    assert c.wait(lambda c: c._counts()[0] >= n + 3)
    assert c.progress[-1] > c.progress[n - 1]
    c.send_float("fov", 180.0)
    c.send_float("env_rotation", 45.0)
    c.send_u32("interactive_samples", 3)
    c.send_string("load_nif", "/some/where/assets.extra")
    assert c.wait(lambda c: c._counts()[2] and np.allclose(c.image(-1)[0, 0], [3, 45, np.float32(np.pi)]))
    p = c.progress
    assert 1 <= sum(1 for a, b in zip(p, p[1:]) if b < a) <= 4
    c.send_bool("stop")
    t.join(30)
    assert not t.is_alive() and res["rc"] == 0, res
    o = res["out"]
    assert o["stopped"] == "1" and 1 <= int(o["restarts"]) <= 4
    assert float(o["exposure"]) == 1.5 and abs(float(o["gamma"]) - 1.8) < 1e-6
    assert abs(float(o["fov"]) - np.pi) < 1e-6 and float(o["rot"]) == 45.0 and o["isamples"] == "3"
    assert o["nif"] == "/some/where/assets.extra"
    assert c.wait_closed(10)                                # the reference's demuxer sees our GoodBye / hang-up
    c.close()


def test_ui_heartbeat_keeps_the_references_client_alive(host):
    """An idle link: the reference's demuxer drops a peer that stays silent (it expects HeartBeat control packets,
    PacketMuxer.cpp:62-77); two 1.4 s steps with nothing to say must leave its ok() true until the server leaves."""
    import threading
    import ui_ref_client as R
    if not R.available():
        pytest.skip("oracle/_ref/libref_uiclient.so not built here (needs /root/reference)")
    port, res = _free_port(), {}
    t = threading.Thread(target=_ui_session, args=(host, port, res), kwargs=dict(steps=2, save=100, step_ms=1400))
    t.start()
    c = R.RefUiClient(port)
    import time
    time.sleep(1.2)
    assert c.ok()
    t.join(30)
    assert res["rc"] == 0 and res["out"]["steps"] == "2"
    assert c.wait(lambda c: c._counts()[2] >= 1, timeout=10)   # the final image is always delivered
    c.close()

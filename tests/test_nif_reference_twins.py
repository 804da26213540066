"""The two ends of the NIF against REFERENCE CODE. The reference has no CPU inference of the network, but it has host-side
twins of the encode and the decode (NifModel::generateInputSamples / makeCoefficients / makeGridCoordsUV and
NifModel::decodeSamples, src/neural_networks/NifModel.cpp:403-443,466-509). They are cut out of the reference source at build
time and compiled unmodified (oracle/_ref/libref_niftwins.so, oracle/nif_twins_ref_driver.cpp); their outputs are committed as
tests/golden/nif_twins.npz (generator: tests/golden/make_nif_twins_golden.py) so the pin also holds where /root/reference is absent.

Pinned by this file: feature order [sin u, sin v, cos u, cos v], coefficients 2^j, the (uv - 1) * 2 mapping, u = row / height as
the vertical coordinate, decode x * max + mean -> exp. NOT pinned (nothing in the tree can): the device graph's binary16
rounding of the argument and of the result (NifModel.cpp:211-213), which the host twin does not model - the oracle and the
GPU apply it, and the comparison is exact wherever that rounding is the identity (power-of-two grids)."""
import numpy as np
import pytest


@pytest.fixture(scope="module")
def twins(golden_dir):
    return np.load(golden_dir / "nif_twins.npz")


@pytest.fixture(scope="module")
def golden_dir():
    import pathlib
    return pathlib.Path(__file__).parent / "golden"


def test_golden_twins_match_the_compiled_reference(O, twins):
    if not O.NifTwinsReference.available():
        pytest.skip("oracle/_ref/libref_niftwins.so not built here (needs /root/reference); the golden file stands in")
    ref = O.NifTwinsReference()
    for tag in "abc":
        h, w, e = (int(x) for x in twins[f"{tag}_shape"])
        f, u, v = ref.encode_grid(h, w, e)
        assert np.array_equal(f, twins[f"{tag}_features"]) and np.array_equal(u, twins[f"{tag}_u"]) and np.array_equal(v, twins[f"{tag}_v"])
    got = ref.decode(twins["dec_x"], float(twins["dec_max"]), twins["dec_mean"], float(twins["dec_eps"]), True)
    assert np.array_equal(got, twins["dec_log"])


def test_oracle_encode_equals_reference_twin(orc, twins):
    # power-of-two grids: every argument p * 2^j is exactly representable in binary16, so the device graph's cast is the
    # identity and the oracle's binary16 features must be the reference's fp32 features rounded to binary16, bit for bit
    for tag in "ac":
        h, w, e = (int(x) for x in twins[f"{tag}_shape"])
        u, v = twins[f"{tag}_u"], twins[f"{tag}_v"]
        assert np.array_equal(u.reshape(h, w)[:, 0], np.arange(h, dtype=np.float32) / h)        # u runs down the rows (vertical)
        assert np.array_equal(v.reshape(h, w)[0], np.arange(w, dtype=np.float32) / w)
        got = orc.nif_encode(u, v, e)
        want = twins[f"{tag}_features"].astype(np.float16)
        assert np.array_equal(got.view(np.uint16), want.view(np.uint16))
    # an arbitrary grid (5 x 7): the first octaves agree to binary16 precision; the higher ones differ exactly by the
    # device graph's binary16 cast of the argument (NifModel.cpp:211-213), which the oracle applies and the host twin does not
    h, w, e = (int(x) for x in twins["b_shape"])
    u, v = twins["b_u"], twins["b_v"]
    got = orc.nif_encode(u, v, e).astype(np.float32)
    ref = twins["b_features"]
    for block in range(4):                                      # [sin u, sin v, cos u, cos v], E entries each
        assert np.abs(got[:, block * e] - ref[:, block * e]).max() < 2e-3
    p = np.stack([(u - 1) * 2, (v - 1) * 2], 1).astype(np.float32)
    for j in range(e):
        arg = (p * np.float32(2.0 ** j)).astype(np.float16).astype(np.float64)
        want = np.concatenate([np.sin(arg), np.cos(arg)], 1).astype(np.float32).astype(np.float16)
        have = np.stack([got[:, j], got[:, e + j], got[:, 2 * e + j], got[:, 3 * e + j]], 1).astype(np.float16)
        assert np.array_equal(have.view(np.uint16), want.view(np.uint16))


def test_oracle_decode_equals_reference_twin(orc, twins):
    x, mean, mx, eps = twins["dec_x"], twins["dec_mean"], float(twins["dec_max"]), np.float32(twins["dec_eps"])
    folded = (mean - eps).astype(np.float32)                    # NifMetaData.cpp:48-53 folds eps into the mean
    np.testing.assert_allclose(orc.nif_decode(x, mx, folded, True), twins["dec_log"], rtol=2e-6)
    np.testing.assert_allclose(orc.nif_decode(x, mx, mean, False), twins["dec_lin"], rtol=1e-6, atol=1e-6)


@pytest.mark.gpu
def test_gpu_encode_and_decode_equal_reference_twins(lib_built, nif_model, twins):
    from ipu_path_trace_b200 import tracer as T
    # encode: the device function the kernels use (table path and arithmetic path), on the power-of-two grids
    t = lib_built.PathTracer(16, 16)
    t.init_nif_weights(nif_model)                               # embedding dimension 12
    for tag in "ac":
        h, w, e = (int(x) for x in twins[f"{tag}_shape"])
        rows = np.stack([twins[f"{tag}_u"], twins[f"{tag}_v"]], 1).astype(np.float32)
        got = t.kat(T.KAT_NIF_FEATURES, rows, 4 * e).astype(np.float16)
        want = twins[f"{tag}_features"].astype(np.float16)
        assert np.array_equal(got.view(np.uint16), want.view(np.uint16))
    t.close()
    # decode: a one-layer network with zero kernel and bias x returns decode(x) for every ray
    x, mean, mx, eps = twins["dec_x"], twins["dec_mean"], float(twins["dec_max"]), np.float32(twins["dec_eps"])
    folded = tuple(float(m) for m in (mean - eps).astype(np.float32))
    for i in (0, 1, 2, 3):
        xi = x[i].astype(np.float16)
        model = lib_built.NifModel(12, mx, folded, True, [{"kernel": np.zeros((48, 3), np.float16), "bias": xi, "relu": False}])
        t = lib_built.PathTracer(16, 16)
        t.init_nif_weights(model)
        got = t.nif_infer(np.float32([0.25, 0.5]), np.float32([0.75, 0.125]), lib_built.NIF_SIMT)
        t.close()
        want = np.exp(xi.astype(np.float32) * np.float32(mx) + (mean - eps).astype(np.float32))
        np.testing.assert_allclose(got, np.broadcast_to(want, got.shape), rtol=3e-6)
        # and against the reference twin's own numbers where x is already a binary16 value
    xh = x.astype(np.float16).astype(np.float32)
    import os
    from oracle import oracle as O
    if O.NifTwinsReference.available():
        ref = O.NifTwinsReference().decode(xh[:4], mx, mean, float(eps), True)
        np.testing.assert_allclose(np.exp(xh[:4] * np.float32(mx) + (mean - eps).astype(np.float32)), ref, rtol=3e-6)

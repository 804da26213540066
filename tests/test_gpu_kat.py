"""Function-level known-answer tests: the device functions the wavefront kernel calls, through
b200pt_kat, against vectors from the unmodified reference (tests/golden/ref_golden.npz).
Tolerance: 1e-5 relative fp32 (north_star); integer/flag outputs exact."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL = 1e-5


@pytest.fixture(scope="module")
def pt(lib_built, nif_model):
    t = lib_built.PathTracer(64, 48)
    t.init_nif_weights(nif_model)
    yield t
    t.close()


def close(a, b, atol=1e-6):
    np.testing.assert_allclose(a, b, rtol=RTOL, atol=atol)


def test_pixel_to_ray(pt, golden):
    from ipu_path_trace_b200 import tracer as T
    n = golden["p2r_x"].size
    w, h, fov = golden["p2r_whf"]
    rows = np.stack([golden["p2r_x"], golden["p2r_y"], np.full(n, w), np.full(n, h), np.full(n, fov)], 1)
    close(pt.kat(T.KAT_PIXEL_TO_RAY, rows, 3), golden["p2r_out"])


def test_sphere_intersect(pt, golden):
    from ipu_path_trace_b200 import tracer as T
    rows = np.concatenate([golden["sph_o"], golden["sph_d"], golden["sph_c"], golden["sph_r"][:, None]], 1)
    got = pt.kat(T.KAT_SPHERE, rows, 1)[:, 0]
    want = golden["sph_t"]
    assert np.array_equal(got > 0, want > 0)
    assert (want > 0).sum() > 50 and (want == 0).sum() > 50
    close(got, want, atol=2e-6)


def test_disc_intersect(pt, golden):
    from ipu_path_trace_b200 import tracer as T
    rows = np.concatenate([golden["disc_o"], golden["disc_d"], golden["disc_n"], golden["disc_c"], golden["disc_r"][:, None]], 1)
    got = pt.kat(T.KAT_DISC, rows, 1)[:, 0]
    want = golden["disc_t"]
    assert np.array_equal(got > 0, want > 0)
    assert (want > 0).sum() > 20
    close(got, want, atol=2e-6)


def test_diffuse(pt, golden):
    from ipu_path_trace_b200 import tracer as T
    n = golden["dif_u1"].size
    rows = np.concatenate([golden["dif_n"], golden["dif_u1"][:, None], golden["dif_u2"][:, None], np.full((n, 1), golden["dif_rr"])], 1)
    got = pt.kat(T.KAT_DIFFUSE, rows, 4)
    # sinf/cosf are glibc's algorithm on the device (pt_device.cuh glibcSinCosf): the sampled direction is the reference's, bit for bit
    assert np.array_equal(got[:, :3].astype(np.float32).view(np.uint32), golden["dif_d"].astype(np.float32).view(np.uint32))
    assert np.array_equal(got[:, 3].astype(np.float32).view(np.uint32), golden["dif_w"].astype(np.float32).view(np.uint32))


def test_reflect(pt, golden):
    from ipu_path_trace_b200 import tracer as T
    got = pt.kat(T.KAT_REFLECT, np.concatenate([golden["refl_d"], golden["refl_n"]], 1), 3)
    close(got, golden["refl_out"])


def test_refract(pt, golden):
    from ipu_path_trace_b200 import tracer as T
    n = golden["refr_rnd"].size
    rows = np.concatenate([golden["refr_d"], golden["refr_n"], np.full((n, 1), golden["refr_ri"]), golden["refr_rnd"][:, None]], 1)
    got = pt.kat(T.KAT_REFRACT, rows, 4)
    assert np.array_equal(got[:, 3].astype(np.int32), golden["refr_flag"])
    assert 0 < golden["refr_flag"].sum() < n
    close(got[:, :3], golden["refr_out"])


def test_roulette(pt, golden):
    from ipu_path_trace_b200 import tracer as T
    n = golden["rou_rnd"].size
    got = pt.kat(T.KAT_ROULETTE, np.stack([golden["rou_rnd"], np.full(n, golden["rou_p"])], 1), 2)
    assert np.array_equal(got[:, 0].astype(np.int32), golden["rou_stop"])
    close(got[:, 1], golden["rou_w"])


def test_escaped_uv(pt, golden, O):
    from ipu_path_trace_b200 import tracer as T
    rec = np.ascontiguousarray(golden["img_records"]).view(O.CONTRIB).reshape(golden["img_records"].shape[0], -1)
    # the ESCAPED record carries the unit direction in clr (codelets.cpp:181-186)
    esc_rows = [(i, j) for i in range(rec.shape[0]) for j in range(rec.shape[1]) if rec[i, j]["type"] == O.C_ESCAPED]
    idx = np.array([i for i, _ in esc_rows]); d = np.stack([rec[i, j]["clr"] for i, j in esc_rows])
    rows = np.concatenate([d, np.full((len(idx), 1), golden["img_azimuth"])], 1)
    got = pt.kat(T.KAT_ESCAPED_UV, rows, 2)
    # against the compiled reference (codelets.cpp:333-347 with glibc's acosf / double atan2): the same bits
    assert np.array_equal(got[:, 0].astype(np.float32).view(np.uint32), golden["img_u"][idx].astype(np.float32).view(np.uint32))
    assert np.array_equal(got[:, 1].astype(np.float32).view(np.uint32), golden["img_v"][idx].astype(np.float32).view(np.uint32))
    # |d.y| = 1 + 1 ulp after normalisation: the reference's acosf gives NaN and the pixel is lost for ever; clamped here
    edge = np.array([[1e-4, np.nextafter(np.float32(1), np.float32(2)), 0.0, 0.0], [0.0, -np.nextafter(np.float32(1), np.float32(2)), 1e-4, 0.0],
                     [0.0, 1.0, 0.0, 0.0], [0.0, -1.0, 0.0, 0.0]], np.float32)
    uv = pt.kat(T.KAT_ESCAPED_UV, edge, 2)
    assert np.isfinite(uv).all() and uv[0, 0] == 0.0 and uv[2, 0] == 0.0 and abs(uv[1, 0] - 1.0) < 1e-6 and abs(uv[3, 0] - 1.0) < 1e-6
    # ... and the oracle stays reference-exact: NaN for the two out-of-domain rows, the GPU's bits for the two in-domain rows
    erec = np.zeros((4, 1), O.CONTRIB)
    erec["clr"][:, 0, :] = edge[:, :3]; erec["weight"] = 1.0; erec["type"] = O.C_ESCAPED
    ou, ov = O.Oracle().escaped_uv(erec, 0.0)
    assert np.isnan(ou[0]) and np.isnan(ou[1])
    assert np.array_equal(ou[2:].view(np.uint32), uv[2:, 0].astype(np.float32).view(np.uint32))
    assert np.array_equal(ov.view(np.uint32), uv[:, 1].astype(np.float32).view(np.uint32))


def test_philox_bit_exact(pt, orc):
    from ipu_path_trace_b200 import tracer as T
    rng = np.random.default_rng(11)
    words = rng.integers(0, 2**32, (256, 6), dtype=np.uint64).astype(np.uint32)
    words[0] = 0
    words[1] = 0xffffffff
    got = pt.kat(T.KAT_PHILOX, words.view(np.float32), 4).view(np.uint32)
    want = np.stack([orc.philox(w[:4], w[4:]) for w in words])
    assert np.array_equal(got, want)
    # Random123 known answers (Salmon et al., kat_vectors): philox4x32-10
    assert [hex(x) for x in want[0]] == ["0x6627e8d5", "0xe169c58d", "0xbc57ac4c", "0x9b00dbd8"]
    assert [hex(x) for x in want[1]] == ["0x408f276d", "0x41c83b0e", "0xa20bc7c6", "0x6d5451fd"]


@pytest.mark.parametrize("noise_type", [0, 1, 2])
def test_aa_noise_bit_exact(pt, orc, noise_type):
    from ipu_path_trace_b200 import tracer as T
    rng = np.random.default_rng(12)
    w = rng.integers(0, 2**32, (4096, 2), dtype=np.uint64).astype(np.uint32)
    rows = np.concatenate([w, np.full((len(w), 1), noise_type, np.uint32)], 1)
    got = pt.kat(T.KAT_AA_NOISE, rows.view(np.float32), 2)
    want = orc.half_to_float(orc.aa_noise(w.reshape(-1), noise_type)).reshape(-1, 2)
    assert np.array_equal(got, want)
    if noise_type == 0:
        assert abs(got.std() - 1.0) < 0.03 and abs(got.mean()) < 0.03
    if noise_type == 2:
        assert np.abs(got).max() <= 3.0


def test_nif_features(pt, orc):
    """Fourier features (NifModel.cpp:185-218): binary16 argument rounding reproduced exactly;
    the sin/cos values may differ from glibc's by at most one binary16 ulp, rarely."""
    from ipu_path_trace_b200 import tracer as T
    rng = np.random.default_rng(13)
    uv = rng.random((8192, 2)).astype(np.float32)
    uv[:3] = [[0, 0], [1, 1], [0.5, 0.25]]
    got = pt.kat(T.KAT_NIF_FEATURES, uv, 48)
    want = orc.nif_encode(uv[:, 0], uv[:, 1], 12).astype(np.float32)
    assert (got == want).mean() > 0.999
    assert np.max(np.abs(got - want)) <= 2.0 ** -10

"""Pins the CPU oracle (oracle/liboracle.so) against vectors produced by the unmodified reference
codelets (tests/golden/ref_golden.npz, made by tests/golden/make_golden.py), and, when the
compiled reference itself is present (oracle/_ref), against it directly on fresh inputs."""
import numpy as np
import pytest


def _records(O, g):
    return np.ascontiguousarray(g["img_records"]).view(O.CONTRIB).reshape(g["img_records"].shape[0], -1)


def test_camera_rays_match_reference_golden(O, orc, golden):
    W, H = (int(x) for x in golden["img_wh"])
    s = O.Settings(width=W, height=H)
    rays = orc.camera_rays(O.make_worklist(W, H), golden["img_noise"].reshape(-1), s)
    assert np.array_equal(rays, golden["img_rays"])


def test_path_records_match_reference_golden(O, orc, golden):
    W, H = (int(x) for x in golden["img_wh"])
    s = O.Settings(width=W, height=H)
    rec = orc.trace_records(golden["img_rays"], golden["img_uniforms"].reshape(-1), 30, s)
    assert rec.tobytes() == _records(O, golden).tobytes()      # bit-exact, every Contribution


def test_escaped_uv_match_reference_golden(O, orc, golden):
    u, v = orc.escaped_uv(_records(O, golden), float(golden["img_azimuth"]))
    assert np.array_equal(u, golden["img_u"]) and np.array_equal(v, golden["img_v"])


def test_accumulate_match_reference_golden(O, orc, golden):
    W, H = (int(x) for x in golden["img_wh"])
    out = orc.light_and_accumulate(_records(O, golden), golden["img_bgr"], O.make_worklist(W, H))
    want = np.ascontiguousarray(golden["img_accum"]).view(O.TRACE_RECORD).reshape(-1)
    assert out.tobytes() == want.tobytes()


def test_forward_throughput_equals_backward_fold(O, orc, golden):
    """The CUDA path carries an RGB throughput forward instead of folding the record stack
    backwards (codelets.cpp:255-298); path_summary is that forward form. It must give the
    reference's accumulated pixel within fp32 rounding."""
    rec = _records(O, golden)
    esc, d, thr, length = orc.path_summary(rec)
    bgr = golden["img_bgr"]
    want = np.ascontiguousarray(golden["img_accum"]).view(O.TRACE_RECORD).reshape(-1)
    rgb = bgr[:, ::-1] * thr
    got = np.where(esc[:, None] == 1, rgb, 0)
    ref = np.stack([want["r"], want["g"], want["b"]], 1)
    np.testing.assert_allclose(got, ref, rtol=1e-5, atol=1e-7)
    assert np.array_equal(length, want["pathLength"])


def test_oracle_vs_compiled_reference_fresh_inputs(O, orc):
    if not O.Reference.available():
        pytest.skip("oracle/_ref not built here (no /root/reference): golden vectors cover it")
    ref = O.Reference()
    rng = np.random.default_rng(7)
    for (W, H, kw) in [(96, 64, {}), (40, 56, dict(roulette_depth=1, stop_prob=0.5, max_path_length=6)),
                       (33, 17, dict(fov_degrees=60.0, refractive_index=1.33, aa_noise_scale=0.7))]:
        s = O.Settings(width=W, height=H, **kw)
        items = O.make_worklist(W, H)
        noise = O.half_bits(rng.standard_normal((len(items), 2)).astype(np.float32))
        per = 3 * s.max_path_length
        uni = O.half_bits(rng.random((len(items), per)).astype(np.float32))
        a = orc.camera_rays(items, noise.reshape(-1), s)
        assert np.array_equal(a, ref.camera_rays(items, noise, s))
        ra = orc.trace_records(a, uni.reshape(-1), per, s)
        rb = ref.ray_trace(a, uni.reshape(-1), s, per_path=per)
        assert ra.tobytes() == rb.tobytes()
        ua, va = orc.escaped_uv(ra, 1.1)
        ub, vb, _ = ref.preprocess_escaped(rb, 1.1)
        assert np.array_equal(ua, ub) and np.array_equal(va, vb)
        bgr = rng.uniform(0, 2, (len(items), 3)).astype(np.float32)
        assert orc.light_and_accumulate(ra, bgr, items).tobytes() == ref.postprocess_and_accumulate(rb, bgr, items).tobytes()


def test_escape_statistics_match_survey_probe(O, orc):
    """320x240 -> escaped fraction ~0.978, mean records ~1.378."""
    s = O.Settings(width=320, height=240)
    esc, uv, thr, length, cam = orc.trace_philox(O.make_worklist(320, 240), 0, s)
    assert abs(esc.mean() - 0.9777) < 0.004
    assert abs(length.mean() - 1.378) < 0.02


def test_escaped_uv_out_of_domain_is_the_references_nan(O, orc):
    """codelets.cpp:333 `acosf(dir.y)`: a normalised direction can have |y| = 1 + 1 ulp, for which the reference's
    result is NaN. The oracle restates exactly that (the oracle does not follow the product's clamp);
    where oracle/_ref exists the compiled reference must agree bit for bit, NaN included."""
    one_up = np.nextafter(np.float32(1), np.float32(2))
    d = np.array([[1e-4, one_up, 0.0], [0.0, -one_up, 1e-4], [0.0, 1.0, 0.0], [0.0, -1.0, 0.0], [0.6, 0.0, 0.8]], np.float32)
    rec = np.zeros((len(d), 1), O.CONTRIB)
    rec["clr"][:, 0, :] = d; rec["weight"] = 1.0; rec["type"] = O.C_ESCAPED
    u, v = orc.escaped_uv(rec, 0.25)
    assert np.isnan(u[0]) and np.isnan(u[1]) and np.all(np.isfinite(u[2:])) and np.all(np.isfinite(v))
    assert u[2] == 0.0 and abs(u[3] - 1.0) < 1e-6 and u[4] == 0.5
    if O.Reference.available():
        ur, vr, _ = O.Reference().preprocess_escaped(rec, 0.25)
        assert np.array_equal(u.view(np.uint32) & 0x7fffffff, ur.view(np.uint32) & 0x7fffffff)   # NaN sign aside
        assert np.array_equal(v.view(np.uint32), vr.view(np.uint32))

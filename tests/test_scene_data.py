"""The scene as data and emissive materials: b200pt_set_scene / orc_set_scene, EMIT records
(codelets.cpp:192-196, backward fold :255-298). The reference hard-codes its scene and has its only emitter commented out
(codelets.cpp:117,142), so the compiled reference pins the built-in scene only; EMIT is checked GPU <-> oracle restatement."""
import numpy as np
import pytest


def _lit_scene(O):
    """The reference scene + its commented-out light, moved close and enlarged so that paths actually end on it, + an
    emissive disc overhead and a second glass sphere: 9 primitives."""
    s = O.reference_scene(with_light=False)
    s.insert(5, dict(shape="sphere", centre=(2.5, 1.5, -5.3), radius=1.2, colour=(0, 0, 0), emission=(30, 25, 20), bsdf="diffuse"))
    s.append(dict(shape="disc", centre=(0, 4.0, -4.0), normal=(0, -2, 0), radius=2.0, colour=(0, 0, 0), emission=(3, 4, 5), bsdf="diffuse"))
    s.append(dict(shape="sphere", centre=(-0.9, -1.2, -2.2), radius=0.3, colour=(.9, .95, 1.0), bsdf="refractive"))
    return s


def test_oracle_scene_as_data_equals_builtin(O, orc):
    s = O.Settings(width=96, height=64, seed=3)
    it = O.make_worklist(96, 64)
    a = orc.trace_philox(it, 1, s)
    try:
        orc.set_scene(O.reference_scene())
        b = orc.trace_philox(it, 1, s)
        assert all(np.array_equal(x, y) for x, y in zip(a, b))
        orc.set_scene(_lit_scene(O))
        esc, uv, thr, length, cam = orc.trace_philox(it, 1, s)
        assert (esc == 2).sum() > 100 and (esc == 1).sum() > 1000 and (esc == 0).sum() > 0
        assert np.all(thr[esc == 2].max(1) > 0) and np.all(uv[esc == 2] == 0)
        assert np.array_equal(cam, a[4])                        # the camera does not depend on the scene
    finally:
        orc.set_scene(None)
    c = orc.trace_philox(it, 1, s)
    assert all(np.array_equal(x, y) for x, y in zip(a, c))      # restored


@pytest.mark.gpu
@pytest.mark.parametrize("variant", [None, "v2"])
def test_gpu_scene_as_data_and_emitters_match_oracle(lib_built, nif_model, O, orc, oracle_nif, monkeypatch, variant):
    if variant:
        monkeypatch.setenv("B200PT_TRACE", variant)
    else:
        monkeypatch.delenv("B200PT_TRACE", raising=False)
    W, H, spp = 160, 120, 8
    s = O.Settings(width=W, height=H, seed=9)
    items = lib_built.create_work_list_for_image(W, H)
    t = lib_built.PathTracer(W, H)
    t.init_nif_weights(nif_model)
    t.init_render_settings(seed=9, samples_per_step=spp)
    t.setup(items)
    try:
        # 1. the built-in scene handed over as data gives the built-in scene's bits
        base = t.trace_wavefront(2)
        t.set_scene(O.reference_scene())
        again = t.trace_wavefront(2)
        assert all(np.array_equal(x, y) for x, y in zip(base, again))
        # 2. a scene with emitters: per-path flags (0 ended, 1 escaped, 2 emitted), lengths and uv bit-exact, radiance 1e-5
        scene = _lit_scene(O)
        t.set_scene(scene)
        orc.set_scene(scene)
        esc, uv, thr, length, cam = t.trace_wavefront(2)
        e2, uv2, t2, l2, cam2 = orc.trace_philox(items, 2, s)
        assert np.array_equal(esc, e2) and np.array_equal(length, l2) and np.array_equal(cam, cam2)
        assert (esc == 2).sum() > 500
        m = esc == 1
        assert np.array_equal(uv[m].view(np.uint32), uv2[m].astype(np.float32).view(np.uint32))
        np.testing.assert_allclose(thr[esc > 0], t2[esc > 0], rtol=1e-5, atol=1e-7)
        # 3. the whole program: image with emitters against the oracle render
        t.path_trace()
        out, ctr = t.read_results()
        want, n_esc, _, _ = orc.render(items, 0, spp, s, oracle_nif, mode=0)
        assert ctr.escaped == n_esc
        assert np.array_equal(out["pathLength"], want["pathLength"])
        a = np.stack([out["r"], out["g"], out["b"]], 1).astype(np.float64)
        b = np.stack([want["r"], want["g"], want["b"]], 1).astype(np.float64)
        assert np.sqrt(np.mean((a - b) ** 2)) / np.mean(b) < 0.01
        assert np.isclose(a, b, rtol=2e-2, atol=1e-3).all(1).mean() > 0.995
        assert b.max() > 30 * spp * 0.5                          # emitter pixels are in the image
        # 4. error contract
        with pytest.raises(lib_built.B200PTError, match="exceed the limit"):
            t.set_scene([scene[0]] * 17)
        with pytest.raises(lib_built.B200PTError, match="non-positive radius"):
            t.set_scene([dict(shape="sphere", centre=(0, 0, -3), radius=0.0)])
        # 5. back to the built-in scene
        t.set_scene(None)
        assert all(np.array_equal(x, y) for x, y in zip(base, t.trace_wavefront(2)))
    finally:
        orc.set_scene(None)
        t.close()

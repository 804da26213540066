"""Stage-level parity of the path-tracing kernels (S0-S4 of SURVEY.md 2.2) through the C ABI."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _mk(lib_built, nif_model, W, H, **kw):
    t = lib_built.PathTracer(W, H, **kw)
    t.init_nif_weights(nif_model)
    return t


def test_trace_fixed_matches_reference_records(lib_built, nif_model, golden, O, orc):
    """RayTraceKernel (codelets.cpp:93-227) on the reference's own inputs: escape flag, record
    count and escaped direction exact for every path; throughput within 1e-5 relative (the fold order differs)."""
    W, H = (int(x) for x in golden["img_wh"])
    t = _mk(lib_built, nif_model, W, H)
    t.init_render_settings()
    rec = np.ascontiguousarray(golden["img_records"]).view(O.CONTRIB).reshape(W * H, -1)
    esc_w, dir_w, thr_w, len_w = orc.path_summary(rec)
    esc, d, thr, length = t.trace_fixed(golden["img_rays"], golden["img_uniforms"], 30)
    agree = (esc == esc_w) & (length == len_w)
    assert agree.all()                   # every branch of every path: sinf/cosf are glibc's, bit for bit (pt_device.cuh)
    m = agree & (esc == 1)
    assert np.array_equal(d[m].view(np.uint32), dir_w[m].astype(np.float32).view(np.uint32))      # escaped directions: the reference's bits
    np.testing.assert_allclose(thr[m], thr_w[m], rtol=1e-5, atol=1e-7)
    assert np.all(thr[esc == 0] == 0)
    t.close()


@pytest.mark.parametrize("cfg", [
    dict(W=320, H=240), dict(W=97, H=61, roulette_depth=1, stop_prob=0.5, max_path_length=5),
    dict(W=128, H=64, aa_noise_type=1, refractive_index=1.33), dict(W=64, H=64, aa_noise_type=2),
])
def test_wavefront_matches_oracle(lib_built, nif_model, O, orc, cfg):
    """Philox stream -> AA noise -> GenerateCameraRays -> path -> escaped uv, one sample, against
    the oracle on the same counters: camera rays, flags, lengths and escaped-ray uv bit-exact."""
    cfg = dict(cfg)
    W, H = cfg.pop("W"), cfg.pop("H")
    s = O.Settings(width=W, height=H, seed=77, env_map_rotation_degrees=33.0, **cfg)
    t = _mk(lib_built, nif_model, W, H, **cfg)
    t.init_render_settings(seed=77, env_map_rotation_degrees=33.0)
    items = O.make_worklist(W, H)
    t.setup(items)
    for sample in (0, 5):
        esc, uv, thr, length, cam = t.trace_wavefront(sample)
        e2, uv2, t2, l2, cam2 = orc.trace_philox(items, sample, s)
        assert np.array_equal(cam, cam2)
        agree = (esc == e2) & (length == l2)
        assert agree.all()
        m = agree & (esc == 1)
        assert np.array_equal(uv[m].view(np.uint32), uv2[m].astype(np.float32).view(np.uint32))      # acosf is glibc's algorithm on the device
        np.testing.assert_allclose(thr[m], t2[m], rtol=1e-5, atol=1e-7)
    t.close()


def test_padding_items_are_skipped(lib_built, nif_model, O):
    """u = v = 65535 marks null work (LoadBalancer.cpp:66-71)."""
    W, H = 40, 24
    t = _mk(lib_built, nif_model, W, H, max_worklist=1024)
    t.init_render_settings(samples_per_step=3)
    items = lib_built.pad_work_list(lib_built.create_work_list_for_image(W, H), 1024)
    t.setup(items)
    t.path_trace()
    out, ctr = t.read_results()
    assert ctr.paths == W * H * 3
    pad = out["u"] == 65535
    assert pad.sum() == 1024 - W * H
    assert np.all(out["sampleCount"][pad] == 0) and np.all(out["r"][pad] == 0)
    assert np.all(out["sampleCount"][~pad] == 3)
    t.close()


def test_empty_worklist(lib_built, nif_model):
    t = _mk(lib_built, nif_model, 16, 16)
    t.init_render_settings(samples_per_step=2)
    t.setup(np.zeros(0, lib_built.TRACE_RECORD))
    t.path_trace()
    out, ctr = t.read_results()
    assert len(out) == 0 and ctr.paths == 0 and ctr.escaped == 0
    t.close()


def test_call_order_errors(lib_built, nif_model):
    t = lib_built.PathTracer(16, 16)
    t.init_render_settings()
    t.setup(lib_built.create_work_list_for_image(16, 16))
    with pytest.raises(lib_built.B200PTError):
        t.path_trace()                      # no NIF weights uploaded
    t.close()
    t = lib_built.PathTracer(16, 16)
    t.init_nif_weights(nif_model)
    with pytest.raises(lib_built.B200PTError):
        t.path_trace()                      # render settings not set
    with pytest.raises(lib_built.B200PTError):
        t.setup(lib_built.create_work_list_for_image(32, 32))   # exceeds max_worklist
    t.close()
    with pytest.raises(lib_built.B200PTError):
        lib_built.PathTracer(16, 16, roulette_depth=0)


@pytest.mark.parametrize("variant", ["v2", "v2p3"])
def test_live_ray_compaction_kernel_equals_default(lib_built, nif_model, monkeypatch, variant):
    """B200PT_TRACE=v2: the trace kernel with live-ray compaction (warp ballot/prefix into a per-block live list) and path
    regeneration (a lane takes the next live ray when its path ends) must reproduce the default kernel: per-item
    sampleCount and pathLength, escaped / segment counters bit for bit, radiance up to the order of the fp32 atomics -
    on a ragged work list with padding items, 9 samples per step (groups of 4 + 1) and two steps."""
    W, H, sps = 200, 130, 9
    wl = lib_built.create_work_list_for_image(W, H)
    pad = np.zeros(37, dtype=wl.dtype); pad["u"] = 65535; pad["v"] = 65535
    wl = np.concatenate([wl[:777], pad, wl[777:]])

    def run(tv):
        if tv:
            monkeypatch.setenv("B200PT_TRACE", tv)
        else:
            monkeypatch.delenv("B200PT_TRACE", raising=False)
        t = lib_built.PathTracer(W, H, max_worklist=len(wl))
        t.init_nif_weights(nif_model)
        t.init_render_settings(seed=5, samples_per_step=sps)
        t.setup(wl)
        t.path_trace(); t.path_trace()
        out, ctr = t.read_results()
        t.close()
        return out, ctr
    a, ca = run(None)
    b, cb = run(variant)
    assert (ca.paths, ca.escaped, ca.path_segments) == (cb.paths, cb.escaped, cb.path_segments)
    assert ca.paths == W * H * sps                      # counters are those of the last path_trace
    assert np.array_equal(a["sampleCount"], b["sampleCount"]) and np.array_equal(a["pathLength"], b["pathLength"])
    assert np.all(b["sampleCount"][777:777 + 37] == 0)
    rgb = lambda x: np.stack([x["r"], x["g"], x["b"]], 1).astype(np.float64)
    np.testing.assert_allclose(rgb(a), rgb(b), rtol=1e-5, atol=1e-6)

"""End-to-end parity of the five programs (PathTracerApp.cpp:479-483) against the oracle render
on the same Philox streams, plus size-independent properties at the BASELINE size."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _rgb(items):
    return np.stack([items["r"], items["g"], items["b"]], 1).astype(np.float64)


def _rrmse(a, b):
    return float(np.sqrt(np.mean((a - b) ** 2)) / np.mean(b))


@pytest.mark.parametrize("impl", ["simt", "tc"])
def test_render_matches_oracle(lib_built, nif_model, O, orc, oracle_nif, impl):
    """C1-like: 160x120 at 16 spp (oracle finishes in seconds). Same RNG counters on both sides,
    so accumulators compare pixel by pixel; converged-image gate: relative RMSE < 1 %."""
    W, H, spp = 160, 120, 16
    nif_impl = lib_built.NIF_SIMT if impl == "simt" else lib_built.NIF_TCGEN05
    t = lib_built.PathTracer(W, H, nif_impl=nif_impl)
    t.init_nif_weights(nif_model)
    t.init_render_settings(seed=1, samples_per_step=spp)
    items = lib_built.create_work_list_for_image(W, H)
    t.setup(items)
    t.path_trace()
    out, ctr = t.read_results()
    want, esc, _, _ = orc.render(items, 0, spp, O.Settings(width=W, height=H), oracle_nif, mode=0)
    assert ctr.paths == W * H * spp
    assert abs(ctr.escaped - esc) <= 1e-3 * esc
    assert np.all(out["sampleCount"] == spp)
    assert abs(int(out["pathLength"].sum()) - int(want["pathLength"].sum())) <= 1e-3 * want["pathLength"].sum()
    assert ctr.path_segments == int(out["pathLength"].sum())
    a, b = _rgb(out), _rgb(want)
    assert _rrmse(a, b) < 0.01
    close = np.isclose(a, b, rtol=2e-2, atol=1e-3).all(1)
    assert close.mean() > 0.995
    t.close()


def test_film_matches_accumulated_image(lib_built, nif_model):
    """AccumulatedImage::accumulate / updateLdrImage (AccumulatedImage.cpp:59-74,23-46):
    hdr(v,u) += (b,g,r)/sampleCount per step; ldr = (hdr*2^exposure)^(1/gamma)*255."""
    W, H = 96, 64
    t = lib_built.PathTracer(W, H)
    t.init_nif_weights(nif_model)
    t.init_render_settings(samples_per_step=4)
    items = lib_built.create_work_list_for_image(W, H)
    t.setup(items)
    hdr = np.zeros((H, W, 3), np.float32)
    steps = 3
    for _ in range(steps):
        t.path_trace()
        out, _ = t.read_results()
        hdr[out["v"], out["u"]] += np.stack([out["b"], out["g"], out["r"]], 1) / out["sampleCount"][:, None].astype(np.float32)
        t.film_accumulate()
        out2, _ = t.read_results()
        assert np.all(out2["sampleCount"] == 0) and np.all(out2["r"] == 0)     # clearInactiveAccumulators
    film = t.film_read(1.0 / steps)
    np.testing.assert_allclose(film, hdr / steps, rtol=1e-5, atol=1e-7)
    ldr = t.film_tonemap(1.0 / steps, exposure=0.5, gamma=2.2)
    want = np.clip(np.power((hdr / steps) * np.float32(2.0 ** 0.5), np.float32(1 / 2.2)) * 255.0, 0, 255)
    assert np.abs(ldr.astype(np.float32) - want).max() <= 1.01
    t.film_clear()
    assert np.all(t.film_read() == 0)
    t.close()


def test_sample_index_continues_and_is_shard_invariant(lib_built, nif_model):
    """The image must not depend on how pixels are dealt to GPUs (RNG keyed on pixel and sample)
    nor on how spp are split into steps."""
    W, H = 128, 80
    full = lib_built.create_work_list_for_image(W, H)

    def render(worklists, steps, sps):
        acc = np.zeros((H, W, 3), np.float64)
        for wl in worklists:
            t = lib_built.PathTracer(W, H, nif_impl=lib_built.NIF_SIMT)
            t.init_nif_weights(nif_model)
            t.init_render_settings(samples_per_step=sps)
            t.setup(wl)
            for _ in range(steps):
                t.path_trace()
            out, _ = t.read_results()
            acc[out["v"], out["u"]] += _rgb(out)
            t.close()
        return acc
    a = render([full], 1, 8)
    b = render(lib_built.shard_tiles(W, H, 3), 2, 4)
    np.testing.assert_allclose(a, b, rtol=1e-4, atol=1e-5)


def test_baseline_size_properties(lib_built, nif_model):
    """C2 size (1104x1000): counts, escape statistics of SURVEY.md 8(d), finiteness, determinism of
    the integer outputs. Oracle cannot run this size in seconds, so properties only."""
    W, H, sps = 1104, 1000, 4
    t = lib_built.PathTracer(W, H)
    t.init_nif_weights(nif_model)
    t.init_render_settings(samples_per_step=sps)
    t.setup(lib_built.create_work_list_for_image(W, H))
    t.path_trace()
    out, ctr = t.read_results()
    assert ctr.paths == W * H * sps
    assert abs(ctr.escaped / ctr.paths - 0.9732) < 0.003
    assert abs(ctr.path_segments / ctr.paths - 1.456) < 0.01
    assert np.all(out["sampleCount"] == sps)
    rgb = _rgb(out)
    assert np.all(np.isfinite(rgb)) and rgb.min() >= 0 and rgb.mean() > 0
    t.set_sample_index(0)
    t.film_accumulate()
    t.path_trace()
    out2, ctr2 = t.read_results()
    assert np.array_equal(out2["pathLength"], out["pathLength"]) and ctr2.escaped == ctr.escaped
    np.testing.assert_allclose(_rgb(out2), rgb, rtol=1e-4, atol=1e-5)
    t.close()


def test_uint16_wrap_like_reference(lib_built, nif_model):
    """TraceRecord.sampleCount is uint16 (TraceRecord.hpp:8-11): read_results wraps like the reference."""
    t = lib_built.PathTracer(8, 8)
    t.init_nif_weights(nif_model)
    t.init_render_settings(samples_per_step=65536 + 3)
    t.setup(lib_built.create_work_list_for_image(8, 8))
    t.path_trace()
    out, ctr = t.read_results()
    assert ctr.paths == 64 * (65536 + 3)
    assert np.all(out["sampleCount"] == 3)
    t.close()


def test_setup_next_and_pinned_buffers_equal_the_serial_programs(lib_built, nif_model):
    """b200pt_setup_next (the reference's second work-list buffer, PathTracerApp.cpp:703-709) staged while path_trace runs,
    with page-locked host buffers (b200pt_host_alloc), must give exactly the records of the serial setup / path_trace /
    read_results sequence - including a different list at every step (here: the two halves of the image alternate)."""
    W, H, sps = 96, 64, 4
    full = lib_built.create_work_list_for_image(W, H)
    lists = [full[: len(full) // 2], full[len(full) // 2:]]

    def serial():
        t = lib_built.PathTracer(W, H, nif_impl=lib_built.NIF_SIMT)
        t.init_nif_weights(nif_model); t.init_render_settings(samples_per_step=sps)
        outs = []
        for k in range(4):
            t.setup(lists[k & 1]); t.path_trace(); outs.append(t.read_results()[0].copy())
        t.close()
        return outs

    def pipelined():
        t = lib_built.PathTracer(W, H, nif_impl=lib_built.NIF_SIMT)
        t.init_nif_weights(nif_model); t.init_render_settings(samples_per_step=sps)
        pins = [lib_built.PinnedRecords(len(l)) for l in lists]
        for p, l in zip(pins, lists):
            p.array[:] = l
        out = lib_built.PinnedRecords(len(lists[0]))
        outs = []
        t.setup(pins[0].array)
        for k in range(4):
            t.path_trace(sync=False)
            if k < 3:
                t.setup_next(pins[(k + 1) & 1].array)
            outs.append(t.read_results(out.array)[0].copy())
        t.close()
        for p in pins + [out]:
            p.close()
        return outs
    a, b = serial(), pipelined()
    for x, y in zip(a, b):
        assert np.array_equal(x["u"], y["u"]) and np.array_equal(x["v"], y["v"])
        assert np.array_equal(x["pathLength"], y["pathLength"]) and np.array_equal(x["sampleCount"], y["sampleCount"])
        np.testing.assert_allclose(_rgb(x), _rgb(y), rtol=1e-5, atol=1e-6)
    assert not np.array_equal(a[0]["v"], a[1]["v"])


def test_setup_next_waits_for_the_unpack_of_the_previous_staged_list(lib_built, nif_model):
    """setup_next(k+1) may be called while the kernel that unpacks staged list k is still QUEUED behind earlier work on the
    context's stream; the upload must not overwrite that list (both use the one staging buffer). The stream is kept busy
    with film clears of a 4K film so that the unpack of list B has certainly not run when list C is staged."""
    W, H, sps = 3840, 2160, 1
    full = lib_built.create_work_list_for_image(W, H)
    n = 4096
    A, B, Cc = (full[i * 100000: i * 100000 + n] for i in range(3))
    t = lib_built.PathTracer(W, H, nif_impl=lib_built.NIF_SIMT)
    t.init_nif_weights(nif_model); t.init_render_settings(samples_per_step=sps)
    pins = [lib_built.PinnedRecords(n) for _ in range(3)]
    for p, l in zip(pins, (A, B, Cc)):
        p.array[:] = l
    t.setup(pins[0].array)
    t.path_trace(sync=False)
    t.setup_next(pins[1].array)
    got_a = t.read_results()[0].copy()
    for _ in range(400):                      # ~100 MB memset each, asynchronous: the stream is busy for milliseconds
        t.film_clear()
    t.path_trace(sync=False)                  # queues unpack(B) behind the clears
    t.setup_next(pins[2].array)               # must wait for that unpack before it overwrites the staging buffer
    got_b = t.read_results()[0].copy()
    t.path_trace(sync=False)
    got_c = t.read_results()[0].copy()
    t.close()
    for p in pins:
        p.close()
    for got, want in ((got_a, A), (got_b, B), (got_c, Cc)):
        assert np.array_equal(got["u"], want["u"]) and np.array_equal(got["v"], want["v"])
        assert np.all(got["sampleCount"] == sps)


def test_two_runs_give_the_same_bits(lib_built, nif_model):
    """The reference folds a pixel's samples in a fixed order (codelets.cpp:235-306): its image is a function of the seed.
    Here samples finish in no particular order, so radiance is accumulated as 64-bit fixed point with integer atomics
    (pt_common.cuh AccumSlot): two renders - also with different wavefront sizes and with the two stream schedules - must
    give bit-identical records and films."""
    W, H, sps = 160, 96, 24
    outs = []
    for kw in (dict(), dict(), dict(samples_per_launch=5), dict(samples_per_launch=24)):
        t = lib_built.PathTracer(W, H, **kw)
        t.init_nif_weights(nif_model); t.init_render_settings(seed=3, samples_per_step=sps)
        t.setup(lib_built.create_work_list_for_image(W, H))
        t.path_trace()
        rec = t.read_results()[0].copy()
        t.film_accumulate()
        outs.append((rec, t.film_read()))
        t.close()
    for rec, film in outs[1:]:
        assert np.array_equal(rec, outs[0][0]) and np.array_equal(film, outs[0][1])
    assert np.isfinite(outs[0][1]).all() and outs[0][1].max() > 0


def test_c5_size_properties(lib_built, nif_model):
    """BASELINE config C5's image (3840x2160 = 8.3 M work items, the largest single-GPU work list): counts, escape
    statistics, finiteness, and the film of one step equals the records (size-independent properties; the oracle needs
    minutes at this size)."""
    W, H, sps = 3840, 2160, 2
    t = lib_built.PathTracer(W, H)
    t.init_nif_weights(nif_model)
    t.init_render_settings(samples_per_step=sps)
    t.setup(lib_built.create_work_list_for_image(W, H))
    t.path_trace()
    out, ctr = t.read_results()
    assert ctr.paths == W * H * sps
    assert 0.96 < ctr.escaped / ctr.paths < 0.985
    assert np.all(out["sampleCount"] == sps) and int(out["pathLength"].sum()) == ctr.path_segments
    rgb = _rgb(out)
    assert np.all(np.isfinite(rgb)) and rgb.min() >= 0
    t.film_accumulate()
    film = t.film_read(1.0)
    want = np.zeros((H, W, 3), np.float32)
    want[out["v"], out["u"]] = np.stack([out["b"], out["g"], out["r"]], 1) / np.float32(sps)
    np.testing.assert_allclose(film, want, rtol=1e-6, atol=1e-8)
    t.close()

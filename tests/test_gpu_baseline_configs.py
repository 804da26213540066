"""Parity on the BASELINE.json configurations themselves (VERDICT r1 "next round" item 1), through the C ABI:

  C1  320x240 at 16 spp, full render on the tcgen05 path, pixel by pixel against the CPU oracle;
  C2  1104x1000 (the headline size) at 2 spp against the oracle, same assertions + exact pathLength per pixel;
  C4  1 048 576 unit directions uniform on the sphere (Philox seed 1234) -> uv -> tcgen05 MLP against the
      oracle's NIF (mode 0) on EVERY direction (not against the SIMT kernel);
  independent streams: a converged GPU image (seed A) against the oracle on other seeds, so a statistically
      wrong estimator shared by both sides through identical random numbers could not hide.

The oracle legs take 5-30 s of host time each (0.1-0.3 M samples/s on the GPU box's cores). NIF weights are the
seeded synthetic ones (the shipped blob is missing from the snapshot): NIF parity stays "unpinned" in that sense."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _rgb(items):
    return np.stack([items["r"], items["g"], items["b"]], 1).astype(np.float64)


def _rrmse(a, b):
    return float(np.sqrt(np.mean((a - b) ** 2)) / np.mean(b))


def _render_gpu(pt, model, W, H, spp, seed=1, sps=None):
    t = pt.PathTracer(W, H, nif_impl=pt.NIF_TCGEN05)
    t.init_nif_weights(model)
    sps = sps or spp
    t.init_render_settings(seed=seed, samples_per_step=sps)
    items = pt.create_work_list_for_image(W, H)
    t.setup(items)
    acc = np.zeros((W * H, 3), np.float64)
    length = np.zeros(W * H, np.int64)
    escaped = paths = 0
    for _ in range(spp // sps):
        t.path_trace()
        out, ctr = t.read_results()
        assert ctr.nif_impl_used == pt.NIF_TCGEN05
        assert np.all(out["sampleCount"] == (sps & 0xffff))
        acc += _rgb(out)
        length += out["pathLength"]
        escaped += ctr.escaped; paths += ctr.paths
        t.film_accumulate()                      # clears the accumulators like clearInactiveAccumulators
    t.close()
    return items, acc, length, escaped, paths


def _compare_with_oracle(pt, model, O, orc, oracle_nif, W, H, spp):
    items, got, length, escaped, paths = _render_gpu(pt, model, W, H, spp)
    want, esc, _, _ = orc.render(items, 0, spp, O.Settings(width=W, height=H), oracle_nif, mode=0)
    assert paths == W * H * spp
    # discrete decisions are bit-exact against the pinned path-tracing restatement
    assert escaped == esc
    assert np.array_equal(length, want["pathLength"].astype(np.int64))
    b = _rgb(want)
    ok = np.isfinite(b).all(1)                   # the reference's acosf(1+ulp) NaN pixels (codelets.cpp:333), if any
    assert ok.mean() > 0.99999 and np.isfinite(got).all()
    a, b = got[ok], b[ok]
    assert _rrmse(a, b) < 0.01
    close = np.isclose(a, b, rtol=2e-2, atol=1e-3 * spp / 16).all(1)
    assert close.mean() >= 0.995, close.mean()
    return _rrmse(a, b), close.mean()


def test_c1_320x240_16spp_matches_oracle(lib_built, nif_model, O, orc, oracle_nif):
    """BASELINE config C1 exactly: 320x240 at 16 spp (1.23 M path samples)."""
    r, c = _compare_with_oracle(lib_built, nif_model, O, orc, oracle_nif, 320, 240, 16)
    print(f"C1 rRMSE {r:.2e}, pixels within tolerance {c:.5f}")


def test_c2_1104x1000_matches_oracle(lib_built, nif_model, O, orc, oracle_nif):
    """BASELINE config C2's image size, 2 spp (2.2 M path samples): every pixel of the headline image."""
    r, c = _compare_with_oracle(lib_built, nif_model, O, orc, oracle_nif, 1104, 1000, 2)
    print(f"C2 rRMSE {r:.2e}, pixels within tolerance {c:.5f}")


def c4_directions(n=1 << 20, seed=1234):
    """BASELINE config C4 / SURVEY 8(d): unit directions uniform on the sphere from a Philox stream seeded 1234."""
    g = np.random.Generator(np.random.Philox(seed))
    d = g.standard_normal((n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True).astype(np.float32)
    return d.astype(np.float32)


def test_c4_one_million_directions_match_oracle(lib_built, nif_model, O, orc, oracle_nif):
    d = c4_directions()
    rec = np.zeros((len(d), 1), O.CONTRIB)
    rec["clr"][:, 0, :] = d; rec["weight"] = 1.0; rec["type"] = O.C_ESCAPED
    u, v = orc.escaped_uv(rec, 0.0)              # codelets.cpp:333-347 (pinned)
    keep = np.isfinite(u)
    assert keep.mean() > 0.99999
    u, v = u[keep], v[keep]
    t = lib_built.PathTracer(64, 48)
    t.init_nif_weights(nif_model)
    got = t.nif_infer(u, v, lib_built.NIF_TCGEN05)
    t.close()
    want = orc.nif_forward(oracle_nif, u, v, mode=0)
    assert np.isfinite(got).all()
    err = np.abs(np.log(got) - np.log(want))
    print(f"C4: {len(u)} directions, max |dlog| {err.max():.2e}, mean {err.mean():.2e}, equal bits {np.mean(got == want):.4f}")
    assert err.max() < 1e-2                     # north_star: NIF outputs within 1e-2
    assert err.mean() < 1e-3


def test_converged_image_with_independent_streams(lib_built, nif_model, O, orc, oracle_nif):
    """SURVEY 8(d) last row: the converged-image gate with INDEPENDENT random streams. A 64x48 image; the GPU
    renders it to convergence twice (seeds 11 and 12, 131072 spp each: the GPU's own noise is measured, not
    assumed); the oracle renders 16 independent batches of 32 spp on seed 13 (different Philox key, so no random
    number is shared with either GPU image). Per-pixel Monte-Carlo noise of a 512-spp oracle image is several per
    cent (the synthetic NIF is rough at pixel scale), so it is bounded separately, as the survey allows: the
    oracle's per-pixel standard error comes from its batches, and the gates are
      * whole-image mean radiance: GPU vs oracle within 0.5 % (noise there is ~0.1 %);
      * 16x16-pixel block means (12 blocks x 3 channels): relative RMSE < 1 %;
      * per-pixel relative RMSE after subtracting the measured oracle and GPU noise power: < 1 %;
      * per-pixel z-scores (difference / standard error): mean ~ 0, spread ~ 1 (t-distributed, 15 dof)."""
    W, H = 64, 48
    hi = 131072
    _, a1, _, _, _ = _render_gpu(lib_built, nif_model, W, H, hi, seed=11, sps=32768)
    _, a2, _, _, _ = _render_gpu(lib_built, nif_model, W, H, hi, seed=12, sps=32768)
    a1 /= hi; a2 /= hi
    gpu_noise = _rrmse(a1, a2) / np.sqrt(2.0)    # per image
    ref = 0.5 * (a1 + a2)                        # noise gpu_noise / sqrt(2)
    assert gpu_noise < 0.006, gpu_noise
    items = lib_built.create_work_list_for_image(W, H)
    K, spp = 16, 32
    batches = []
    for k in range(K):
        out, _, _, _ = orc.render(items, k * spp, spp, O.Settings(width=W, height=H, seed=13), oracle_nif, mode=0)
        batches.append(_rgb(out) / spp)
    batches = np.stack(batches)
    o = batches.mean(0)
    se2 = batches.var(0, ddof=1) / K             # squared standard error of the oracle mean, per pixel and channel
    m = float(np.mean(ref))
    # whole image
    assert abs(o.mean() / ref.mean() - 1.0) < 5e-3, o.mean() / ref.mean()
    # block means
    def blocks(x):
        return x.reshape(H // 16, 16, W // 16, 16, 3).mean((1, 3))
    blk = _rrmse(blocks(o.reshape(H, W, 3)), blocks(ref.reshape(H, W, 3)))
    # noise-corrected per-pixel rRMSE
    raw = float(np.mean((o - ref) ** 2))
    corrected = raw - float(np.mean(se2)) - 0.5 * (gpu_noise * m) ** 2
    bias = np.sqrt(max(0.0, corrected)) / m
    z = (o - ref) / np.sqrt(se2 + 0.5 * (gpu_noise * m) ** 2 + 1e-30)
    print(f"independent streams: gpu self-noise {gpu_noise:.4f}, oracle noise {np.sqrt(np.mean(se2)) / m:.4f}, raw rRMSE {np.sqrt(raw) / m:.4f}, "
          f"block rRMSE {blk:.4f}, noise-corrected rRMSE {bias:.4f}, z mean {z.mean():.3f} std {z.std():.3f}")
    assert blk < 0.01, blk
    assert bias < 0.01, bias
    assert abs(z.mean()) < 0.06 and 0.85 < z.std() < 1.3, (z.mean(), z.std())

"""NIF MLP parity (S5-S7): the SIMT kernel and the tcgen05 kernel against the CPU oracle.
Tolerance per north_star: 1e-2. It is applied to the decoded LOG-radiance (x*max+mean, before
exp), i.e. 1e-2 relative in linear radiance. Weights are seeded synthetic (the shipped blob is
missing from the snapshot): parity against the shipped Keras weights stays unpinned."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pt(lib_built, nif_model):
    t = lib_built.PathTracer(64, 48)
    t.init_nif_weights(nif_model)
    yield t
    t.close()


def _uv(n, seed):
    rng = np.random.default_rng(seed)
    return rng.random(n).astype(np.float32), rng.random(n).astype(np.float32)


def test_simt_matches_ref_faithful_oracle(pt, lib_built, orc, oracle_nif):
    u, v = _uv(3000, 21)
    got = pt.nif_infer(u, v, lib_built.NIF_SIMT)
    want = orc.nif_forward(oracle_nif, u, v, mode=0)
    assert np.max(np.abs(np.log(got) - np.log(want))) < 2e-3


@pytest.mark.parametrize("n", [1, 127, 128, 129, 1000, 148 * 128 + 5, 100003])
def test_tcgen05_matches_oracle(pt, lib_built, orc, oracle_nif, n):
    u, v = _uv(n, 22 + n)
    got = pt.nif_infer(u, v, lib_built.NIF_TCGEN05)
    assert np.all(np.isfinite(got))
    m = min(n, 4096)
    sel = np.random.default_rng(1).choice(n, m, replace=False)
    want = orc.nif_forward(oracle_nif, u[sel], v[sel], mode=0)
    err = np.abs(np.log(got[sel]) - np.log(want))
    assert err.max() < 1e-2, err.max()
    want64 = orc.nif_forward(oracle_nif, u[sel], v[sel], mode=1)
    assert np.abs(np.log(got[sel]) - np.log(want64)).max() < 1e-2


def test_tcgen05_equals_simt_everywhere(pt, lib_built):
    """Size-independent property at microbenchmark scale (C4: 1M directions): both kernels agree."""
    u, v = _uv(1 << 20, 5)
    a = pt.nif_infer(u, v, lib_built.NIF_TCGEN05)
    b = pt.nif_infer(u, v, lib_built.NIF_SIMT)
    assert np.abs(np.log(a) - np.log(b)).max() < 1e-2


def test_tcgen05_is_deterministic(pt, lib_built):
    u, v = _uv(50000, 6)
    a = pt.nif_infer(u, v, lib_built.NIF_TCGEN05)
    b = pt.nif_infer(u, v, lib_built.NIF_TCGEN05)
    assert np.array_equal(a, b)


def test_auto_uses_tensor_cores_for_shipped_topology(pt, lib_built):
    u, v = _uv(1000, 7)
    pt.nif_infer(u, v, lib_built.NIF_AUTO)
    # nif_impl_used is reported by path_trace; run one to read it
    pt.init_render_settings(samples_per_step=1)
    pt.setup(lib_built.create_work_list_for_image(64, 48))
    pt.path_trace()
    _, ctr = pt.read_results()
    assert ctr.nif_impl_used == lib_built.NIF_TCGEN05


def test_other_topologies(lib_built, O, orc):
    """Layer count, width, embedding dimension and the skip position come from the weights file
    (NifModel.cpp:299-326), not from constants."""
    for kw in [dict(hidden=64, hidden_layers=3, concat_at=2, embedding_dimension=4),
               dict(hidden=128, hidden_layers=4, concat_at=0, embedding_dimension=8),
               dict(hidden=320, hidden_layers=6, concat_at=4, embedding_dimension=12),
               dict(hidden=256, hidden_layers=2, concat_at=1, embedding_dimension=10)]:
        model = lib_built.synthetic_model(seed=9, **kw)
        nif = O.NifLayers(model.embedding_dimension, model.max, model.mean, model.log_tone_map, model.layers)
        t = lib_built.PathTracer(16, 16)
        t.init_nif_weights(model)
        u, v = _uv(700, 8)
        want = orc.nif_forward(nif, u, v, mode=0)
        for impl in (lib_built.NIF_SIMT, lib_built.NIF_AUTO):
            got = t.nif_infer(u, v, impl)
            assert np.abs(np.log(got) - np.log(want)).max() < 1e-2, (kw, impl)
        t.close()


def test_linear_output_no_tone_map(lib_built, O, orc):
    model = lib_built.synthetic_model(seed=10)
    model.log_tone_map = False
    nif = O.NifLayers(model.embedding_dimension, model.max, model.mean, False, model.layers)
    t = lib_built.PathTracer(16, 16)
    t.init_nif_weights(model)
    u, v = _uv(500, 9)
    got = t.nif_infer(u, v)
    np.testing.assert_allclose(got, orc.nif_forward(nif, u, v, mode=0), atol=1e-2)
    t.close()


@pytest.mark.parametrize("mode", ["ss1", "ss2", "ts5", "ts10", "tp5", "tp10", "mc5", "mc10"])
def test_every_kernel_variant_matches_oracle(lib_built, nif_model, orc, oracle_nif, mode, monkeypatch):
    """B200PT_NIF_MODE selects the tcgen05 kernel variant at init_nif_weights time (activations in shared
    or tensor memory, single CTA / CTA pair / weight multicast between two CTAs, 5 or 10 K-steps per chunk).
    Every variant must give the default variant's result bit for bit and match the oracle."""
    n = 2 * 148 * 128 + 77          # more than one tile per CTA, odd tile count, ragged last tile
    u, v = _uv(n, 31)
    monkeypatch.delenv("B200PT_NIF_MODE", raising=False)
    t0 = lib_built.PathTracer(16, 16)
    t0.init_nif_weights(nif_model)
    base = t0.nif_infer(u, v, lib_built.NIF_TCGEN05)
    t0.close()
    monkeypatch.setenv("B200PT_NIF_MODE", mode)
    t = lib_built.PathTracer(16, 16)
    t.init_nif_weights(nif_model)
    got = t.nif_infer(u, v, lib_built.NIF_TCGEN05)
    t.close()
    assert np.array_equal(got, base)
    sel = np.random.default_rng(2).choice(n, 2048, replace=False)
    want = orc.nif_forward(oracle_nif, u[sel], v[sel], mode=0)
    assert np.abs(np.log(got[sel]) - np.log(want)).max() < 1e-2

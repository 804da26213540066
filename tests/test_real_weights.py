"""NIF parity on REAL Keras weights (north_star gate two: "NIF outputs within 1e-2 of the shipped Keras fp16 weights").

The reference's weights blob (nif_models/urban_alley_01_4k_fp16_yuv/assets.extra/converted.hdf5) is missing from the snapshot
(/root/reference/.MISSING_LARGE_BLOBS:1), so the `shipped` cases SKIP today - and run unchanged the moment the file appears in
one of tests/h5_via_host.py:CANDIDATE_ASSETS (or $B200PT_NIF_ASSETS points at an assets.extra directory). The `h5_writer`
cases run the very same code on a file written by tests/h5_writer.py so that the path itself (HDF5 subset reader ->
model_config walk -> oracle / GPU) is exercised in every run. Until a real file is read: HDF5 container parity and NIF
arithmetic parity stay UNPINNED."""
import numpy as np
import pytest

import h5_via_host as HV


def _assets(kind, tmp_path):
    if kind == "shipped":
        d = HV.shipped_assets()
        if d is None:
            pytest.skip("converted.hdf5 of the shipped model is not present (reference snapshot: .MISSING_LARGE_BLOBS:1)")
        return d
    import json
    import shutil
    from h5_writer import write_keras_nif
    from ipu_path_trace_b200.nif import synthetic_model
    from test_host_cpu import SHIPPED_METADATA
    d = tmp_path / "assets.extra"
    d.mkdir()
    shutil.copy(SHIPPED_METADATA, d / "nif_metadata.txt")
    write_keras_nif(str(d / "converted.hdf5"), synthetic_model(seed=5))
    return str(d)


@pytest.mark.parametrize("kind", ["h5_writer", "shipped"])
def test_real_weights_load_and_oracle_is_finite(kind, tmp_path, O, orc):
    model, meta, topo = HV.load_keras_nif(_assets(kind, tmp_path))
    print(f"{kind}: topology {topo}; embedding {meta.embedding_dimension}, hidden {meta.hidden_size}, max {meta.max}, mean {meta.mean}")
    assert model.layers[0]["kernel"].shape[0] == 4 * meta.embedding_dimension
    assert model.layers[-1]["kernel"].shape[1] == 3
    assert all(l["kernel"].shape[1] == meta.hidden_size for l in model.layers[:-1])
    nif = O.NifLayers(model.embedding_dimension, model.max, model.mean, model.log_tone_map, model.layers)
    rng = np.random.default_rng(4)
    u, v = rng.random(2000).astype(np.float32), rng.random(2000).astype(np.float32)
    a = orc.nif_forward(nif, u, v, 0)
    b = orc.nif_forward(nif, u, v, 1)
    assert np.isfinite(a).all() and (a > 0).all()
    assert np.abs(np.log(a) - np.log(b)).max() < 2e-2            # binary16 activations vs exact arithmetic, same weights


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["h5_writer", "shipped"])
def test_real_weights_gpu_matches_oracle(kind, tmp_path, lib_built, O, orc):
    model, meta, topo = HV.load_keras_nif(_assets(kind, tmp_path))
    nif = O.NifLayers(model.embedding_dimension, model.max, model.mean, model.log_tone_map, model.layers)
    t = lib_built.PathTracer(64, 48)
    t.init_nif_weights(model)
    rng = np.random.default_rng(5)
    n = 200000
    u, v = rng.random(n).astype(np.float32), rng.random(n).astype(np.float32)
    got = t.nif_infer(u, v, lib_built.NIF_AUTO)
    t.close()
    want = orc.nif_forward(nif, u, v, 0)
    err = np.abs(np.log(got) - np.log(want))
    print(f"{kind}: {topo}: max |dlog| {err.max():.2e} mean {err.mean():.2e}")
    assert err.max() < 1e-2

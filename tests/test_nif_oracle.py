"""The NIF CPU restatement (oracle/nif_oracle.cpp) against an independent numpy statement of
NifModel.cpp:185-245,299-326. PARITY UNPINNED by the reference: it ships neither NIF test
vectors nor (in this snapshot) the weights; see DESIGN.md."""
import numpy as np


def np_features(u, v, E):
    pu = ((u.astype(np.float32) - np.float32(1)) * np.float32(2)).astype(np.float32)
    pv = ((v.astype(np.float32) - np.float32(1)) * np.float32(2)).astype(np.float32)
    c = (2.0 ** np.arange(E)).astype(np.float32)
    xu = (pu[:, None] * c).astype(np.float32).astype(np.float16).astype(np.float64)
    xv = (pv[:, None] * c).astype(np.float32).astype(np.float16).astype(np.float64)
    f = np.concatenate([np.sin(xu), np.sin(xv), np.cos(xu), np.cos(xv)], 1)
    return f.astype(np.float32).astype(np.float16)


def np_forward_fp64(model, u, v):
    feat = np_features(u, v, model.embedding_dimension).astype(np.float64)
    x = feat
    for l in model.layers:
        k = l["kernel"].astype(np.float64)
        if x.shape[1] != k.shape[0]:
            x = np.concatenate([x, feat], 1)
        x = x @ k
        if l.get("bias") is not None:
            x = x + l["bias"].astype(np.float64)
        if l["relu"]:
            x = np.maximum(x, 0)
    o = x[:, :3] * np.float64(np.float32(model.max)) + np.array(model.mean, np.float32).astype(np.float64)
    return np.exp(o) if model.log_tone_map else o


def test_features_bit_exact(orc):
    rng = np.random.default_rng(3)
    u = rng.random(4096).astype(np.float32); v = rng.random(4096).astype(np.float32)
    u[:4] = [0.0, 1.0, 0.5, 0.25]; v[:4] = [1.0, 0.0, 0.5, 0.75]
    got = orc.nif_encode(u, v, 12)
    assert np.array_equal(got.view(np.uint16), np_features(u, v, 12).view(np.uint16))


def test_feature_order_and_values(orc):
    # NifModel.cpp:216 / :427-430: [sin(u c), sin(v c), cos(u c), cos(v c)]
    f = orc.nif_encode(np.array([0.75], np.float32), np.array([0.5], np.float32), 2).astype(np.float32)[0]
    pu, pv = (0.75 - 1) * 2, (0.5 - 1) * 2
    want = [np.sin(pu), np.sin(2 * pu), np.sin(pv), np.sin(2 * pv), np.cos(pu), np.cos(2 * pu), np.cos(pv), np.cos(2 * pv)]
    np.testing.assert_allclose(f, want, atol=1e-3)


def test_forward_fp64_mode_matches_numpy(orc, nif_model, oracle_nif):
    rng = np.random.default_rng(4)
    u = rng.random(512).astype(np.float32); v = rng.random(512).astype(np.float32)
    got = orc.nif_forward(oracle_nif, u, v, mode=1)
    np.testing.assert_allclose(got, np_forward_fp64(nif_model, u, v), rtol=2e-6)


def test_ref_faithful_mode_within_budget_of_fp64(orc, oracle_nif):
    """binary16 rounding of every Dense output stays well inside the 1e-2 budget north_star gives."""
    rng = np.random.default_rng(5)
    u = rng.random(2048).astype(np.float32); v = rng.random(2048).astype(np.float32)
    a = orc.nif_forward(oracle_nif, u, v, mode=0)
    b = orc.nif_forward(oracle_nif, u, v, mode=1)
    assert np.max(np.abs(np.log(a) - np.log(b))) < 1e-2


def test_skip_concat_autodetect(O, orc):
    """NifModel.cpp:303-308: a layer whose kernel has more rows than the activations gets
    concat(activations, features)."""
    rng = np.random.default_rng(6)
    E = 2
    layers = [dict(kernel=rng.normal(0, .3, (8, 16)).astype(np.float16), bias=None, relu=True),
              dict(kernel=rng.normal(0, .3, (24, 16)).astype(np.float16), bias=rng.normal(0, .1, 16).astype(np.float16), relu=True),
              dict(kernel=rng.normal(0, .3, (16, 3)).astype(np.float16), bias=None, relu=False)]
    nif = O.NifLayers(E, 1.0, (0.0, 0.0, 0.0), False, layers)
    u = rng.random(64).astype(np.float32); v = rng.random(64).astype(np.float32)

    class M:
        embedding_dimension = E; max = 1.0; mean = (0.0, 0.0, 0.0); log_tone_map = False
    M.layers = layers
    np.testing.assert_allclose(orc.nif_forward(nif, u, v, mode=1), np_forward_fp64(M, u, v), rtol=1e-5, atol=1e-6)


def test_flops_formula(orc, nif_model, oracle_nif):
    # NifModel.cpp:122-134; nominal topology -> 1 089 283
    assert orc.nif_flops_per_eval(oracle_nif) == nif_model.flops_per_eval() == 1089283


def test_fast_cpu_nif_matches_the_scalar_restatement(orc, oracle_nif):
    """oracle/nif_fast.cpp (mode 2: the batched AVX GEMM that bench.py's CPU legs TIME) states the same network
    as the scalar restatement (mode 0, the checker): decoded log-radiance within 1e-3 (it is in fact equal on almost
    every ray: both round every Dense output to binary16's significand; only fp32 summation order / FMA differ),
    including a ragged last batch and a batch of one."""
    rng = np.random.default_rng(11)
    for n in (1, 7, 384, 5000):
        u = rng.random(n).astype(np.float32); v = rng.random(n).astype(np.float32)
        a = orc.nif_forward(oracle_nif, u, v, 0)
        b = orc.nif_forward(oracle_nif, u, v, 2)
        assert np.all(np.isfinite(b))
        assert np.abs(np.log(a) - np.log(b)).max() < 1e-3
    assert (a == b).mean() > 0.99


def test_cpu_arm_uses_all_host_threads_under_torchrun_env(orc):
    """torch.distributed.run exports OMP_NUM_THREADS=1; the CPU arm lifts it."""
    import os
    n = orc.use_all_host_threads()
    assert n == len(os.sched_getaffinity(0))

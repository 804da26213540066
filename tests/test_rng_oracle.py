"""The counter-based RNG shared by the oracle and the CUDA path (the reference's IPU hardware
RNG cannot be reproduced, north_star): Philox4x32-10 against the Random123 known answers."""
import numpy as np


def test_philox_random123_kat(orc):
    kat = [
        ([0] * 4, [0] * 2, [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]),
        ([0xffffffff] * 4, [0xffffffff] * 2, [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]),
        ([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0],
         [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]),
    ]
    for c, k, want in kat:
        assert list(orc.philox(c, k)) == want


def test_aa_noise_moments(orc):
    rng = np.random.default_rng(1)
    w = rng.integers(0, 2**32, 200000, dtype=np.uint64).astype(np.uint32)
    n = orc.half_to_float(orc.aa_noise(w, 0))
    assert abs(n.mean()) < 0.01 and abs(n.std() - 1.0) < 0.01        # poprand::normal(0,1), PathTracerApp.cpp:34-37
    u = orc.half_to_float(orc.aa_noise(w, 1))
    assert u.min() >= -1 and u.max() <= 1 and abs(u.std() - 1 / np.sqrt(3)) < 0.01   # uniform(-1,1) :30-33
    # truncated normal :38-41 is a conditional distribution (re-draw), not a clamp: nothing piles up at the bounds
    # (a clamp leaves 0.27 % of the samples at exactly +-3; binary16 rounding of values just below 3 leaves ~1e-5),
    # the variance is that of N(0,1) given |x| <= 3, and pairs that needed no re-draw are the plain normal's
    w = rng.integers(0, 2**32, 2000000, dtype=np.uint64).astype(np.uint32)
    n = orc.half_to_float(orc.aa_noise(w, 0)).reshape(-1, 2)
    t = orc.half_to_float(orc.aa_noise(w, 2)).reshape(-1, 2)
    assert np.abs(t).max() <= 3.0
    assert np.mean(np.abs(t) == 3.0) < 1e-4
    assert abs(t.var() - 0.97334) < 0.004 and abs(t.mean()) < 0.003
    kept = np.abs(n).max(axis=1) < 3.0            # strict: binary16 rounding maps values a hair beyond 3 onto 3.0
    assert np.array_equal(t[kept], n[kept]) and 0.003 < 1 - kept.mean() < 0.008
    redrawn = t[~kept]
    assert abs(redrawn.var() - 0.97334) < 0.1 and abs(redrawn.mean()) < 0.05      # the replacement pairs are themselves truncated normals


def test_half_conversion_matches_numpy(orc):
    rng = np.random.default_rng(2)
    x = np.concatenate([rng.normal(0, 100, 5000), rng.normal(0, 1e-4, 5000), [0, -0.0, 65504, 65520, 1e-8, np.inf]]).astype(np.float32)
    assert np.array_equal(orc.float_to_half(x), x.astype(np.float16).view(np.uint16))
    bits = np.arange(0, 65536, dtype=np.uint16)
    f = orc.half_to_float(bits)
    want = bits.view(np.float16).astype(np.float32)
    ok = ~np.isnan(want)
    assert np.array_equal(f[ok], want[ok])

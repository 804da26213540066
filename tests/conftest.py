"""Shared fixtures. `-m "not gpu"` runs on CPU (oracle vs golden, host logic, ABI surface);
`-m gpu` runs the parity tests proper through the C ABI on a B200."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200)")


@pytest.fixture(scope="session")
def golden():
    return np.load(os.path.join(ROOT, "tests", "golden", "ref_golden.npz"))


@pytest.fixture(scope="session")
def orc():
    from oracle import oracle as O
    O.build(ref=True)
    return O.Oracle()


@pytest.fixture(scope="session")
def O():
    from oracle import oracle as O
    return O


@pytest.fixture(scope="session")
def lib_built():
    import ipu_path_trace_b200 as pt
    if not os.path.exists(pt.LIB_PATH):
        pt.build()
    return pt


@pytest.fixture(scope="session")
def nif_model(lib_built):
    """Seeded synthetic weights in the shipped model's topology (the real blob is missing
    from the reference snapshot)."""
    return lib_built.synthetic_model()


def to_oracle_nif(O, model):
    return O.NifLayers(model.embedding_dimension, model.max, model.mean, model.log_tone_map, model.layers)


@pytest.fixture(scope="session")
def oracle_nif(O, nif_model):
    return to_oracle_nif(O, nif_model)

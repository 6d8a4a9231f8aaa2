import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    """The CPU oracle (test infrastructure; built on demand with gcc)."""
    from oracle import oracle as o
    o.build()
    return o


@pytest.fixture(scope="session")
def nb():
    import nbody_b200
    return nbody_b200


def norm_rel_err(test, ref):
    """SURVEY.md 8(c) parity metric: max_i |x_i^test - x_i^ref|_2 / max_i |x_i^ref|_2
    over the xyz columns."""
    import numpy as np
    d = np.linalg.norm(np.asarray(test, dtype=np.float64)[:, :3] - np.asarray(ref, dtype=np.float64)[:, :3], axis=1)
    r = np.linalg.norm(np.asarray(ref, dtype=np.float64)[:, :3], axis=1)
    return float(d.max() / max(r.max(), 1e-300))

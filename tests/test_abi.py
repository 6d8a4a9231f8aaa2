"""CPU-side checks of the C-ABI boundary: the library builds/loads, exports every
symbol include/nbody.h declares, and fails loudly (no fallback) without a GPU."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "nbody.h")).read()
    return sorted(set(re.findall(r"NBODY_API\s+[\w\s\*]+?\b(nbody_\w+)\s*\(", text)))


def test_header_and_binding_agree(nb):
    from nbody_b200 import _lib
    assert header_symbols() == sorted(_lib.SYMBOLS)


def test_library_exports_every_declared_symbol(nb):
    from nbody_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        _lib.build()
    L = ctypes.CDLL(_lib.LIB_PATH)
    for sym in header_symbols():
        assert hasattr(L, sym), sym
    L.nbody_version.restype = ctypes.c_char_p
    assert b"sm_100a" in L.nbody_version()


def test_strerror_and_invalid_arguments(nb):
    from nbody_b200 import _lib
    L = _lib.lib()
    assert L.nbody_strerror(0) == b"ok"
    assert L.nbody_strerror(_lib.NBODY_EINVAL) == b"invalid argument"
    ctx = ctypes.c_void_p()
    assert L.nbody_create(ctypes.byref(ctx), 7, 1, None, 16) == _lib.NBODY_EINVAL      # bad precision
    assert L.nbody_create(ctypes.byref(ctx), 0, 0, None, 16) == _lib.NBODY_EINVAL      # no gpus
    assert L.nbody_create(ctypes.byref(ctx), 0, 1, None, 0) == _lib.NBODY_EINVAL       # zero capacity
    assert L.nbody_create_rank(ctypes.byref(ctx), 0, 2, 2, 0, None, 16) == _lib.NBODY_EINVAL  # rank >= world
    assert L.nbody_step(None, 0.1, 1.0, 0.1, 1) == _lib.NBODY_EINVAL
    assert L.nbody_upload(None, 1, None, None, None, None) == _lib.NBODY_EINVAL
    L.nbody_destroy(None)  # no-op


def test_no_cpu_fallback_without_gpu(nb):
    """Without a CUDA device the product path must raise, never compute on the CPU."""
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        pytest.skip("a GPU is present")
    with pytest.raises(nb.NBodyError) as e:
        nb.Engine(64)
    assert e.value.status == -2  # NBODY_ECUDA
    ps = nb.PhysicsSystem()          # the device context is created on first use ...
    ps.addBody(nb.physics.Body(1.0, nb.physics.V3(0, 0, 0), nb.physics.V3(0, 0, 0)))
    with pytest.raises(nb.NBodyError):
        ps.integrate(0.008)          # ... and every step needs it: no host path computes gravity
    with pytest.raises(nb.NBodyError):
        nb.PhysicsSystem(collisions=True).engine


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "n-body-sim_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cc", ".js")):
                src = open(os.path.join(dirpath, f), errors="ignore").read()
                assert not re.search(r"import\s+oracle|from\s+oracle|from\s+\.+oracle|liboracle|oracle/|oracle\.py", src), (dirpath, f)

"""ctypes front-end of the CPU oracle (oracle/nbody_oracle.c).

TEST INFRASTRUCTURE ONLY.  Importers allowed: tests/, __graft_entry__.smoke(),
and bench.py's cpu_baseline / --impl reference legs.  The product package
(n-body-sim_b200/) never imports this module.  PARITY UNPINNED: see the header
of nbody_oracle.c.

All arrays are float64, C-contiguous, in the host layout of include/nbody.h:
pos_mass[n,4] = x,y,z,m; vel[n,4]; acc[n,4]; temp[n].
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")

_dp = ctypes.POINTER(ctypes.c_double)
_ip = ctypes.POINTER(ctypes.c_int64)


def build(force: bool = False) -> str:
    """Compile liboracle.so with the committed Makefile (gcc only, seconds)."""
    src = os.path.join(_HERE, "nbody_oracle.c")
    stale = (not os.path.exists(_LIB_PATH)) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src)
    if force or stale:
        subprocess.run(["make", "-C", _HERE, "-s"], check=True)
    return _LIB_PATH


_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(_LIB_PATH)
        sz, d, i = ctypes.c_size_t, ctypes.c_double, ctypes.c_int
        L.oracle_compute_gravity.argtypes = [sz, _dp, _dp, d, d]
        L.oracle_compute_gravity_mt.argtypes = [sz, _dp, _dp, d, d]
        L.oracle_kick1_drift.argtypes = [sz, _dp, _dp, _dp, d]
        L.oracle_kick2.argtypes = [sz, _dp, _dp, _dp, d, d]
        L.oracle_integrate.argtypes = [sz, _dp, _dp, _dp, _dp, d, d, d, d, i]
        L.oracle_acc_subset.argtypes = [sz, _dp, sz, _ip, _dp, d, d]
        L.oracle_energy.argtypes = [sz, _dp, _dp, d, d, _dp, _dp, _dp]
        L.oracle_max_threads.restype = i
        for f in ("oracle_compute_gravity", "oracle_compute_gravity_mt", "oracle_kick1_drift",
                  "oracle_kick2", "oracle_integrate", "oracle_acc_subset", "oracle_energy"):
            getattr(L, f).restype = None
        _lib = L
    return _lib


def _f64(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float64)
    if shape is not None:
        assert a.shape == shape, (a.shape, shape)
    return a


def _p(a):
    return a.ctypes.data_as(_dp) if a is not None else None


def compute_gravity(pos_mass, G, softening, mt=False):
    """acc[n,4] of computeGravity (idx:264-291).  mt=True: courtesy OpenMP variant."""
    pm = _f64(pos_mass)
    n = pm.shape[0]
    acc = np.zeros((n, 4))
    fn = lib().oracle_compute_gravity_mt if mt else lib().oracle_compute_gravity
    fn(n, _p(pm), _p(acc), float(G), float(softening))
    return acc


def integrate(pos_mass, vel, acc, temp, dt, G, softening, heat_decay=0.998, nsteps=1):
    """nsteps x integrate(dt) (idx:247-262 minus idx:256).  Returns new copies."""
    pm, v, a = _f64(pos_mass).copy(), _f64(vel).copy(), _f64(acc).copy()
    n = pm.shape[0]
    t = None if temp is None else _f64(temp, (n,)).copy()
    lib().oracle_integrate(n, _p(pm), _p(v), _p(a), _p(t), float(dt), float(G), float(softening),
                           float(heat_decay), int(nsteps))
    return pm, v, a, t


def kick1_drift(pos_mass, vel, acc, dt):
    pm, v, a = _f64(pos_mass).copy(), _f64(vel).copy(), _f64(acc)
    lib().oracle_kick1_drift(pm.shape[0], _p(pm), _p(v), _p(a), float(dt))
    return pm, v


def kick2(vel, acc, temp, dt, heat_decay=0.998):
    v, a = _f64(vel).copy(), _f64(acc)
    t = None if temp is None else _f64(temp).copy()
    lib().oracle_kick2(v.shape[0], _p(v), _p(a), _p(t), float(dt), float(heat_decay))
    return v, t


def acc_subset(pos_mass, idx, G, softening):
    """acc[k,4] of the bodies in idx against all n bodies, reference order."""
    pm = _f64(pos_mass)
    ix = np.ascontiguousarray(idx, dtype=np.int64)
    out = np.zeros((ix.shape[0], 4))
    lib().oracle_acc_subset(pm.shape[0], _p(pm), ix.shape[0], ix.ctypes.data_as(_ip), _p(out),
                            float(G), float(softening))
    return out


def energy(pos_mass, vel, G, softening):
    """(kinetic, potential, momentum[3]); diagnostics the reference lacks."""
    pm, v = _f64(pos_mass), _f64(vel)
    ke, pe = ctypes.c_double(), ctypes.c_double()
    mom = np.zeros(3)
    lib().oracle_energy(pm.shape[0], _p(pm), _p(v), float(G), float(softening),
                        ctypes.byref(ke), ctypes.byref(pe), _p(mom))
    return ke.value, pe.value, mom


def max_threads() -> int:
    return int(lib().oracle_max_threads())

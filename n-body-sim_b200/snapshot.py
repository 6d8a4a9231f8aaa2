"""NBODYSN1 snapshot files (include/nbody.h, "snapshots"): host-side reader/writer.

The reference saves nothing; this is the fixture / checkpoint wire format
of SURVEY.md 8(f) row 3.  64-byte little-endian header, then pos_mass[4n], vel[4n],
acc[4n], [temp[n]], [radius[n]] as float32 or float64.  libnbody.so reads and writes the
same bytes (nbody_snapshot_save / nbody_snapshot_load); this module lets tools and tests
make or inspect a file without a GPU.
"""
from __future__ import annotations

import struct
from typing import Optional

import numpy as np

MAGIC = b"NBODYSN1"
HEADER = struct.Struct("<8sQQdddIIQ")
assert HEADER.size == 64


def write(path: str, pos_mass, vel, acc=None, temp=None, radius=None, *, step: int = 0, dt: float = 0.0,
          G: float = 0.0, softening: float = 0.0, dtype=None) -> None:
    dtype = np.dtype(dtype or np.asarray(pos_mass).dtype)
    if dtype not in (np.dtype(np.float32), np.dtype(np.float64)):
        raise ValueError("snapshot arrays are float32 or float64")
    pm = np.ascontiguousarray(pos_mass, dtype=dtype)
    n = pm.shape[0]
    v = np.ascontiguousarray(vel, dtype=dtype)
    a = np.zeros((n, 4), dtype) if acc is None else np.ascontiguousarray(acc, dtype=dtype)
    if pm.shape != (n, 4) or v.shape != (n, 4) or a.shape != (n, 4):
        raise ValueError("pos_mass, vel, acc must be [n,4]")
    flags = (1 if temp is not None else 0) | (2 if radius is not None else 0)
    with open(path, "wb") as f:
        f.write(HEADER.pack(MAGIC, n, int(step), float(dt), float(G), float(softening),
                            0 if dtype == np.float32 else 1, flags, 0))
        for arr in (pm, v, a):
            f.write(arr.astype("<" + dtype.str[1:], copy=False).tobytes())
        for extra in (temp, radius):
            if extra is not None:
                e = np.ascontiguousarray(extra, dtype=dtype)
                if e.shape != (n,):
                    raise ValueError("temp and radius must be [n]")
                f.write(e.tobytes())


def read(path: str) -> dict:
    with open(path, "rb") as f:
        raw = f.read()
    if len(raw) < HEADER.size:
        raise ValueError(f"{path}: too short for an NBODYSN1 header")
    magic, n, step, dt, G, eps, prec, flags, _ = HEADER.unpack_from(raw)
    if magic != MAGIC or prec not in (0, 1) or flags & ~3:
        raise ValueError(f"{path}: not an NBODYSN1 snapshot")
    dtype = np.dtype("<f4" if prec == 0 else "<f8")
    counts = [4 * n, 4 * n, 4 * n] + ([n] if flags & 1 else []) + ([n] if flags & 2 else [])
    if len(raw) != HEADER.size + sum(counts) * dtype.itemsize:
        raise ValueError(f"{path}: size does not match its header")
    out = {"n": n, "step": step, "dt": dt, "G": G, "softening": eps, "dtype": dtype, "temp": None, "radius": None}
    off = HEADER.size
    names = ["pos_mass", "vel", "acc"] + (["temp"] if flags & 1 else []) + (["radius"] if flags & 2 else [])
    for name, cnt in zip(names, counts):
        arr = np.frombuffer(raw, dtype=dtype, count=cnt, offset=off).copy()
        out[name] = arr.reshape(n, 4) if cnt == 4 * n else arr
        off += cnt * dtype.itemsize
    return out

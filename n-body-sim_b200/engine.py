"""Engine: thin, typed wrapper over the C ABI (include/nbody.h).

Host arrays are numpy, float32 or float64 according to the engine precision,
in the step-contract layout: pos_mass[n,4] = x,y,z,m; vel[n,4]; acc[n,4];
temp[n].  All compute happens in libnbody.so on the GPU.
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional, Sequence

import numpy as np

from . import _lib
from ._lib import NBODY_F32, NBODY_F64, NBodyError, check


def _ptr(a: Optional[np.ndarray]):
    return None if a is None else ctypes.c_void_p(a.ctypes.data)


class Engine:
    """One simulation context on one or more B200s.

    precision: "f32" (production) or "f64" (reference-matching mode).
    n_gpus / device_ids: single-process multi-GPU (the N-API host model).
    rank/world/nccl_id: one-process-per-GPU model (see dist.create_rank_engine).
    """

    def __init__(self, capacity: int, precision: str = "f32", n_gpus: int = 1,
                 device_ids: Optional[Sequence[int]] = None, *, rank: Optional[int] = None,
                 world: Optional[int] = None, device: int = 0, nccl_id: Optional[bytes] = None):
        if precision not in ("f32", "f64"):
            raise ValueError("precision must be 'f32' or 'f64'")
        self.precision = precision
        self.dtype = np.float32 if precision == "f32" else np.float64
        self._L = _lib.lib()
        self._ctx = ctypes.c_void_p()
        prec = NBODY_F32 if precision == "f32" else NBODY_F64
        if rank is None:
            ids = None
            if device_ids is not None:
                ids = (ctypes.c_int * n_gpus)(*device_ids)
            st = self._L.nbody_create(ctypes.byref(self._ctx), prec, n_gpus, ids, capacity)
        else:
            buf = None
            if nccl_id is not None:
                if len(nccl_id) != _lib.NCCL_ID_BYTES:
                    raise ValueError("nccl_id must be 128 bytes")
                buf = ctypes.create_string_buffer(bytes(nccl_id), _lib.NCCL_ID_BYTES)
            st = self._L.nbody_create_rank(ctypes.byref(self._ctx), prec, rank, world, device,
                                           ctypes.cast(buf, ctypes.c_void_p) if buf else None, capacity)
        if st != _lib.NBODY_OK:
            self._ctx = ctypes.c_void_p()
            check(st)
        self.n = 0
        self.capacity = int(capacity)

    # -- lifetime ----------------------------------------------------------------
    def close(self) -> None:
        if getattr(self, "_ctx", None) and self._ctx.value:
            self._L.nbody_destroy(self._ctx)
            self._ctx = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # -- state ---------------------------------------------------------------------
    def _arr(self, a, shape, name):
        if a is None:
            return None
        a = np.ascontiguousarray(a, dtype=self.dtype)
        if a.shape != shape:
            raise ValueError(f"{name} must have shape {shape}, got {a.shape}")
        return a

    def upload(self, pos_mass, vel, acc=None, temp=None) -> None:
        pm = np.ascontiguousarray(pos_mass, dtype=self.dtype)
        if pm.ndim != 2 or pm.shape[1] != 4:
            raise ValueError("pos_mass must be [n,4]")
        n = pm.shape[0]
        v = self._arr(vel, (n, 4), "vel")
        a = self._arr(acc, (n, 4), "acc")
        t = self._arr(temp, (n,), "temp")
        check(self._L.nbody_upload(self._ctx, n, _ptr(pm), _ptr(v), _ptr(a), _ptr(t)), self._ctx)
        self.n = n

    def download(self, pos_mass=True, vel=True, acc=True, temp=False, out=None, own_only=False):
        """Returns dict of arrays.  `out` may hold preallocated arrays to fill.  own_only: pos_mass too is read back for
        the own range only (nbody_download_own; in rank mode the ranks together read every body once)."""
        n = self.n
        out = dict(out or {})
        want = {"pos_mass": pos_mass, "vel": vel, "acc": acc, "temp": temp}
        for k, w in want.items():
            if w and k not in out:
                out[k] = np.zeros((n,) if k == "temp" else (n, 4), dtype=self.dtype)
        for k, a in out.items():
            if a.dtype != self.dtype or not a.flags.c_contiguous:
                raise ValueError(f"out[{k}] must be contiguous {self.dtype}")
        fn = self._L.nbody_download_own if own_only else self._L.nbody_download
        check(fn(self._ctx, _ptr(out.get("pos_mass")), _ptr(out.get("vel")), _ptr(out.get("acc")), _ptr(out.get("temp"))), self._ctx)
        return out

    # -- hot path ------------------------------------------------------------------
    def step(self, dt: float, G: float, softening: float, nsteps: int = 1) -> None:
        check(self._L.nbody_step(self._ctx, dt, G, softening, nsteps), self._ctx)

    def step_host(self, pos_mass, vel, acc, temp, dt: float, G: float, softening: float, nsteps: int = 1) -> None:
        """One frame of the JS shim in one call (nbody_step_host): the arrays (contiguous, the engine's dtype;
        `temp` may be None) are uploaded as a new body list, advanced by nsteps x integrate(dt) and overwritten
        IN PLACE with the new state; the state also stays on the device."""
        n = pos_mass.shape[0]
        for name, a, shape in (("pos_mass", pos_mass, (n, 4)), ("vel", vel, (n, 4)), ("acc", acc, (n, 4)), ("temp", temp, (n,))):
            if a is None and name == "temp":
                continue
            if not isinstance(a, np.ndarray) or a.dtype != self.dtype or not a.flags.c_contiguous or a.shape != shape or not a.flags.writeable:
                raise ValueError(f"{name} must be a writable contiguous {np.dtype(self.dtype).name} array of shape {shape}")
        check(self._L.nbody_step_host(self._ctx, n, _ptr(pos_mass), _ptr(vel), _ptr(acc), _ptr(temp), dt, G, softening, nsteps), self._ctx)
        self.n = n

    def kick1_drift(self, dt: float) -> None:
        check(self._L.nbody_kick1_drift(self._ctx, dt), self._ctx)

    def force(self, G: float, softening: float) -> None:
        check(self._L.nbody_force(self._ctx, G, softening), self._ctx)

    def kick2(self, dt: float, heat_decay: float = 0.998) -> None:
        check(self._L.nbody_kick2(self._ctx, dt, heat_decay), self._ctx)

    def set_heat_decay(self, heat_decay: float) -> None:
        check(self._L.nbody_set_heat_decay(self._ctx, heat_decay), self._ctx)

    # -- collision broad phase ------------------------------------
    def set_radius(self, radius) -> None:
        """Body.radius (idx:221) of the uploaded bodies; enables the contact stage."""
        r = self._arr(radius, (self.n,), "radius")
        check(self._L.nbody_set_radius(self._ctx, _ptr(r)), self._ctx)

    def contacts(self, margin: float = 0.0) -> np.ndarray:
        """Pairs (i, j), i < j, with |pos_j - pos_i|^2 < (r_i + r_j + margin)^2, in the
        order resolveCollisions visits them (idx:301-313).  uint32 [k, 2]."""
        cap = max(1024, getattr(self, "_pair_cap", 0))
        while True:
            out = np.empty((cap, 2), dtype=np.uint32)
            cnt = ctypes.c_size_t()
            st = self._L.nbody_contacts(self._ctx, float(margin), ctypes.c_void_p(out.ctypes.data), cap, ctypes.byref(cnt))
            if st == _lib.NBODY_ENOMEM and cnt.value > cap:
                cap = int(cnt.value * 1.25) + 16
                continue
            check(st, self._ctx)
            self._pair_cap = cap
            return out[:cnt.value].copy()

    def contact_stats(self) -> dict:
        """What the last contacts() call did: bodies flagged by the force kernel's
        nearest-neighbour tracking, bodies scanned by the pair kernel, pairs found."""
        st = (ctypes.c_longlong * 4)()
        check(self._L.nbody_contact_stats(self._ctx, st), self._ctx)
        return {"flagged": st[0], "scanned": st[1], "pairs": st[2], "used_flags": bool(st[3]), "candidates_only": st[3] == 2}

    # -- the page's complete step on the device (response included) ---------
    def world_enable(self, seed: int = 0x5EED0001, max_bodies: int = 300) -> None:
        """The context becomes the page's `physics` object: ids, types, radii, contactPairs and a seeded Math.random
        stream live on the device (fp64, one GPU; csrc/page_world.cu)."""
        check(self._L.nbody_world_enable(self._ctx, int(seed), int(max_bodies)), self._ctx)

    def world_config(self, fracture_threshold: float = 25.0, min_fragment_mass: float = 0.2, merge_time: float = 0.5) -> None:
        check(self._L.nbody_world_config(self._ctx, fracture_threshold, min_fragment_mass, merge_time), self._ctx)

    def world_set_bodies(self, id, type, radius) -> None:
        """Body.id / .type / .radius of the uploaded bodies; empties the contact map."""
        i = np.ascontiguousarray(id, dtype=np.float64)
        t = np.ascontiguousarray(type, dtype=np.int32)
        r = np.ascontiguousarray(radius, dtype=np.float64)
        if i.shape != (self.n,) or t.shape != (self.n,) or r.shape != (self.n,):
            raise ValueError("id, type and radius must have one entry per body")
        check(self._L.nbody_world_set_bodies(self._ctx, _ptr(i), _ptr(t), _ptr(r)), self._ctx)

    def world_step(self, dt: float, G: float, softening: float, nsteps: int = 1, collisions: bool = True) -> None:
        """nsteps x the reference's complete integrate(dt) (idx:247-262, resolveCollisions included), on the device."""
        check(self._L.nbody_world_step(self._ctx, dt, G, softening, int(nsteps), int(bool(collisions))), self._ctx)

    def world_sync(self) -> dict:
        """Synchronise and adopt the device's body count; returns id, type, radius, counters, contact count, rng state."""
        n, nc, rng = ctypes.c_size_t(), ctypes.c_size_t(), ctypes.c_ulonglong()
        check(self._L.nbody_world_sync(self._ctx, ctypes.byref(n), None, None, None, None, ctypes.byref(nc), ctypes.byref(rng)), self._ctx)
        self.n = n.value
        ids, ty, rad = np.zeros(self.n), np.zeros(self.n, np.int32), np.zeros(self.n)
        cnt = (ctypes.c_longlong * 5)()
        check(self._L.nbody_world_sync(self._ctx, None, _ptr(ids), _ptr(ty), _ptr(rad), cnt, None, None), self._ctx)
        return {"n": self.n, "id": ids, "type": ty, "radius": rad, "n_contacts": nc.value, "rng": rng.value,
                "counters": dict(zip(("merges", "fractures", "overlaps", "flashes", "explosions"), list(cnt)))}

    def world_frame(self, dt: float, G: float, softening: float, nsteps: int = 2, collisions: bool = True, out: dict | None = None) -> dict:
        """The page's frame (idx:873-883): nsteps x the complete integrate(dt), then everything visuals.update reads,
        through one launch and one synchronise.  `out` (a previous call's result) is refilled in place when it is large enough."""
        cap = self.capacity
        if out is None or len(out["_buf"]["temp"]) < cap:
            buf = {"pos_mass": np.zeros((cap, 4)), "vel": np.zeros((cap, 4)), "acc": np.zeros((cap, 4)), "temp": np.zeros(cap),
                   "id": np.zeros(cap), "type": np.zeros(cap, np.int32), "radius": np.zeros(cap)}
        else:
            buf = out["_buf"]
        n, nc, rng = ctypes.c_size_t(), ctypes.c_size_t(), ctypes.c_ulonglong()
        cnt = (ctypes.c_longlong * 5)()
        check(self._L.nbody_world_frame(self._ctx, dt, G, softening, int(nsteps), int(bool(collisions)), cap, ctypes.byref(n),
                                        _ptr(buf["pos_mass"]), _ptr(buf["vel"]), _ptr(buf["acc"]), _ptr(buf["temp"]), _ptr(buf["id"]),
                                        _ptr(buf["type"]), _ptr(buf["radius"]), cnt, ctypes.byref(nc), ctypes.byref(rng)), self._ctx)
        self.n = n.value
        res = {k: v[:self.n] for k, v in buf.items()}
        res.update(n=self.n, n_contacts=nc.value, rng=rng.value, _buf=buf,
                   counters=dict(zip(("merges", "fractures", "overlaps", "flashes", "explosions"), list(cnt))))
        return res

    def world_contacts(self) -> np.ndarray:
        """contactPairs as rows (lower id, higher id, accumulated time), sorted."""
        cap = 2048
        out = np.zeros((cap, 3))
        cnt = ctypes.c_size_t()
        check(self._L.nbody_world_contacts(self._ctx, _ptr(out), cap, ctypes.byref(cnt)), self._ctx)
        rows = out[:cnt.value]
        return rows[np.lexsort((rows[:, 1], rows[:, 0]))] if len(rows) else rows

    # -- renderer feed: Body.history and ribbon trails -----------------
    def history_enable(self, depth: int = 80) -> None:
        """Keep the last `depth` positions of every body on the device (CONFIG.trailLength, idx:58)."""
        check(self._L.nbody_history_enable(self._ctx, int(depth)), self._ctx)
        self._hist_depth = int(depth)

    def history_push(self) -> None:
        """body.history.unshift(pos.clone()); pop the oldest (idx:564-565), for every body."""
        check(self._L.nbody_history_push(self._ctx), self._ctx)

    def history_remap(self, source) -> None:
        """After the body list changed and was uploaded again: body i continues the ring of body
        source[i] of the previous list; source[i] < 0 is a body the ring has not seen (filled at the
        next push, idx:540)."""
        src = np.ascontiguousarray(source, dtype=np.int32)
        if src.shape != (self.n,):
            raise ValueError(f"source must have one entry per body ({self.n})")
        check(self._L.nbody_history_remap(self._ctx, _ptr(src)), self._ctx)

    def history(self) -> np.ndarray:
        """Body.history of every body: [n, depth, 3], newest first."""
        out = np.empty((self.n, self._hist_depth, 3), dtype=self.dtype)
        check(self._L.nbody_history_download(self._ctx, _ptr(out)), self._ctx)
        return out

    def trail_ribbons(self, camera, download: bool = True) -> Optional[np.ndarray]:
        """The ribbon vertex arrays VisualSystem.update builds (idx:567-592): float32
        [n, depth, 6].  download=False builds them on the device only."""
        cam = (ctypes.c_double * 3)(*[float(x) for x in camera])
        out = np.empty((self.n, self._hist_depth, 6), dtype=np.float32) if download else None
        check(self._L.nbody_trail_ribbons(self._ctx, cam, _ptr(out)), self._ctx)
        return out

    # -- headless body renderer ------------------------------------------
    def set_materials(self, type=None, seed=None) -> None:
        """Body.type per body (0 rock, 1 metal, 2 ice or the names) and each material's `seed` uniform (idx:492)."""
        t = s = None
        if type is not None:
            names = {"rock": 0, "metal": 1, "ice": 2}
            t = np.ascontiguousarray([names.get(x, x) for x in type], dtype=np.int32)
            if t.shape != (self.n,):
                raise ValueError("type must have one entry per body")
        if seed is not None:
            s = np.ascontiguousarray(seed, dtype=np.float32)
            if s.shape != (self.n,):
                raise ValueError("seed must have one entry per body")
        check(self._L.nbody_set_materials(self._ctx, _ptr(t), _ptr(s)), self._ctx)

    def render(self, width: int, height: int, camera=(0.0, 80.0, 150.0), target=(0.0, 0.0, 0.0), fov_deg: float = 45.0,
               ids: bool = False, download: bool = True):
        """The planet shaders (idx:83-202) on ray-cast spheres: float32 [height, width, 4]
        (gl_FragColor before tone mapping); with ids=True also int32 [height, width]."""
        cam = (ctypes.c_double * 3)(*[float(x) for x in camera])
        tgt = (ctypes.c_double * 3)(*[float(x) for x in target])
        ok = width > 0 and height > 0                      # bad sizes: let the library report them
        img = np.empty((height, width, 4), dtype=np.float32) if (download and ok) else None
        idb = np.empty((height, width), dtype=np.int32) if (ids and download and ok) else None
        check(self._L.nbody_render(self._ctx, int(width), int(height), cam, tgt, float(fov_deg), _ptr(img), _ptr(idb)), self._ctx)
        return (img, idb) if ids else img

    # -- snapshots ------------------------------------------------------------------------
    def save_snapshot(self, path: str, step: int = 0, dt: float = 0.0, G: float = 0.0, softening: float = 0.0) -> None:
        check(self._L.nbody_snapshot_save(self._ctx, os.fsencode(path), int(step), dt, G, softening), self._ctx)

    def load_snapshot(self, path: str) -> dict:
        step, dt, G, eps = ctypes.c_ulonglong(), ctypes.c_double(), ctypes.c_double(), ctypes.c_double()
        check(self._L.nbody_snapshot_load(self._ctx, os.fsencode(path), ctypes.byref(step), ctypes.byref(dt),
                                          ctypes.byref(G), ctypes.byref(eps)), self._ctx)
        self.n = self.info()["n"]
        return {"step": step.value, "dt": dt.value, "G": G.value, "softening": eps.value}

    # -- diagnostics / plumbing ------------------------------------------------------
    def energy(self, G: float, softening: float):
        ke, pe = ctypes.c_double(), ctypes.c_double()
        mom = (ctypes.c_double * 3)()
        check(self._L.nbody_energy(self._ctx, G, softening, ctypes.byref(ke), ctypes.byref(pe), mom),
              self._ctx)
        return ke.value, pe.value, np.array(list(mom))

    def sync(self) -> None:
        check(self._L.nbody_sync(self._ctx), self._ctx)

    def set_stream(self, cuda_stream: int) -> None:
        check(self._L.nbody_set_stream(self._ctx, ctypes.c_void_p(cuda_stream or None)), self._ctx)

    def enable_timing(self, on: bool = True) -> None:
        check(self._L.nbody_enable_timing(self._ctx, int(on)), self._ctx)

    def read_timing(self):
        """(force-kernel launches, summed device ms) since the last read."""
        n, ms = ctypes.c_longlong(), ctypes.c_double()
        check(self._L.nbody_read_timing(self._ctx, ctypes.byref(n), ctypes.byref(ms)), self._ctx)
        return n.value, ms.value

    def info(self) -> dict:
        inf = _lib.Info()
        check(self._L.nbody_get_info(self._ctx, ctypes.byref(inf)), self._ctx)
        return {k: getattr(inf, k) for k, _ in _lib.Info._fields_}


STATIC_MAX_BODIES = 16384   # include/nbody.h NBODY_STATIC_MAX_BODIES


def accel_f32_static(pos_mass_dev: int, acc_dev: int, n: int, G: float, softening: float, stream: int = 0) -> None:
    """nbody_accel_f32_static (include/nbody.h): fp32 accelerations of n <= STATIC_MAX_BODIES bodies between two DEVICE buffers
    (addresses on the current device), queued on `stream`, with no allocation anywhere: the scratch is a static device array."""
    err = ctypes.create_string_buffer(256)
    st = _lib.lib().nbody_accel_f32_static(ctypes.c_void_p(pos_mass_dev), ctypes.c_void_p(acc_dev), int(n), float(G), float(softening),
                                           ctypes.c_void_p(stream), err, len(err))
    if st != _lib.NBODY_OK:
        raise NBodyError(st, err.value.decode() or "nbody_accel_f32_static")


def device_count() -> int:
    n = _lib.lib().nbody_device_count()
    if n < 0:
        raise NBodyError(n, "cannot query CUDA devices")
    return n


def nccl_unique_id() -> bytes:
    buf = ctypes.create_string_buffer(_lib.NCCL_ID_BYTES)
    check(_lib.lib().nbody_nccl_unique_id(ctypes.cast(buf, ctypes.c_void_p)))
    return buf.raw

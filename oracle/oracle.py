"""ctypes front-end of the CPU oracle (oracle/nbody_oracle.c).

TEST INFRASTRUCTURE ONLY.  Importers allowed: tests/, __graft_entry__.smoke(),
and bench.py's cpu_baseline / --impl reference legs.  The product package
(n-body-sim_b200/) never imports this module.  Parity: pinned to the reference's source text
(oracle/js_reference.py), not to V8 -- see the header of nbody_oracle.c.

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
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith(".c") or f == "Makefile"]
    stale = (not os.path.exists(_LIB_PATH)) or os.path.getmtime(_LIB_PATH) < max(os.path.getmtime(f) for f in srcs)
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
        # collision stage (collide_oracle.c)
        wp, u64 = ctypes.c_void_p, ctypes.c_uint64
        _i32p, _u32p = ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_uint32)
        L.oracle_world_create.argtypes = [u64, sz]
        L.oracle_world_create.restype = wp
        L.oracle_world_destroy.argtypes = [wp]
        L.oracle_world_destroy.restype = None
        L.oracle_world_seed.argtypes = [wp, u64]
        L.oracle_world_seed.restype = None
        L.oracle_world_config.argtypes = [wp, d, d, d, d]
        L.oracle_world_config.restype = None
        L.oracle_world_add.argtypes = [wp, d, _dp, _dp, i, d]
        L.oracle_world_add.restype = None
        L.oracle_world_clear.argtypes = [wp]
        L.oracle_world_clear.restype = None
        L.oracle_world_size.argtypes = [wp]
        L.oracle_world_size.restype = sz
        L.oracle_world_contacts.argtypes = [wp]
        L.oracle_world_contacts.restype = sz
        L.oracle_world_counters.argtypes = [wp, ctypes.POINTER(ctypes.c_long)]
        L.oracle_world_counters.restype = None
        L.oracle_world_get_contacts.argtypes = [wp, _dp]
        L.oracle_world_get_contacts.restype = None
        L.oracle_world_rng.argtypes = [wp]
        L.oracle_world_rng.restype = ctypes.c_uint64
        L.oracle_world_get.argtypes = [wp, _dp, _dp, _dp, _dp, _dp, _i32p, _dp]
        L.oracle_world_get.restype = None
        L.oracle_world_resolve_collisions.argtypes = [wp, d]
        L.oracle_world_resolve_collisions.restype = None
        L.oracle_world_integrate.argtypes = [wp, d, d, d, i, i]
        L.oracle_world_integrate.restype = None
        L.oracle_trail_ribbons.argtypes = [sz, i, i, _dp, _dp, _dp, ctypes.POINTER(ctypes.c_float)]
        L.oracle_trail_ribbons.restype = None
        _fp = ctypes.POINTER(ctypes.c_float)
        L.oracle_render.argtypes = [sz, _fp, _fp, _fp, ctypes.POINTER(ctypes.c_int), _fp, i, i, _dp, _dp, d, _fp,
                                    ctypes.POINTER(ctypes.c_int32)]
        L.oracle_render.restype = None
        L.oracle_planet_fragment.argtypes = [_fp, _fp, _fp, i, ctypes.c_float, ctypes.c_float, _fp, _fp]
        L.oracle_planet_fragment.restype = None
        L.oracle_snoise.argtypes = [_fp]
        L.oracle_snoise.restype = ctypes.c_float
        for f in ("oracle_overlap_pairs", "oracle_overlap_pairs_mt"):
            getattr(L, f).argtypes = [sz, _dp, _dp, d, _u32p, sz]
            getattr(L, f).restype = sz
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


# ---- collision stage (collide_oracle.c) ---------------------------

TYPES = ("rock", "metal", "ice")   # CONFIG.materials, idx:70-74


def overlap_pairs(pos_mass, radius, margin=0.0, mt=True):
    """(i,j) pairs, i<j, in the reference's visiting order, whose spheres overlap by the
    test of idx:309-313 widened by `margin`.  uint32 [k,2]."""
    pm, r = _f64(pos_mass), _f64(radius)
    n = pm.shape[0]
    fn = lib().oracle_overlap_pairs_mt if mt else lib().oracle_overlap_pairs
    cap = 1 << 16
    while True:
        out = np.zeros((cap, 2), dtype=np.uint32)
        k = fn(n, _p(pm), _p(r), float(margin), out.ctypes.data_as(ctypes.POINTER(ctypes.c_uint32)), cap)
        if k <= cap:
            return out[:k].copy()
        cap = int(k)


class World:
    """The reference's PhysicsSystem with its COMPLETE integrate(dt) (idx:233-443):
    gravity, collisions, merge, fracture.  Math.random is a splitmix64 stream seeded here."""

    def __init__(self, seed=0x5EED0001, max_bodies=300):
        self._L = lib()
        self._w = ctypes.c_void_p(self._L.oracle_world_create(int(seed), int(max_bodies)))

    def __del__(self):
        if getattr(self, "_w", None):
            self._L.oracle_world_destroy(self._w)
            self._w = None

    def seed(self, seed):
        """Re-seed the Math.random stand-in (sequential splitmix64)."""
        self._L.oracle_world_seed(self._w, int(seed))

    def config(self, fractureThreshold=25.0, minFragmentMass=0.2, mergeTime=0.5, heatDecay=0.998):
        self._L.oracle_world_config(self._w, fractureThreshold, minFragmentMass, mergeTime, heatDecay)

    def add(self, mass, pos, vel, type="rock", temp=0.0):
        p, v = _f64(pos, (3,)), _f64(vel, (3,))
        self._L.oracle_world_add(self._w, float(mass), _p(p), _p(v), TYPES.index(type), float(temp))

    def clear(self):
        self._L.oracle_world_clear(self._w)

    def __len__(self):
        return int(self._L.oracle_world_size(self._w))

    @property
    def n_contacts(self):
        return int(self._L.oracle_world_contacts(self._w))

    def contacts(self):
        """contactPairs as rows (lower id, higher id, accumulated time), sorted."""
        rows = np.zeros((self.n_contacts, 3))
        if len(rows):
            self._L.oracle_world_get_contacts(self._w, _p(rows))
            rows = rows[np.lexsort((rows[:, 1], rows[:, 0]))]
        return rows

    @property
    def rng(self):
        return int(self._L.oracle_world_rng(self._w))

    def counters(self):
        c = (ctypes.c_long * 5)()
        self._L.oracle_world_counters(self._w, c)
        return dict(zip(("merge", "fracture", "overlap", "flash", "explosion"), list(c)))

    def state(self):
        n = len(self)
        pm, v, a = np.zeros((n, 4)), np.zeros((n, 4)), np.zeros((n, 4))
        t, r, ident = np.zeros(n), np.zeros(n), np.zeros(n)
        ty = np.zeros(n, dtype=np.int32)
        self._L.oracle_world_get(self._w, _p(pm), _p(v), _p(a), _p(t), _p(r),
                                 ty.ctypes.data_as(ctypes.POINTER(ctypes.c_int)), _p(ident))
        return {"pos_mass": pm, "vel": v, "acc": a, "temp": t, "radius": r, "type": ty, "id": ident}

    def resolve_collisions(self, dt):
        self._L.oracle_world_resolve_collisions(self._w, float(dt))

    def integrate(self, dt, G, softening, collisions=True, nsteps=1):
        self._L.oracle_world_integrate(self._w, float(dt), float(G), float(softening), int(bool(collisions)), int(nsteps))


# ---- renderer feed (visual_oracle.c) --------------------------------

def history_push(hist, pos):
    """body.history.unshift(pos.clone()); pop (idx:564-565).  hist [n,depth,3] newest first, or
    None for bodies seen for the first time: depth copies of pos, then the unshift (idx:540, 564)."""
    pos = _f64(pos)[:, :3]
    out = np.empty_like(hist)
    out[:, 1:] = hist[:, :-1]
    out[:, 0] = pos
    return out


def history_init(pos, depth=80):
    """A body seen for the first time: trailLength copies of its position (idx:540)."""
    return np.repeat(_f64(pos)[:, None, :3], depth, axis=1).copy()


def trail_ribbons(hist, radius, camera, trail_length=None):
    """float32 [n, depth, 6]: the ribbon vertex arrays of idx:567-592."""
    h = _f64(hist)
    n, depth, _ = h.shape
    r, cam = _f64(radius, (n,)), _f64(camera, (3,))
    out = np.zeros((n, depth, 6), dtype=np.float32)
    lib().oracle_trail_ribbons(n, depth, int(trail_length or depth), _p(h), _p(r), _p(cam),
                               out.ctypes.data_as(ctypes.POINTER(ctypes.c_float)))
    return out


def render(pos_mass, radius, temp, type, seed, width, height, camera=(0.0, 80.0, 150.0), target=(0.0, 0.0, 0.0), fov_deg=45.0):
    """The planet shaders (idx:83-202) on ray-cast spheres, fp32 like GLSL.  Inputs are rounded to
    float32 first (WebGL uniforms).  Returns (rgba float32 [h,w,4], ids int32 [h,w])."""
    f32 = lambda a: np.ascontiguousarray(a, dtype=np.float32)
    pm, r, t, sd = f32(pos_mass), f32(radius), f32(temp), f32(seed)
    n = pm.shape[0]
    ty = np.ascontiguousarray(type, dtype=np.int32)
    assert pm.shape == (n, 4) and r.shape == (n,) and t.shape == (n,) and sd.shape == (n,) and ty.shape == (n,)
    cam, tgt = _f64(camera, (3,)), _f64(target, (3,))
    rgba = np.zeros((height, width, 4), dtype=np.float32)
    ids = np.zeros((height, width), dtype=np.int32)
    fp = ctypes.POINTER(ctypes.c_float)
    lib().oracle_render(n, pm.ctypes.data_as(fp), r.ctypes.data_as(fp), t.ctypes.data_as(fp),
                        ty.ctypes.data_as(ctypes.POINTER(ctypes.c_int)), sd.ctypes.data_as(fp), int(width), int(height),
                        _p(cam), _p(tgt), float(fov_deg), rgba.ctypes.data_as(fp), ids.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)))
    return rgba, ids


def planet_fragment(vpos, vnormal, vview, type, temperature, seed, sun=(50.0, 50.0, 50.0)):
    """gl_FragColor of planetFragmentShader (idx:164-201) for one fragment, float32[4]."""
    fp = ctypes.POINTER(ctypes.c_float)
    a = [np.ascontiguousarray(x, dtype=np.float32) for x in (vpos, vnormal, vview, sun)]
    rgba = np.zeros(4, dtype=np.float32)
    lib().oracle_planet_fragment(a[0].ctypes.data_as(fp), a[1].ctypes.data_as(fp), a[2].ctypes.data_as(fp), int(type), float(temperature),
                                 float(seed), a[3].ctypes.data_as(fp), rgba.ctypes.data_as(fp))
    return rgba


def snoise(v):
    """snoise(vec3) of the planet shader (idx:114-162), float32."""
    a = np.ascontiguousarray(v, dtype=np.float32)
    return np.float32(lib().oracle_snoise(a.ctypes.data_as(ctypes.POINTER(ctypes.c_float))))

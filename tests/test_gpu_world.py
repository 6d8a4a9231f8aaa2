"""The page's COMPLETE step on the device (csrc/page_world.cu; SURVEY.md 8f row 1 with the response): integrate(dt) with
resolveCollisions, mergeBodies, fractureBodies, the Body constructor and addBody (/root/reference/index.html:208-262, 293-443)
in one kernel.  The collision stage restates the reference's fp64 statements one rounded operation at a time, so with gravity
switched off (G = 0: both sides then integrate identically) every bit must equal the CPU oracle's -- positions, velocities,
temperatures, radii of merged bodies (Math.pow), ids and fragments drawn from the shared Math.random stream, contact times,
counters -- step after step through merges and fractures.  With gravity the usual fp64 gate applies."""
import time

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
TYPES = ("rock", "metal", "ice")
KEYS = ("pos_mass", "vel", "temp", "radius", "type", "id")


def crowd(oracle, n, seed, box, speed, world_seed=0xC0FFEE, max_bodies=300):
    rng = np.random.default_rng(seed)
    w = oracle.World(seed=world_seed, max_bodies=max_bodies)
    for _ in range(n):
        w.add(float(rng.uniform(0.3, 3.0)), rng.uniform(-box, box, 3), rng.normal(0, speed, 3), TYPES[int(rng.integers(0, 3))],
              float(rng.uniform(0, 5)))
    return w


def mirror(nb, w, max_bodies=300, capacity=1024):
    """A device world holding exactly the oracle world's bodies and random stream."""
    st = w.state()
    e = nb.Engine(capacity, precision="f64")
    e.upload(st["pos_mass"], st["vel"], st["acc"], st["temp"])
    e.world_enable(seed=w.rng, max_bodies=max_bodies)
    e.world_set_bodies(st["id"], st["type"], st["radius"])
    return e


def device_state(e):
    info = e.world_sync()
    out = e.download(temp=True)
    out.update(radius=info["radius"], type=info["type"], id=info["id"])
    return out, info


def assert_same_bits(e, w, what):
    got, info = device_state(e)
    want = w.state()
    assert info["n"] == len(w), what
    for k in KEYS:
        assert np.array_equal(got[k], want[k]), (what, k)
    assert info["rng"] == w.rng, what
    c = w.counters()
    assert info["counters"] == {"merges": c["merge"], "fractures": c["fracture"], "overlaps": c["overlap"], "flashes": c["flash"],
                                "explosions": c["explosion"]}, what
    assert np.array_equal(e.world_contacts(), w.contacts()), what
    return info


@pytest.mark.parametrize("speed,nsteps,merge_time,expect", [(0.3, 90, 0.5, "overlaps"), (3.0, 60, 0.012, "merges"), (8.0, 30, 0.5, "fractures")])
def test_world_collision_stage_bit_identical(nb, oracle, speed, nsteps, merge_time, expect):
    """Resting and bouncing contacts (1,524 responses in 90 calls), merges (CONFIG.mergeTime lowered: without gravity nothing
    holds two bodies together for half a second) and fractures."""
    w = crowd(oracle, 70, 3, 5.0, speed)
    w.config(mergeTime=merge_time)
    e = mirror(nb, w)
    e.world_config(merge_time=merge_time)
    sizes = set()
    for step in range(nsteps):
        w.integrate(0.008, 0.0, 0.5, collisions=True, nsteps=1)        # G = 0: the step is drift + collisions on both sides
        e.world_step(0.008, 0.0, 0.5, 1, collisions=True)
        info = assert_same_bits(e, w, f"step {step}")
        sizes.add(info["n"])
    assert info["counters"][expect] > 0 and (expect == "overlaps" or len(sizes) > 1)
    print(f"speed {speed}: {info['counters']} after {nsteps} calls, {info['n']} bodies, {info['n_contacts']} contacts")
    e.close()


def test_world_several_steps_per_launch_and_the_300_body_shift(nb, oracle):
    """nsteps calls in ONE launch equal nsteps launches, and addBody's shift (idx:240-242) drops the oldest bodies when
    fragments push the list past CONFIG.maxBodies."""
    w = crowd(oracle, 80, 5, 4.0, 9.0, max_bodies=90)
    e = mirror(nb, w, max_bodies=90)
    for block in range(5):
        w.integrate(0.008, 0.0, 0.5, collisions=True, nsteps=6)
        e.world_step(0.008, 0.0, 0.5, 6, collisions=True)
        info = assert_same_bits(e, w, f"block {block}")
    assert info["counters"]["fractures"] > 0 and info["n"] == 90      # the list hit the cap and stayed there
    e.close()


def test_world_with_gravity_default_scene(nb, oracle):
    """The reference's own default scene (151 bodies, spheres overlapping from the first call) through 200 complete integrate
    calls = 100 frames of animate(): same bodies, same events, state within the fp64 gate's reach of the oracle's."""
    s = nb.scenes.galaxy_default()
    w = oracle.World(seed=0x5EED0001, max_bodies=300)
    rng = np.random.default_rng(1)
    for i in range(s.n):
        w.add(s.pos_mass[i, 3], s.pos_mass[i, :3], s.vel[i, :3], "metal" if i == 0 else TYPES[int(rng.integers(0, 3))], s.temp[i])
    e = mirror(nb, w)
    for frame in range(100):
        w.integrate(s.dt, s.G, s.softening, collisions=True, nsteps=2)
        e.world_step(s.dt, s.G, s.softening, 2, collisions=True)
    got, info = device_state(e)
    want = w.state()
    c = w.counters()
    assert info["n"] == len(w) and info["counters"]["overlaps"] == c["overlap"] and info["counters"]["merges"] == c["merge"]
    assert np.array_equal(got["id"], want["id"]) and np.array_equal(got["type"], want["type"])
    scale = np.abs(want["pos_mass"][:, :3]).max()
    assert np.abs(got["pos_mass"][:, :3] - want["pos_mass"][:, :3]).max() < 1e-9 * scale
    assert np.abs(got["vel"][:, :3] - want["vel"][:, :3]).max() < 1e-9 * np.abs(want["vel"][:, :3]).max()
    assert np.allclose(got["temp"], want["temp"], rtol=1e-9, atol=1e-12)
    print(f"default scene, 200 calls with collisions: {info['counters']}, {info['n']} bodies")
    # the frame the page draws: two complete steps per launch, state resident
    e.sync()
    t0 = time.perf_counter()
    frames = 200
    for _ in range(frames):
        e.world_step(s.dt, s.G, s.softening, 2, collisions=True)
    e.world_sync()
    us = (time.perf_counter() - t0) / frames * 1e6
    print(f"complete frame (2 x integrate with collisions, {info['n']} bodies, state on the device): {us:.1f} us")
    e.close()


def test_world_frame_is_step_plus_sync_plus_download(nb, oracle):
    """nbody_world_frame = nbody_world_step + nbody_world_sync + nbody_download through one launch and one synchronise: the
    same bits as the three calls, through merges and fractures (the body count changes under the caller's arrays)."""
    w = crowd(oracle, 80, 5, 4.0, 9.0, max_bodies=90)
    a, b = mirror(nb, w, max_bodies=90), mirror(nb, w, max_bodies=90)
    out, sizes = None, set()
    for block in range(6):
        w.integrate(0.008, 0.0, 0.5, collisions=True, nsteps=2)
        a.world_step(0.008, 0.0, 0.5, 2, collisions=True)
        got, info = device_state(a)
        out = b.world_frame(0.008, 0.0, 0.5, 2, collisions=True, out=out)
        want = w.state()
        assert out["n"] == info["n"] == len(w) == b.n
        for k in KEYS:
            assert np.array_equal(out[k], got[k]) and np.array_equal(out[k], want[k]), (block, k)
        assert np.array_equal(out["acc"], a.download()["acc"])
        assert out["rng"] == info["rng"] == w.rng and out["counters"] == info["counters"] and out["n_contacts"] == info["n_contacts"]
        sizes.add(out["n"])
    assert len(sizes) > 1 and out["counters"]["fractures"] > 0
    same = b.world_frame(0.008, 0.0, 0.5, 0)                          # nsteps = 0: just the record
    for k in KEYS + ("acc",):
        assert np.array_equal(same[k], out[k])
    d = b.download(temp=True)                                         # ... and the context's own arrays hold the same state
    assert np.array_equal(d["pos_mass"], out["pos_mass"]) and np.array_equal(d["temp"], out["temp"])
    # arrays too short for the device's list: an error, never a partial frame
    import ctypes
    n = ctypes.c_size_t()
    pm = np.zeros((4, 4))
    rc = b._L.nbody_world_frame(b._ctx, 0.008, 0.0, 0.5, 0, 1, 4, ctypes.byref(n), pm.ctypes.data, None, None, None, None, None, None, None, None, None)
    assert rc == nb._lib.NBODY_ENOMEM and n.value == out["n"] and not pm.any()
    a.close()
    b.close()


def test_world_frame_timing_default_scene(nb, oracle):
    s = nb.scenes.galaxy_default()
    w = oracle.World(seed=0x5EED0001, max_bodies=300)
    rng = np.random.default_rng(1)
    for i in range(s.n):
        w.add(s.pos_mass[i, 3], s.pos_mass[i, :3], s.vel[i, :3], "metal" if i == 0 else TYPES[int(rng.integers(0, 3))], s.temp[i])
    e = mirror(nb, w)
    out = None
    for _ in range(20):
        out = e.world_frame(s.dt, s.G, s.softening, 2, out=out)
    t0 = time.perf_counter()
    frames = 300
    for _ in range(frames):
        out = e.world_frame(s.dt, s.G, s.softening, 2, out=out)
    us = (time.perf_counter() - t0) / frames * 1e6
    print(f"complete frame with host arrays out (2 x integrate with collisions, {out['n']} bodies): {us:.1f} us")
    assert np.isfinite(out["pos_mass"]).all()
    e.close()


def test_world_without_collisions_equals_the_step_contract(nb, oracle):
    s = nb.scenes.galaxy_default()
    p, v, a, t = oracle.integrate(s.pos_mass, s.vel, np.zeros_like(s.vel), s.temp, s.dt, s.G, s.softening, nsteps=4)
    with nb.Engine(300, precision="f64") as e:
        e.upload(s.pos_mass, s.vel, None, s.temp)
        e.world_enable(seed=1, max_bodies=300)
        e.world_set_bodies(np.arange(s.n) / 1000.0, np.zeros(s.n, np.int32), np.full(s.n, 0.5))
        e.world_step(s.dt, s.G, s.softening, 4, collisions=False)
        e.world_sync()
        o = e.download(temp=True)
    for got, want in ((o["pos_mass"], p), (o["vel"], v), (o["acc"], a)):
        assert np.linalg.norm(got[:, :3] - want[:, :3], axis=1).max() < 1e-12 * np.linalg.norm(want[:, :3], axis=1).max()
    assert np.array_equal(o["temp"], t)


def test_world_pow_third_is_correctly_rounded(nb):
    """Body.radius = Math.pow(3m / (4 pi rho), 1/3) (idx:221) for bodies created on the device: the double-double evaluation
    must be the CORRECTLY ROUNDED x^c, c = the double nearest 1/3 (checked against 60-digit decimal arithmetic).  The host
    libm's pow -- the oracle's, and Python's -- is itself one ulp off that value for ~0.07 % of arguments (measured here too);
    V8's Math.pow is an fdlibm-derived routine with its own last-bit behaviour."""
    import ctypes
    import math
    from decimal import Decimal, getcontext
    from fractions import Fraction
    from nbody_b200 import _lib
    getcontext().prec = 60
    rng = np.random.default_rng(7)
    x = np.concatenate([10.0 ** rng.uniform(-12, 12, 6000), rng.uniform(0.01, 1000.0, 6000), [0.0, 1.0, 8.0, 27.0, 1e-300, 1e300]])
    y = np.zeros_like(x)
    with nb.Engine(16, precision="f64") as e:
        _lib.check(_lib.lib().nbody_selftest_pow_third(e._ctx, ctypes.c_void_p(x.ctypes.data), ctypes.c_void_p(y.ctypes.data), len(x)), e._ctx)
    c = Fraction(1.0 / 3.0)
    exact = np.array([0.0 if v == 0.0 else float((Decimal(float(v)).ln() * Decimal(c.numerator) / Decimal(c.denominator)).exp()) for v in x])
    libm = np.array([math.pow(v, 1.0 / 3.0) for v in x])
    bad = np.flatnonzero(y != exact)
    assert len(bad) == 0, (len(bad), x[bad[:5]], y[bad[:5]], exact[bad[:5]])
    off = int((libm != exact).sum())
    print(f"pow_third: {len(x)} arguments correctly rounded; the host libm's pow differs from the correctly rounded value for {off} of them")
    assert off < 0.005 * len(x) and np.abs(libm - exact).max() <= np.spacing(exact).max()


def test_world_error_paths(nb):
    with nb.Engine(64, precision="f32") as e:
        with pytest.raises(nb.NBodyError):
            e.world_enable()                          # fp64 only
    with nb.Engine(64, precision="f64") as e:
        with pytest.raises(nb.NBodyError):
            e.world_enable(max_bodies=4096)           # beyond the kernel's arrays / the capacity
        e.world_enable(max_bodies=64)
        e.upload(np.ones((3, 4)), np.zeros((3, 4)))
        with pytest.raises(nb.NBodyError):
            e.world_step(0.01, 1.0, 0.1, 1)           # before world_set_bodies

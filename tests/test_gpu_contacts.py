"""GPU tests of the collision stage: the device broad phase
(nbody_set_radius / nbody_contacts through the C ABI) against the oracle's restatement
of the overlap loop of resolveCollisions (idx:301-313), and the complete reference step
(gravity + collisions, merge, fracture) through the host mirror against the oracle.

Pair lists are index work: the bar is bit-exact (same pairs, same order).
"""
import math
import os

import numpy as np
import pytest

from .conftest import norm_rel_err

pytestmark = pytest.mark.gpu


def radius_of(mass, density=1.0):
    return (3 * mass / (4 * math.pi * density)) ** (1 / 3)   # Body.radius, idx:221


def crowd(n, seed, box, mass=(0.05, 0.4)):
    rng = np.random.default_rng(seed)
    pm = np.concatenate([rng.uniform(-box, box, (n, 3)), rng.uniform(*mass, (n, 1))], axis=1)
    vel = np.zeros((n, 4))
    vel[:, :3] = rng.normal(0, 0.5, (n, 3))
    return pm, vel, radius_of(pm[:, 3])


def test_f64_default_scene_pairs_exact(nb, oracle):
    s = nb.scenes.galaxy_default()
    dens = np.where(np.arange(s.n) == 0, 3.0, 1.0)
    rad = radius_of(s.pos_mass[:, 3], dens) * 1.5            # inflated so the scene has contacts
    with nb.Engine(s.n, precision="f64") as e:
        e.upload(s.pos_mass, s.vel)
        with pytest.raises(nb.NBodyError):
            e.contacts()                                     # no radii yet
        e.set_radius(rad)
        for margin in (0.0, 0.75, 5.0):
            got = e.contacts(margin)
            ref = oracle.overlap_pairs(s.pos_mass, rad, margin, mt=False)
            assert np.array_equal(got, ref), margin
        assert len(e.contacts(0.0)) > 0
        assert not e.contact_stats()["used_flags"]           # fp64: every body is scanned


@pytest.mark.parametrize("n,box,eps,mass", [(4096, 12.0, 0.5, (0.05, 0.4)), (5000, 9.0, 0.05, (0.05, 0.4)),
                                            (20000, 25.0, 0.0, (0.05, 0.4)), (5000, 10.0, 0.2, (0.2, 0.2))])
def test_f32_pairs_exact_with_and_without_flags(nb, oracle, n, box, eps, mass):
    """fp32 context: the fp32 pre-test only prefilters; the deciding comparison is fp64 on
    the stored fp32 state, so the set equals the oracle's on the same rounded inputs.  The
    list must be the same whether or not the force sweep's nearest-neighbour flags are used."""
    pm, vel, rad = crowd(n, 7 + n, box, mass)            # equal masses: the equal-mass variants of both sweeps
    pm32, vel32, rad32 = pm.astype(np.float32), vel.astype(np.float32), rad.astype(np.float32)
    ref_in = (pm32.astype(np.float64), rad32.astype(np.float64))
    with nb.Engine(n, precision="f32") as e:
        e.upload(pm32, vel32)
        e.set_radius(rad32)
        a = e.contacts(0.0)                                  # no force pass yet: full scan
        st_a = e.contact_stats()
        assert not st_a["used_flags"] and st_a["scanned"] == n
        ref = oracle.overlap_pairs(*ref_in, 0.0)
        assert len(ref) > 20
        assert np.array_equal(a, ref)
        e.force(1.0, eps)                                    # NEAR sweep (ZERO_SOFT variant when eps == 0)
        for margin in (0.0, 0.2):
            b = e.contacts(margin)
            st = e.contact_stats()
            assert st["used_flags"] and st["flagged"] < n
            assert st["candidates_only"] == (2 * st["flagged"] < n)      # partners sought among the candidates only
            assert np.array_equal(b, oracle.overlap_pairs(*ref_in, margin)), margin
            os.environ["NBODY_CONTACTS_COMPACT"] = "0"                   # candidates against every body: same list
            try:
                assert np.array_equal(e.contacts(margin), b) and not e.contact_stats()["candidates_only"]
            finally:
                del os.environ["NBODY_CONTACTS_COMPACT"]
        acc_near = e.download()["acc"]
        # the flags go stale when bodies move
        e.kick1_drift(0.01)
        e.contacts(0.0)
        assert not e.contact_stats()["used_flags"]
    # tracking the nearest neighbour must not change a single bit of the force
    with nb.Engine(n, precision="f32") as e:
        e.upload(pm32, vel32)
        e.force(1.0, eps)
        assert np.array_equal(e.download()["acc"], acc_near)


def test_f32_two_galaxies_131072_pairs_exact(nb, oracle):
    """Config 3's scene at full size, radii from the masses (central bodies dominate r_max)."""
    s = nb.scenes.two_galaxies(131072)
    pm, vel, _ = s.astype(np.float32)
    rad = radius_of(pm[:, 3].astype(np.float64)).astype(np.float32) * 4.0
    with nb.Engine(s.n, precision="f32") as e:
        e.upload(pm, vel)
        e.set_radius(rad)
        e.step(s.dt, s.G, s.softening, 2)                    # the last force pass of a step leaves valid flags
        got = e.contacts(0.0)
        st = e.contact_stats()
        out = e.download()
    ref = oracle.overlap_pairs(out["pos_mass"].astype(np.float64), rad.astype(np.float64), 0.0)
    assert st["used_flags"] and st["flagged"] < s.n // 2 and st["candidates_only"]   # candidates x candidates
    assert len(ref) > 0
    assert np.array_equal(got, ref)


def test_contacts_capacity_protocol(nb, oracle):
    from nbody_b200 import _lib
    import ctypes
    pm, vel, rad = crowd(2048, 3, 5.0)
    with nb.Engine(2048, precision="f64") as e:
        e.upload(pm, vel)
        e.set_radius(rad)
        full = e.contacts(0.0)
        assert len(full) > 4
        small = np.zeros((2, 2), dtype=np.uint32)
        cnt = ctypes.c_size_t()
        st = e._L.nbody_contacts(e._ctx, 0.0, ctypes.c_void_p(small.ctypes.data), 2, ctypes.byref(cnt))
        assert st == _lib.NBODY_ENOMEM and cnt.value == len(full) and not small.any()
        st = e._L.nbody_contacts(e._ctx, -1.0, ctypes.c_void_p(small.ctypes.data), 2, ctypes.byref(cnt))
        assert st == _lib.NBODY_EINVAL
        e.upload(pm, vel)                                    # a new body list drops the radii
        with pytest.raises(nb.NBodyError):
            e.contacts()


def test_multi_gpu_contacts_identical(nb, oracle):
    if nb.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    pm, vel, rad = crowd(6000, 21, 10.0)
    pm32, vel32, rad32 = pm.astype(np.float32), vel.astype(np.float32), rad.astype(np.float32)
    ref = oracle.overlap_pairs(pm32.astype(np.float64), rad32.astype(np.float64), 0.1)
    with nb.Engine(6000, precision="f32", n_gpus=2) as e:
        e.upload(pm32, vel32)
        e.set_radius(rad32)
        e.force(1.0, 0.3)
        assert np.array_equal(e.contacts(0.1), ref)
        assert e.contact_stats()["used_flags"]


# ---- the complete reference step through the host mirror --------------------------------

def _build(nb, oracle, bodies, seed, precision="f64", margin=0.5):
    ph = nb.physics
    ph.seed_random(seed)
    ps = ph.PhysicsSystem(precision=precision, collisions=True, contact_margin=margin)
    w = oracle.World(seed=seed, max_bodies=ph.CONFIG.maxBodies)
    for (m, p, v, ty, t) in bodies:
        ps.addBody(ph.Body(m, ph.V3(*p), ph.V3(*v), ty, t))
        w.add(m, p, v, ty, t)
    return ps, w


def _compare(ps, st, tol):
    assert len(ps.bodies) == st["pos_mass"].shape[0]
    pm = np.array([[*b.pos, b.mass] for b in ps.bodies])
    vel = np.array([[*b.vel, 0.0] for b in ps.bodies])
    assert np.array_equal(pm[:, 3], st["pos_mass"][:, 3])
    assert norm_rel_err(pm, st["pos_mass"]) < tol
    assert norm_rel_err(vel, st["vel"]) < tol
    np.testing.assert_allclose([b.temperature for b in ps.bodies], st["temp"], rtol=1e-9, atol=1e-12)


@pytest.mark.parametrize("speed", [0.3, 8.0])
def test_full_step_with_collisions_f64(nb, oracle, speed):
    """integrate(dt) COMPLETE (idx:247-262 with idx:256): device gravity + device overlap
    search + host response, against the oracle's restatement, merges and fractures included."""
    ph = nb.physics
    rng = np.random.default_rng(11)
    types = ("rock", "metal", "ice")
    bodies = [(float(rng.uniform(0.3, 3.0)), tuple(rng.uniform(-7, 7, 3)), tuple(rng.normal(0, speed, 3)),
               types[int(rng.integers(0, 3))], float(rng.uniform(0, 5))) for _ in range(120)]
    ps, w = _build(nb, oracle, bodies, seed=99)
    steps = 120 if speed < 1 else 40
    for _ in range(steps):
        ps.integrate(0.008)
        w.integrate(0.008, ph.CONFIG.G, ph.CONFIG.softening, True, 1)
    c = w.counters()
    assert (c["overlap"], c["merge"], c["fracture"]) == (ps.stats["overlap"], ps.stats["merge"], ps.stats["fracture"])
    assert c["merge"] > 0 if speed < 1 else c["fracture"] > 0
    _compare(ps, w.state(), 1e-10)
    ph.seed_random(None)


def test_default_galaxy_complete_step_f64(nb, oracle):
    """The reference's own default scene with its own complete step, 2 sub-steps per frame."""
    ph = nb.physics
    ph.seed_random(5)
    ps = ph.PhysicsSystem(precision="f64", collisions=True)
    ph.resetScenario(ps, "galaxy")
    w = oracle.World(seed=0, max_bodies=300)
    for b in ps.bodies:
        w.add(b.mass, list(b.pos), list(b.vel), b.type, b.temperature)
    ph.seed_random(1234)
    w.seed(1234)
    for _ in range(100):
        ps.integrate(0.008, nsteps=2)                        # animate(), idx:877-879
        w.integrate(0.008, 0.5, 0.5, True, 2)
    assert w.counters()["overlap"] == ps.stats["overlap"] > 0
    _compare(ps, w.state(), 1e-10)
    ph.seed_random(None)

"""Collision stage on the CPU: pins for the oracle's restatement
of resolveCollisions / mergeBodies / fractureBodies (idx:293-443; the reference has no
tests, so analytic known answers are the pins), and the host mirror
(nbody_b200.physics.PhysicsSystem) against the oracle with the device replaced by a
stand-in built from oracle functions (test infrastructure)."""
import math

import numpy as np
import pytest


def _state_equal(ps, st):
    assert len(ps.bodies) == st["pos_mass"].shape[0]
    for i, b in enumerate(ps.bodies):
        assert (b.pos.x, b.pos.y, b.pos.z, b.mass) == tuple(st["pos_mass"][i]), i
        assert (b.vel.x, b.vel.y, b.vel.z) == tuple(st["vel"][i, :3]), i
        assert (b.acc.x, b.acc.y, b.acc.z) == tuple(st["acc"][i, :3]), i
        assert b.temperature == st["temp"][i], i
        assert b.radius == st["radius"][i], i


# ---- oracle pins ---------------------------------------------------------------------

def test_head_on_restitution_known_answer(oracle):
    # equal masses, overlap, closing speed 2v along x: normal impulse with restitution 0.2
    # reverses the relative velocity and scales it by 0.2 (idx:328-330, 361-362); no tangent
    m, v = 2.0, 3.0
    w = oracle.World()
    r = (3 * m / (4 * math.pi)) ** (1 / 3)
    w.add(m, [-0.5 * r, 0, 0], [v, 0, 0])
    w.add(m, [0.5 * r, 0, 0], [-v, 0, 0])
    s0 = w.state()
    assert s0["radius"][0] == pytest.approx(r, rel=1e-15)
    w.resolve_collisions(0.008)
    s = w.state()
    np.testing.assert_allclose(s["vel"][:, 0], [-0.2 * v, 0.2 * v], rtol=1e-15)
    # impactEnergy = 0.5*mu*van^2 = 0.5*(m/2)*(2v)^2 = m v^2 ; each body heats by (E/m)*0.2 (idx:332-336)
    np.testing.assert_allclose(s["temp"], [0.2 * v * v] * 2, rtol=1e-15)
    # positional correction: 0.8*(minDist-dist) split by inverse mass (idx:350-352)
    overlap = 2 * r - r
    np.testing.assert_allclose(s["pos_mass"][:, 0], [-0.5 * r - 0.4 * overlap, 0.5 * r + 0.4 * overlap], rtol=1e-15)
    assert w.n_contacts == 1 and w.counters()["overlap"] == 1


def test_separating_pair_only_counts_contact_time(oracle):
    w = oracle.World()
    w.add(1.0, [0, 0, 0], [-1, 0, 0])
    w.add(1.0, [0.5, 0, 0], [1, 0, 0])
    s0 = w.state()
    w.resolve_collisions(0.1)
    s = w.state()
    for k in ("pos_mass", "vel", "temp"):
        assert np.array_equal(s[k], s0[k])             # velAlongNormal >= 0: nothing moves (idx:327)
    assert w.n_contacts == 1
    w2 = oracle.World()
    w2.add(1.0, [0, 0, 0], [0, 0, 0])
    w2.add(1.0, [50, 0, 0], [0, 0, 0])
    w2.resolve_collisions(0.1)
    assert w2.n_contacts == 0


def test_merge_after_contact_time_conserves_mass_and_momentum(oracle):
    # slow approach, Q small: the pair merges only when the accumulated contact time exceeds
    # mergeTime = 0.5 while the bodies are closing (idx:327, 340)
    def world():
        w = oracle.World()
        w.add(3.0, [0, 0, 0], [0.01, 0.02, 0], "rock", 5.0)
        w.add(1.0, [0.3, 0, 0], [-0.01, 0, 0.03], "ice", 1.0)
        return w
    w = world()
    w.resolve_collisions(0.1)                           # 0.1 < mergeTime: bounce, no merge
    assert len(w) == 2 and w.counters()["merge"] == 0 and w.n_contacts == 1
    w = world()
    s0 = w.state()
    p0 = (s0["pos_mass"][:, 3:4] * s0["vel"][:, :3]).sum(0)
    w.resolve_collisions(0.6)                           # 0.6 > mergeTime on a closing pair: merge
    s = w.state()
    assert len(w) == 1 and w.counters()["merge"] == 1 and w.n_contacts == 0
    assert s["pos_mass"][0, 3] == 4.0
    np.testing.assert_allclose(s["pos_mass"][0, 3] * s["vel"][0, :3], p0, rtol=1e-12, atol=1e-15)
    assert s["type"][0] == 0                            # the heavier body's type (idx:403)
    assert s["radius"][0] == pytest.approx((3 * 4.0 / (4 * math.pi)) ** (1 / 3), rel=1e-15)
    # impact heating happens before the merge, then the mass-weighted mean (idx:335-336, 402)
    assert s["temp"][0] > (5.0 * 3 + 1.0 * 1) / 4


def test_fracture_fragments(oracle):
    # Q = E/(mA+mB) > fractureThreshold = 25 (idx:348): equal masses need closing speed > sqrt(200)
    w = oracle.World(seed=7)
    w.add(4.0, [0, 0, 0], [20, 0, 0], "rock", 3.0)
    w.add(4.0, [0.9, 0.1, 0], [-20, 0, 0], "metal", 9.0)
    w.resolve_collisions(0.008)
    c = w.counters()
    assert c["fracture"] == 1 and c["merge"] == 0
    s = w.state()
    n = len(w)
    assert 1 <= n <= 17 and c["explosion"] == n       # floor(3 + U*3*severity), severity <= 5 (idx:417-418)
    assert s["pos_mass"][:, 3].sum() <= 8.0 * (1 + 1e-15)
    assert (s["pos_mass"][:, 3] >= 0.2).all()          # minFragmentMass (idx:427)
    E = 0.5 * 2.0 * 40.0 ** 2
    assert (s["temp"] >= 9.0 + E / 8.0 * 0.1 - 1e-9).all()   # idx:435 (impact heating came first)
    assert (s["acc"] == 0).all()                        # new Body: acc = 0 (idx:217)


def test_max_bodies_shift(oracle):
    w = oracle.World(max_bodies=3)
    for k in range(5):
        w.add(1.0 + k, [10.0 * k, 0, 0], [0, 0, 0])
    assert len(w) == 3 and list(w.state()["pos_mass"][:, 3]) == [3.0, 4.0, 5.0]   # idx:240-242


def test_overlap_pairs_matches_reference_order(oracle):
    rng = np.random.default_rng(3)
    n = 400
    pm = np.concatenate([rng.uniform(-6, 6, (n, 3)), rng.uniform(0.5, 2.5, (n, 1))], axis=1)
    rad = (3 * pm[:, 3] / (4 * math.pi)) ** (1 / 3)
    for margin in (0.0, 0.3):
        a = oracle.overlap_pairs(pm, rad, margin, mt=False)
        b = oracle.overlap_pairs(pm, rad, margin, mt=True)
        assert np.array_equal(a, b) and len(a) > 50
        keys = a[:, 0].astype(np.int64) * n + a[:, 1]
        assert (np.diff(keys) > 0).all() and (a[:, 0] < a[:, 1]).all()
        d = np.linalg.norm(pm[a[:, 1], :3] - pm[a[:, 0], :3], axis=1)
        assert (d < rad[a[:, 0]] + rad[a[:, 1]] + margin + 1e-12).all()


# ---- host mirror vs oracle -------------------------------------------------------------

class OracleDevice:
    """Stand-in for nbody_b200.Engine built from oracle functions, so the mirror's
    host logic (marshalling, candidate-pair response, margin retry) runs without a GPU."""
    dtype = np.float64

    def __init__(self, oracle):
        self.o = oracle
        self.heat = 0.998

    def set_heat_decay(self, h):
        self.heat = h

    def upload(self, pm, vel, acc=None, temp=None):
        self.pm, self.vel = pm.copy(), vel.copy()
        self.acc = np.zeros_like(vel) if acc is None else acc.copy()
        self.temp = None if temp is None else temp.copy()
        self.n = pm.shape[0]

    def set_radius(self, r):
        self.radius = np.asarray(r, dtype=np.float64).copy()

    def kick1_drift(self, dt):
        self.pm, self.vel = self.o.kick1_drift(self.pm, self.vel, self.acc, dt)

    def force(self, G, eps):
        self.acc = self.o.compute_gravity(self.pm, G, eps)

    def kick2(self, dt, heat):
        self.vel, self.temp = self.o.kick2(self.vel, self.acc, self.temp, dt, heat)

    def contacts(self, margin=0.0):
        return self.o.overlap_pairs(self.pm, self.radius, margin, mt=False)

    def download(self, pos_mass=True, vel=True, acc=True, temp=False):
        return {"pos_mass": self.pm.copy(), "vel": self.vel.copy(), "acc": self.acc.copy(), "temp": self.temp.copy()}


def _mirror_and_world(nb, oracle, bodies, seed, margin=0.5):
    ph = nb.physics
    ph.seed_random(seed)
    ps = ph.PhysicsSystem(precision="f64", collisions=True, contact_margin=margin)
    ps._engine_obj = OracleDevice(oracle)
    w = oracle.World(seed=seed, max_bodies=ph.CONFIG.maxBodies)
    for (m, p, v, ty, t) in bodies:
        ps.addBody(ph.Body(m, ph.V3(*p), ph.V3(*v), ty, t))
        w.add(m, p, v, ty, t)
    return ps, w


def _crowd(n, seed, box, speed):
    rng = np.random.default_rng(seed)
    types = ("rock", "metal", "ice")
    return [(float(rng.uniform(0.3, 3.0)), tuple(rng.uniform(-box, box, 3)), tuple(rng.normal(0, speed, 3)),
             types[int(rng.integers(0, 3))], float(rng.uniform(0, 5))) for _ in range(n)]


@pytest.mark.parametrize("speed,margin", [(0.3, 0.5), (8.0, 0.5), (8.0, 1e-3)])
def test_mirror_full_step_equals_oracle(nb, oracle, speed, margin):
    """integrate(dt) with collisions: the mirror (candidate pairs + sequential response)
    reproduces the oracle's full double loop bit for bit, merges and fractures included,
    even when the margin is so small that the response has to be redone (margin retry)."""
    ph = nb.physics
    ps, w = _mirror_and_world(nb, oracle, _crowd(120, 11, 7.0, speed), seed=99, margin=margin)
    for step in range(120 if speed < 1 else 40):
        ps.integrate(0.008)
        w.integrate(0.008, ph.CONFIG.G, ph.CONFIG.softening, True, 1)
        _state_equal(ps, w.state())
    c = w.counters()
    assert c["overlap"] == ps.stats["overlap"] and c["merge"] == ps.stats["merge"] and c["fracture"] == ps.stats["fracture"]
    if speed < 1:
        assert c["merge"] > 0
    if speed > 1:
        assert c["fracture"] > 0
    if margin < 0.01:
        assert ps.stats["margin_retries"] > 0
    assert len(ps.contactPairs) == w.n_contacts
    ph.seed_random(None)


def test_mirror_default_galaxy_with_collisions(nb, oracle):
    """The reference's own default scene (idx:749-759), complete step, 200 calls."""
    ph = nb.physics
    ph.seed_random(5)
    ps = ph.PhysicsSystem(precision="f64", collisions=True)
    ps._engine_obj = OracleDevice(oracle)
    ph.resetScenario(ps, "galaxy")
    w = oracle.World(seed=0, max_bodies=300)
    for b in ps.bodies:
        w.add(b.mass, list(b.pos), list(b.vel), b.type, b.temperature)
    ph.seed_random(1234)
    w.seed(1234)
    for _ in range(200):
        ps.integrate(0.008)
        w.integrate(0.008, 0.5, 0.5, True, 1)
    _state_equal(ps, w.state())
    assert w.counters()["overlap"] > 0                  # overlaps exist from step 0
    ph.seed_random(None)


def test_full_loop_and_pair_list_agree(nb, oracle):
    """resolveCollisions(dt) with pairs=None is the reference's own double loop."""
    ph = nb.physics
    bodies = _crowd(60, 5, 4.0, 2.0)
    a, _ = _mirror_and_world(nb, oracle, bodies, seed=3)
    b, w = _mirror_and_world(nb, oracle, bodies, seed=3)
    ph.seed_random(42)
    a.resolveCollisions(0.008, None)
    ph.seed_random(42)
    w.seed(42)
    pm = np.array([[*q.pos, q.mass] for q in b.bodies])
    rad = np.array([q.radius for q in b.bodies])
    b.resolveCollisions(0.008, oracle.overlap_pairs(pm, rad, 2.0))
    w.resolve_collisions(0.008)
    _state_equal(a, w.state())
    _state_equal(b, w.state())
    ph.seed_random(None)

"""The drop-in itself, executed: the reference page's own `physics` object (Body, PhysicsSystem, CONFIG run from
index.html's text, oracle/js_reference.py) with `PhysicsSystem.integrate` replaced by THIS repository's
n-body-sim_b200/addon/physics_shim.js (run from its text the same way), calling an addon-shaped object with typed arrays.
No Node exists here, so the N-API boundary is stood in for by `Addon` below: the addon's method names and argument
conventions over an engine -- the CPU oracle (tests marked live, no GPU) or nbody_b200.Engine (GPU tests).
The unchanged page next to it is the reference; both sides draw the same Math.random stream."""
import numpy as np
import pytest

from oracle import js_reference as jr

live = pytest.mark.skipif(not jr.available(), reason="/root/reference/index.html is not on this box")
KEYS = ("pos_mass", "vel", "acc", "temp", "radius", "type")


class OracleEngine:
    """nbody_b200.Engine's methods from oracle functions (fp64), for the CPU run of the shim."""

    def __init__(self, oracle, capacity):
        self.o, self.heat, self.capacity = oracle, 0.998, capacity

    def close(self):
        self.pm = None

    def set_heat_decay(self, h):
        self.heat = h

    def upload(self, pm, vel, acc=None, temp=None):
        assert pm.shape[0] <= self.capacity
        self.pm, self.vel = np.array(pm, dtype=np.float64), np.array(vel, dtype=np.float64)
        self.acc = np.zeros_like(self.vel) if acc is None else np.array(acc, dtype=np.float64)
        self.temp = None if temp is None else np.array(temp, dtype=np.float64)

    def set_radius(self, r):
        self.radius = np.array(r, dtype=np.float64)

    def step(self, dt, G, eps, nsteps):
        self.pm, self.vel, self.acc, self.temp = self.o.integrate(self.pm, self.vel, self.acc, self.temp, dt, G, eps, self.heat, nsteps)

    def step_host(self, pm, vel, acc, temp, dt, G, eps, nsteps):
        self.upload(pm, vel, acc, temp)
        self.step(dt, G, eps, nsteps)
        pm[:], vel[:], acc[:] = self.pm, self.vel, self.acc
        if temp is not None:
            temp[:] = self.temp

    def kick1_drift(self, dt):
        self.pm, self.vel = self.o.kick1_drift(self.pm, self.vel, self.acc, dt)

    def force(self, G, eps):
        self.acc = self.o.compute_gravity(self.pm, G, eps)

    def kick2(self, dt, heat):
        self.vel, self.temp = self.o.kick2(self.vel, self.acc, self.temp, dt, heat)

    def contacts(self, margin=0.0):
        return self.o.overlap_pairs(self.pm, self.radius, margin, mt=False)

    def download(self, pos_mass=True, vel=True, acc=True, temp=False):
        return {"pos_mass": self.pm, "vel": self.vel, "acc": self.acc, "temp": self.temp}


class Addon:
    """The N-API addon's surface (n-body-sim_b200/addon/nbody_napi.cc) over an engine factory."""

    def __init__(self, make_engine):
        self.make_engine, self.calls, self.created = make_engine, {}, 0

    def _count(self, name):
        self.calls[name] = self.calls.get(name, 0) + 1

    def create(self, precision, nGpus, capacity):
        self.created += 1
        return self.make_engine(precision, int(capacity))

    def destroy(self, ctx):
        ctx.close()

    def setHeatDecay(self, ctx, h):
        ctx.set_heat_decay(h)

    def upload(self, ctx, pm, vel, acc, temp):
        self._count("upload")
        ctx.upload(pm.reshape(-1, 4), vel.reshape(-1, 4), None if acc is None else acc.reshape(-1, 4), temp)

    def download(self, ctx, pm, vel, acc, temp):
        self._count("download")
        out = ctx.download(temp=True)
        for dst, key in ((pm, "pos_mass"), (vel, "vel"), (acc, "acc"), (temp, "temp")):
            if dst is not None:
                dst[:] = out[key].reshape(-1)

    def step(self, ctx, dt, G, eps, nsteps):
        self._count("step")
        ctx.step(dt, G, eps, nsteps)

    def stepHost(self, ctx, pm, vel, acc, temp, dt, G, eps, nsteps):
        self._count("stepHost")
        ctx.step_host(pm.reshape(-1, 4), vel.reshape(-1, 4), acc.reshape(-1, 4), temp, dt, G, eps, nsteps)

    def kick1Drift(self, ctx, dt):
        ctx.kick1_drift(dt)

    def force(self, ctx, G, eps):
        self._count("force")
        ctx.force(G, eps)

    def kick2(self, ctx, dt, heat):
        ctx.kick2(dt, heat)

    def setRadius(self, ctx, r):
        ctx.set_radius(r)

    def contacts(self, ctx, margin, pairs_out):
        self._count("contacts")
        pairs = ctx.contacts(margin)
        if 2 * len(pairs) <= pairs_out.length:
            pairs_out[:2 * len(pairs)] = pairs.reshape(-1)
        return len(pairs)


def _pages(seed, bodies, collisions):
    """The unchanged page and the page that gets the shim, same bodies, same Math.random stream."""
    out = []
    for _ in range(2):
        r = jr.Reference(seed=seed, collisions=collisions)
        if bodies == "galaxy":
            r.resetScenario("galaxy")
        else:
            for (m, p, v, ty, t) in bodies:
                r.add(m, p, v, ty, t)
        out.append(r)
    return out


def _crowd(n, seed, box, speed):
    rng = np.random.default_rng(seed)
    return [(float(rng.uniform(0.3, 3.0)), tuple(rng.uniform(-box, box, 3)), tuple(rng.normal(0, speed, 3)),
             jr.TYPES[int(rng.integers(0, 3))], float(rng.uniform(0, 5))) for _ in range(n)]


def _assert_same(a, b, exact, tol=1e-12):
    sa, sb = a.state(), b.state()
    assert len(sa["temp"]) == len(sb["temp"])
    for k in KEYS:
        if exact or k in ("radius", "type"):
            assert np.array_equal(sa[k], sb[k]), k
        else:
            scale = max(1.0, float(np.abs(sb[k]).max()))
            assert float(np.abs(sa[k] - sb[k]).max()) <= tol * scale, k


# ---------------------------------------------------------------- CPU: the oracle behind the addon surface

@live
def test_shim_gravity_only_is_one_addon_call_per_integrate(oracle):
    page, dropin = _pages(5, "galaxy", collisions=False)
    addon = Addon(lambda prec, cap: OracleEngine(oracle, cap))
    handle = jr.install_shim(dropin, addon, collisions=False)
    for frame in range(5):                                  # animate(): subSteps x integrate(dt / subSteps), idx:877-879
        page.integrate(0.008, 2)
        dropin.integrate(0.008, 2)
        _assert_same(dropin, page, exact=True)
    assert addon.calls == {"stepHost": 10} and addon.created == 1
    assert handle.ctx is not None
    handle.dispose()


@live
def test_shim_binds_bodies_to_the_arrays_only_when_the_list_changes(oracle):
    """Body.pos/vel/acc (and mass, temperature) are views of the addon's typed arrays, re-bound only on
    spawn (idx:835), the 300-body shift of addBody (idx:240-242), a collision that edits the list (idx:383-389) and reset
    (idx:745); a gravity-only integrate() is one addon call with no per-body work in between."""
    page, dropin = _pages(7, _crowd(40, 11, 40.0, 0.2), collisions=False)
    addon = Addon(lambda prec, cap: OracleEngine(oracle, cap))
    handle = jr.install_shim(dropin, addon, collisions=False)
    for _ in range(6):
        page.integrate(0.008, 2)
        dropin.integrate(0.008, 2)
        _assert_same(dropin, page, exact=True)
    assert handle.binds == 1 and addon.calls == {"stepHost": 12}
    # the Vector3 objects are still the page's own; their fields read and write the arrays the addon works on
    b = dropin.physics.bodies[3]
    pos_object = b.pos
    b.pos.x = 123.5
    b.vel.addScaledVector(b.acc, 2.0)
    b.temperature = 7.25
    pb = page.physics.bodies[3]
    pb.pos.x = 123.5
    pb.vel.addScaledVector(pb.acc, 2.0)
    pb.temperature = 7.25
    page.integrate(0.008)
    dropin.integrate(0.008)
    _assert_same(dropin, page, exact=True)
    assert handle.binds == 1 and dropin.physics.bodies[3].pos is pos_object
    # spawn (idx:835): addBody -> one re-bind at the next call, none after
    for r in (page, dropin):
        r.add(2.0, (5.0, 1.0, -3.0), (0.1, 0.0, 0.2), "ice", 4.0)
    for _ in range(3):
        page.integrate(0.008)
        dropin.integrate(0.008)
        _assert_same(dropin, page, exact=True)
    assert handle.binds == 2
    # the 300-body shift (idx:240-242): same array, same length, other bodies
    for r in (page, dropin):
        r.CONFIG.maxBodies = len(r.physics.bodies)
        r.add(1.5, (-8.0, 2.0, 9.0), (0.0, 0.3, 0.0), "metal", 0.0)
    gone = dropin.physics.bodies        # (the list object is the same one)
    assert len(page.physics.bodies) == 41 == len(dropin.physics.bodies)
    page.integrate(0.008, 2)
    dropin.integrate(0.008, 2)
    _assert_same(dropin, page, exact=True)
    assert handle.binds == 3 and dropin.physics.bodies is gone
    # reset (idx:745): a new array
    for r in (page, dropin):
        r.CONFIG.maxBodies = 300
        r.resetScenario("galaxy")
    page.integrate(0.008, 2)
    dropin.integrate(0.008, 2)
    _assert_same(dropin, page, exact=True)
    assert handle.binds == 4 and len(dropin.physics.bodies) == 151
    # dispose: the bodies get plain numbers back
    keep = dropin.physics.bodies[5]
    x = keep.pos.x
    handle.dispose()
    keep.pos.x = x + 1.0
    assert keep.pos.x == x + 1.0 and dropin.state()["pos_mass"][5, 0] == x + 1.0


@live
def test_shim_releases_bodies_that_left_the_list(oracle):
    """A Body removed by a merge keeps its last values as plain numbers: a stale reference cannot alias another body's row."""
    page, dropin = _pages(42, _crowd(70, 3, 5.0, 0.3), collisions=True)
    addon = Addon(lambda prec, cap: OracleEngine(oracle, cap))
    handle = jr.install_shim(dropin, addon)
    before = list(dropin.physics.bodies)
    for _ in range(80):
        page.integrate(0.008)
        dropin.integrate(0.008)
    _assert_same(dropin, page, exact=True)
    now = set(map(id, dropin.physics.bodies))
    left = [b for b in before if id(b) not in now]
    assert left and handle.binds > 1
    snapshot = dropin.state()
    for b in left:
        b.pos.x += 1000.0                               # writes to a released body touch nothing else
        b.temperature = 1e9
    after = dropin.state()
    for k in KEYS:
        assert np.array_equal(snapshot[k], after[k]), k


@live
@pytest.mark.parametrize("mode", [None, "gpu-broad-phase"])
@pytest.mark.parametrize("speed,nsteps", [(0.3, 80), (8.0, 25)])
def test_shim_with_the_pages_own_collision_stage(oracle, mode, speed, nsteps):
    """kick-1/drift, force and kick-2 behind the addon; resolveCollisions, mergeBodies, fractureBodies, addBody are the
    page's own code on the page's own Body objects (idx:256): identical to the unchanged page through merges/fractures."""
    page, dropin = _pages(42, _crowd(70, 3, 5.0, speed), collisions=True)
    addon = Addon(lambda prec, cap: OracleEngine(oracle, cap))
    opts = {} if mode is None else {"collisions": mode}
    jr.install_shim(dropin, addon, **opts)
    sizes = set()
    for step in range(nsteps):
        page.integrate(0.008)
        dropin.integrate(0.008)
        _assert_same(dropin, page, exact=True)
        sizes.add(len(page.physics.bodies))
    assert len(sizes) > 1                                   # bodies merged or fractured on the way
    assert page.physics.contactPairs.size == dropin.physics.contactPairs.size
    assert page.Math.draws == dropin.Math.draws
    if mode == "gpu-broad-phase":
        assert addon.calls["contacts"] == nsteps


@live
def test_shim_skips_the_host_stage_when_nothing_overlaps(oracle):
    bodies = [(1.0, (10.0 * k, 0.0, 0.0), (0.0, 0.1 * k, 0.0), "rock", 60.0) for k in range(6)]
    page, dropin = _pages(1, bodies, collisions=True)
    addon = Addon(lambda prec, cap: OracleEngine(oracle, cap))
    jr.install_shim(dropin, addon, collisions="gpu-broad-phase")
    called = []
    original = dropin.physics.resolveCollisions
    dropin.physics.resolveCollisions = lambda dt: called.append(dt) or original(dt)
    for _ in range(4):
        page.integrate(0.008)
        dropin.integrate(0.008)
    _assert_same(dropin, page, exact=True)
    assert called == [] and addon.calls["download"] == 4    # one read-back per step, bodies never left the device in between


@live
def test_shim_grows_its_context_and_compute_gravity(oracle):
    page, dropin = _pages(9, _crowd(12, 1, 30.0, 0.1), collisions=False)
    addon = Addon(lambda prec, cap: OracleEngine(oracle, cap))
    jr.install_shim(dropin, addon, collisions=False, capacity=8)     # smaller than the body list: ensure() re-creates
    page.integrate(0.008, 3)
    dropin.integrate(0.008, 3)
    _assert_same(dropin, page, exact=True)
    assert addon.created == 2
    page.physics.computeGravity()
    dropin.physics.computeGravity()                                 # idx:264-291 through addon.force
    _assert_same(dropin, page, exact=True)


# ---------------------------------------------------------------- GPU: libnbody.so behind the addon surface
# (the GPU box has no /root/reference: the page object is stood in for by plain records with the fields the shim
#  touches -- bodies[*].pos/vel/acc/mass/temperature/radius, contactPairs, resolveCollisions -- and the CPU oracle,
#  itself pinned to the page above and in test_reference_source.py, is the reference)

class _Body:
    def __init__(self, pm, vel, temp, radius):
        self.pos, self.vel, self.acc = jr.Vector3(*map(float, pm[:3])), jr.Vector3(*map(float, vel[:3])), jr.Vector3(0, 0, 0)
        self.mass, self.temperature, self.radius = float(pm[3]), float(temp), float(radius)


class _Page:
    def __init__(self, scene, seed=1):
        class Physics:
            pass
        self.physics = Physics()
        rad = (3 * scene.pos_mass[:, 3] / (4 * np.pi)) ** (1 / 3)
        self.physics.bodies = jr.JSArray(_Body(scene.pos_mass[i], scene.vel[i], scene.temp[i], rad[i]) for i in range(scene.n))
        self.physics.contactPairs = jr.JSMap()
        self.physics.contactPairs.set("stale", 1.0)
        self.resolve_calls = []
        self.physics.resolveCollisions = self.resolve_calls.append      # the page's stage, here a recorder (gravity-only parity)
        self.CONFIG = jr.JSObject({"G": scene.G, "softening": scene.softening, "heatDecay": 0.998, "maxBodies": 300})
        self.Math = jr._Math(seed)

    def arrays(self):
        b = self.physics.bodies
        return (np.array([[x.pos.x, x.pos.y, x.pos.z, x.mass] for x in b]), np.array([[x.vel.x, x.vel.y, x.vel.z, 0] for x in b]),
                np.array([[x.acc.x, x.acc.y, x.acc.z, 0] for x in b]), np.array([x.temperature for x in b]))


def _err(a, b):
    return float(np.linalg.norm(a[:, :3] - b[:, :3], axis=1).max() / max(np.linalg.norm(b[:, :3], axis=1).max(), 1e-300))


@pytest.mark.gpu
@pytest.mark.parametrize("prec,tol", [("f64", 1e-12), ("f32", 3e-5)])
@pytest.mark.parametrize("mode", [False, None, "gpu-broad-phase"])
def test_shim_over_the_engine(nb, oracle, prec, tol, mode):
    """physics_shim.js (executed from its text) driving libnbody.so through the addon surface: gravity-only mode is one
    stepHost per integrate; the default and gpu-broad-phase modes are kick1Drift + force, the page's collision stage (here a
    recorder, so that the oracle's gravity-only integrate is the reference), kick2."""
    s = nb.scenes.galaxy_default()
    page = _Page(s)
    addon = Addon(lambda p, cap: nb.Engine(cap, precision=p))
    handle = jr.install_shim(page, addon, precision=prec, **({} if mode is None else {"collisions": mode}))
    p, v, a, t = s.pos_mass, s.vel, np.zeros_like(s.vel), s.temp
    for frame in range(3):
        for _ in range(2):                                  # animate(): subSteps = 2, idx:877-879
            page.physics.integrate(s.dt)
        p, v, a, t = oracle.integrate(p, v, a, t, s.dt, s.G, s.softening, nsteps=2)
        gp, gv, ga, gt = page.arrays()
        assert _err(gp, p) < tol * (frame + 1) and _err(gv, v) < tol * (frame + 1) and _err(ga, a) < tol * (frame + 1)
        if prec == "f64":
            assert np.array_equal(gt, t)
    if mode is False:
        assert addon.calls == {"stepHost": 6} and page.resolve_calls == []
    elif mode is None:
        assert page.resolve_calls == [s.dt] * 6 and addon.calls["force"] == 6
    else:   # the default scene has overlapping spheres from step 0: the host stage runs
        assert addon.calls["contacts"] == 6 and len(page.resolve_calls) > 0
    assert handle.ctx.n == s.n
    handle.dispose()


@pytest.mark.gpu
def test_shim_over_the_engine_skips_host_stage_without_overlaps(nb, oracle):
    s = nb.scenes.plummer(64)
    page = _Page(s)
    for b in page.physics.bodies:
        b.radius = 1e-6
    addon = Addon(lambda p, cap: nb.Engine(cap, precision=p))
    jr.install_shim(page, addon, collisions="gpu-broad-phase")
    for _ in range(3):
        page.physics.integrate(s.dt)
    p, v, a, t = oracle.integrate(s.pos_mass, s.vel, np.zeros_like(s.vel), s.temp, s.dt, s.G, s.softening, nsteps=3)
    gp, gv, ga, _ = page.arrays()
    assert _err(gp, p) < 1e-12 and _err(gv, v) < 1e-12 and _err(ga, a) < 1e-12
    assert page.resolve_calls == [] and page.physics.contactPairs.size == 0     # idx:376-380: the map is emptied

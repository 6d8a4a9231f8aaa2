"""GPU tests of the renderer feed and snapshots through the C ABI:
device-side Body.history ring, ribbon-trail vertex builder (idx:563-592) against the
oracle, NBODYSN1 snapshot save/load against the host module."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

CAMERA = (0.0, 80.0, 150.0)   # camera.position.set(0, 80, 150), idx:705


def _radius(s):
    return (3 * s.pos_mass[:, 3] / (4 * np.pi)) ** (1 / 3)


def test_history_ring_and_ribbons_f64_bit_exact(nb, oracle):
    """Default scene, 200 frames of 2 sub-steps (animate(), idx:877-881): the device ring
    equals history.unshift/pop on the downloaded positions, wrap-around included, and the
    ribbon vertices equal the oracle's bit for bit (fp64 JS arithmetic, Float32Array store)."""
    s = nb.scenes.galaxy_default()
    rad = _radius(s)
    with nb.Engine(s.n, precision="f64") as e:
        e.upload(s.pos_mass, s.vel)
        e.set_radius(rad)
        with pytest.raises(nb.NBodyError):
            e.history_push()                                  # not enabled yet
        e.history_enable(80)
        with pytest.raises(nb.NBodyError):
            e.history()                                       # nothing pushed yet
        hist = None
        for frame in range(200):
            e.step(s.dt, s.G, s.softening, 2)
            e.history_push()
            pos = e.download(vel=False, acc=False)["pos_mass"]
            hist = oracle.history_push(oracle.history_init(pos, 80), pos) if hist is None else oracle.history_push(hist, pos)
            if frame in (0, 1, 79, 80, 81, 199):
                assert np.array_equal(e.history(), hist), frame
                assert np.array_equal(e.trail_ribbons(CAMERA), oracle.trail_ribbons(hist, rad, CAMERA)), frame
        # an upload with the same n keeps the ring; a different n refills it (idx:540)
        out = e.download()
        e.upload(out["pos_mass"], out["vel"], out["acc"])
        e.set_radius(rad)
        assert np.array_equal(e.history(), hist)
        e.upload(out["pos_mass"][:100], out["vel"][:100])
        with pytest.raises(nb.NBodyError):
            e.history()
        e.history_push()
        assert np.array_equal(e.history(), oracle.history_init(out["pos_mass"][:100], 80))


def test_ribbons_equal_reference_source_vectors(nb):
    """The ribbon kernel against vertices produced by the page's OWN VisualSystem.update (idx:563-592, executed
    by oracle/js_reference.py; tests/golden/make_reference_golden.py): the reference's histories are replayed
    into the device ring oldest first, then one nbody_trail_ribbons call must reproduce the page's Float32Array."""
    import os
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_js_page_galaxy_ribbons.npz"))
    hist, rad = g["history"], g["radius"]
    n, depth, _ = hist.shape
    pm = np.zeros((n, 4))
    pm[:, 3] = 1.0
    with nb.Engine(n, precision="f64") as e:
        e.upload(np.concatenate([hist[:, depth - 1], pm[:, 3:]], axis=1), np.zeros((n, 4)))
        e.set_radius(rad)
        e.history_enable(depth)
        for k in range(depth - 1, -1, -1):
            pm[:, :3] = hist[:, k]
            e.upload(pm, np.zeros((n, 4)))
            e.set_radius(rad)
            e.history_push()
        assert np.array_equal(e.history(), hist)
        assert np.array_equal(e.trail_ribbons(tuple(g["camera"])), g["ribbons"])


def test_ribbons_f32_close(nb, oracle):
    s = nb.scenes.two_galaxies(4096)
    pm, vel, _ = s.astype(np.float32)
    rad = _radius(s).astype(np.float32) + 0.3
    with nb.Engine(s.n, precision="f32") as e:
        e.upload(pm, vel)
        e.set_radius(rad)
        e.history_enable(16)
        for _ in range(20):
            e.step(s.dt, s.G, s.softening, 1)
            e.history_push()
        hist = e.history()
        rib = e.trail_ribbons(CAMERA)
        assert e.trail_ribbons(CAMERA, download=False) is None
    ref = oracle.trail_ribbons(hist.astype(np.float64), rad.astype(np.float64), CAMERA)
    # fp32 cross products of ~200-unit vectors with ~1e-2 steps: absolute error far below a ribbon width
    assert np.abs(rib - ref).max() < 2e-3 * float(rad.max())
    mid = 0.5 * (rib[:, :, :3] + rib[:, :, 3:])
    np.testing.assert_allclose(mid, hist, rtol=0, atol=1e-4)


@pytest.mark.parametrize("precision", ["f64", "f32"])
def test_snapshot_save_load(nb, oracle, tmp_path, precision):
    s = nb.scenes.galaxy_default()
    dt = np.float32 if precision == "f32" else np.float64
    pm, vel, temp = s.astype(dt)
    rad = _radius(s).astype(dt)
    p1, p2 = str(tmp_path / "a.nbs"), str(tmp_path / "b.nbs")
    with nb.Engine(s.n, precision=precision) as e:
        e.upload(pm, vel, None, temp)
        e.set_radius(rad)
        e.step(s.dt, s.G, s.softening, 3)
        want = e.download(temp=True)
        e.save_snapshot(p1, step=3, dt=s.dt, G=s.G, softening=s.softening)
    d = nb.snapshot.read(p1)
    assert (d["n"], d["step"], d["dt"], d["G"], d["softening"], d["dtype"]) == (s.n, 3, s.dt, s.G, s.softening, np.dtype(dt))
    for k in ("pos_mass", "vel", "acc", "temp"):
        assert np.array_equal(d[k], want[k]), k
    assert np.array_equal(d["radius"], rad)
    # a file written by the host module loads into a fresh context and continues identically
    nb.snapshot.write(p2, d["pos_mass"], d["vel"], d["acc"], d["temp"], d["radius"], step=3, dt=s.dt, G=s.G, softening=s.softening)
    assert open(p1, "rb").read() == open(p2, "rb").read()
    with nb.Engine(s.n + 9, precision=precision) as e:
        meta = e.load_snapshot(p2)
        assert meta == {"step": 3, "dt": s.dt, "G": s.G, "softening": s.softening} and e.n == s.n
        e.step(meta["dt"], meta["G"], meta["softening"], 2)
        got = e.download(temp=True)
        e.contacts(0.0)                                       # the radii came with the file
    with nb.Engine(s.n, precision=precision) as e:
        e.upload(pm, vel, None, temp)
        e.step(s.dt, s.G, s.softening, 3)
        e.sync()
        e.step(s.dt, s.G, s.softening, 2)
        straight = e.download(temp=True)
    for k in ("pos_mass", "vel", "acc", "temp"):
        assert np.array_equal(got[k], straight[k]), k        # checkpoint/resume is exact


def test_snapshot_errors(nb, tmp_path):
    from nbody_b200 import _lib
    s = nb.scenes.galaxy_default()
    p = str(tmp_path / "a.nbs")
    nb.snapshot.write(p, s.pos_mass, s.vel)
    with nb.Engine(s.n, precision="f32") as e:
        with pytest.raises(nb.NBodyError) as ei:
            e.load_snapshot(p)                                # float64 file, float32 context
        assert ei.value.status == _lib.NBODY_EINVAL
    with nb.Engine(s.n - 1, precision="f64") as e:
        with pytest.raises(nb.NBodyError):
            e.load_snapshot(p)                                # n > capacity
    with nb.Engine(s.n, precision="f64") as e:
        with pytest.raises(nb.NBodyError) as ei:
            e.load_snapshot(str(tmp_path / "missing.nbs"))
        assert ei.value.status == _lib.NBODY_EIO
        raw = open(p, "rb").read()
        open(p, "wb").write(raw[:-16])
        with pytest.raises(nb.NBodyError) as ei:
            e.load_snapshot(p)
        assert ei.value.status == _lib.NBODY_EIO
        with pytest.raises(nb.NBodyError):
            e.save_snapshot(p)                                # nothing uploaded


def test_visual_system_mirror(nb, oracle):
    """The page's frame loop through the mirrors (animate(), idx:873-883): physics.integrate x subSteps,
    then visuals.update -- histories, ribbon vertices and a rendered frame against the oracle."""
    ph = nb.physics
    ph.seed_random(77)
    ps = ph.PhysicsSystem(precision="f64")
    ph.resetScenario(ps, "galaxy")
    vis = nb.VisualSystem(ps)
    hist = None
    for frame in range(6):
        ps.integrate(ph.CONFIG.dt / ph.CONFIG.subSteps, nsteps=ph.CONFIG.subSteps)
        trails = vis.update(ps.bodies, ph.CONFIG.dt)
        pos = np.array([[*b.pos, b.mass] for b in ps.bodies])
        hist = oracle.history_push(oracle.history_init(pos, 80), pos) if hist is None else oracle.history_push(hist, pos)
    radius = np.array([b.radius for b in ps.bodies])
    assert np.array_equal(vis.histories(), hist)
    assert np.array_equal(trails, oracle.trail_ribbons(hist, radius, vis.camera_position))
    img, ids = vis.render(160, 90, ids=True)
    types = np.array([nb.visuals.TYPE_INDEX[b.type] for b in ps.bodies], dtype=np.int32)
    seeds = np.array([vis.seeds[b.id] for b in ps.bodies], dtype=np.float32)
    ref, ref_ids = oracle.render(pos, radius, [b.temperature for b in ps.bodies], types, seeds, 160, 90)
    assert np.array_equal(ids, ref_ids) and np.array_equal(img, ref)
    assert (ids == 0).any()                                   # the 1000-degree metal centre is in view (idx:750)
    vis.garbageCollect(ps.bodies[:10])
    assert len(vis.seeds) == 10
    ph.seed_random(None)


@pytest.mark.parametrize("precision", ["f64", "f32"])
def test_history_remap_follows_bodies(nb, oracle, precision):
    """A Body object carries its history through list changes (idx:383-389, 835); the device ring
    follows with nbody_history_remap: survivors keep their samples under their new index, new
    bodies start with depth copies of their position (idx:540), wrap-around included."""
    dt = np.float64 if precision == "f64" else np.float32
    rng = np.random.default_rng(5)
    n, depth = 300, 12
    pm = rng.normal(0, 10, (n, 4)).astype(dt)
    vel = np.zeros((n, 4), dt)
    vel[:, :3] = rng.normal(0, 1, (n, 3))
    with nb.Engine(512, precision=precision) as e:
        e.upload(pm, vel)
        with pytest.raises(nb.NBodyError):
            e.history_remap(np.arange(n))                     # no ring yet
        e.history_enable(depth)
        hist = None
        for frame in range(17):                               # past one wrap of the ring
            e.kick1_drift(0.01)
            e.history_push()
            pos = e.download(vel=False, acc=False)["pos_mass"]
            hist = oracle.history_push(oracle.history_init(pos, depth), pos) if hist is None else oracle.history_push(hist, pos)
        assert np.array_equal(e.history(), hist)
        # the list changes twice before the renderer looks again: 40 bodies leave, then 25 new ones join
        out = e.download()
        keep = np.sort(rng.choice(n, n - 40, replace=False))
        e.upload(out["pos_mass"][keep], out["vel"][keep])
        new_pm = rng.normal(0, 10, (25, 4)).astype(dt)
        pm2 = np.concatenate([new_pm[:10], out["pos_mass"][keep], new_pm[10:]])
        vel2 = np.concatenate([np.zeros((10, 4), dt), out["vel"][keep], np.zeros((15, 4), dt)])
        e.upload(pm2, vel2)
        source = np.concatenate([np.full(10, -1), keep, np.full(15, -1)])
        with pytest.raises(nb.NBodyError):
            e.history_remap(np.where(source < 0, n, source))  # outside the previous list
        with pytest.raises(ValueError):
            e.history_remap(source[:-1])
        e.history_remap(source)
        with pytest.raises(nb.NBodyError):
            e.history()                                       # new rows are filled by the next push
        e.kick1_drift(0.01)
        e.history_push()
        pos = e.download(vel=False, acc=False)["pos_mass"]
        carried = np.where(source[:, None, None] >= 0, hist[np.maximum(source, 0)], oracle.history_init(pos, depth))
        hist2 = oracle.history_push(carried, pos)
        assert np.array_equal(e.history(), hist2)
        # same length, different identity: a permutation
        perm = rng.permutation(len(source))
        e.upload(pos[perm], vel2[perm])
        e.history_remap(perm)
        e.history_push()
        assert np.array_equal(e.history(), oracle.history_push(hist2[perm], pos[perm]))
        # without a remap a changed n starts every ring over, as before
        e.upload(pos[:50], vel2[:50])
        e.history_push()
        assert np.array_equal(e.history(), oracle.history_init(pos[:50], depth))


def test_visual_system_keeps_trails_through_merges(nb, oracle):
    """The page's frame loop with its own collision stage: bodies merge and fracture between
    frames, and every surviving Body keeps its own history (keyed by body.id like meshMap /
    trailMap, idx:505-506, 540, 564-565)."""
    ph = nb.physics
    ph.seed_random(21)
    rng = np.random.default_rng(3)
    ps = ph.PhysicsSystem(precision="f64", collisions=True)
    for _ in range(90):
        ps.addBody(ph.Body(float(rng.uniform(0.5, 3.0)), ph.V3(*rng.uniform(-6, 6, 3)), ph.V3(*rng.normal(0, 0.4, 3)), "rock", 0.0))
    vis = nb.VisualSystem(ps)
    depth = ph.CONFIG.trailLength
    ref = {}                                                  # body.id -> history, the way Body objects hold it
    counts = set()
    for frame in range(100):
        ps.integrate(ph.CONFIG.dt / ph.CONFIG.subSteps, nsteps=ph.CONFIG.subSteps)
        vis.update(ps.bodies, ph.CONFIG.dt)
        for b in ps.bodies:
            p = np.array([[*b.pos]])
            h = ref.get(b.id)
            ref[b.id] = oracle.history_push(oracle.history_init(p, depth) if h is None else h, p)
        counts.add(len(ps.bodies))
        if frame % 9 == 0 or frame == 99:
            want = np.concatenate([ref[b.id] for b in ps.bodies])
            assert np.array_equal(vis.histories(), want), frame
    assert len(counts) > 3 and ps.stats["merge"] > 0          # the list really changed along the way
    radius = np.array([b.radius for b in ps.bodies])
    assert np.array_equal(vis.trails, oracle.trail_ribbons(want, radius, vis.camera_position))
    ph.seed_random(None)

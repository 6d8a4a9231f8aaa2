"""The oracle pinned against the reference's OWN source text.

No JS engine exists in the build image or on the GPU box, so oracle/js_reference.py
executes `/root/reference/index.html`'s `CONFIG`, `Body`, `PhysicsSystem` (idx:52-75,
208-444) and `resetScenario` (idx:744-764) through a JavaScript-subset transliterator.

* "live" tests run where the reference is present (the build container): they execute the
  page's statements and require the C oracle (nbody_oracle.c, collide_oracle.c) and the host
  mirror to agree BIT FOR BIT, and they re-derive the committed vectors;
* "fixture" tests run anywhere (the GPU box has no /root/reference): the committed
  tests/golden/ref_js_*.npz (made by tests/golden/make_reference_golden.py) against the oracle.
The CUDA path is held to the same vectors in tests/test_gpu_parity.py.
"""
import importlib.util
import os

import numpy as np
import pytest

from oracle import js_reference as jr

GOLD = os.path.join(os.path.dirname(__file__), "golden")
live = pytest.mark.skipif(not jr.available(), reason="/root/reference/index.html is not on this box")
STATE_KEYS = ("pos_mass", "vel", "acc", "temp")


def _make_golden():
    spec = importlib.util.spec_from_file_location("make_reference_golden", os.path.join(GOLD, "make_reference_golden.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _run(js, entry="f", *args):
    ns = {"JSArray": jr.JSArray, "JSObject": jr.JSObject, "JSGetter": jr.JSGetter, "_bind": jr._bind, "Map": jr.JSMap, "Set": jr.JSSet,
          "_div": jr._div, "Math": jr._Math(1)}
    exec(jr.transliterate(js), ns)
    return ns[entry](*args)


# ---------------------------------------------------------------- the JS subset itself

def test_subset_precedence_and_association():
    assert _run("function f(a, b, c) { return a + b * c - a / b; }", "f", 2.0, 4.0, 8.0) == 2.0 + 4.0 * 8.0 - 2.0 / 4.0
    assert _run("function f(a, b) { return -(1 + a) * b; }", "f", 0.2, 3.0) == (-(1 + 0.2)) * 3.0
    # left-to-right: (f * m) * v, never f * (m * v)  (idx:282-288)
    f, m, v = 0.1, 0.3, 0.7
    assert _run("function f(f_, m, v) { return f_ * m * v; }", "f", f, m, v) == (f * m) * v
    assert _run("function f(a, b) { return a > b ? a - b : b - a; }", "f", 1.0, 3.0) == 2.0
    assert _run("function f(a) { return a || 7; }", "f", 0) == 7 and _run("function f(a) { return !a && 1 === 1; }", "f", 0) is True


def test_subset_division_by_zero_is_ieee():
    assert _run("function f(a) { return a / 0; }", "f", 1.0) == float("inf")
    assert _run("function f(a) { return a / 0; }", "f", -1.0) == float("-inf")
    r = _run("function f(a) { return a / 0; }", "f", 0.0)
    assert r != r


def test_subset_loops_collections_and_arrows():
    js = """
    function f(n) {
        const seen = new Set(); const m = new Map(); let total = 0;
        for (let i = 0; i < n; i++) { if (i === 2) continue; seen.add(i); m.set(i, (m.get(i) || 0) + i * 2); total += i; }
        for (const k of m.keys()) { if (!seen.has(k + 1)) m.delete(k); }
        const xs = [5, 6, 7, 8].filter((_, idx) => !seen.has(idx));
        return [total, m.size, xs.length, xs[0], [3, 1].sort().join('-')];
    }"""
    total, msize, nx, x0, key = _run(js, "f", 4)
    assert (total, nx, x0, key) == (0 + 1 + 3, 1, 7, (1, 3))
    assert msize == 1                                   # keys 1, 3 lose their successor test; 0 keeps it


@pytest.mark.parametrize("js", [
    "function f() { while (true) {} }", "function f(a) { a = b = 1; return a; }", "function f(n) { for (let i = 0; i < n; i += 2) {} }",
    "function f(n) { for (let i = 0; i < n - 1; i++) { if (i) continue; } }", "function f(a) { a.x = function () { return this.y; }.bind(a); }", "function f() { g = 1; }", "function f(a) { return `t${a}`; }",
])
def test_subset_rejects_what_it_does_not_cover(js):
    with pytest.raises(SyntaxError):
        exec(jr.transliterate(js), {})


# ---------------------------------------------------------------- live: the page's statements vs the oracle

@live
def test_live_every_method_is_translated():
    r = jr.Reference()
    for cls, names in ((r.Body, ("__init__", "updateThermal")),
                       (r.PhysicsSystem, ("__init__", "addBody", "integrate", "computeGravity", "resolveCollisions", "mergeBodies", "fractureBodies"))):
        for n in names:
            assert callable(getattr(cls, n)), n
    c = r.CONFIG      # idx:53-67
    assert (c.G, c.softening, c.dt, c.subSteps, c.maxBodies, c.heatDecay, c.fractureThreshold, c.minFragmentMass, c.mergeTime) == \
        (0.5, 0.5, 0.016, 2, 300, 0.998, 25, 0.2, 0.5)
    assert [c.materials[t].density for t in jr.TYPES] == [1.0, 3.0, 0.5]


@live
@pytest.mark.parametrize("scene_name,n,nsteps", [("galaxy_default", 151, 6), ("plummer", 96, 6), ("two_galaxies", 64, 4)])
def test_live_gravity_only_step_bit_exact(nb, oracle, scene_name, n, nsteps):
    """integrate(dt) of the page (idx:247-262, idx:256 a no-op) == oracle_integrate, every step, every bit."""
    s = nb.scenes.galaxy_default() if scene_name == "galaxy_default" else getattr(nb.scenes, scene_name)(n)
    r = jr.Reference(collisions=False)
    r.CONFIG.G, r.CONFIG.softening = s.G, s.softening
    for i in range(s.n):
        r.add(s.pos_mass[i, 3], s.pos_mass[i, :3], s.vel[i, :3], "rock", s.temp[i])
    p, v, a, t = s.pos_mass, s.vel, np.zeros_like(s.vel), s.temp
    for _ in range(nsteps):
        r.integrate(s.dt)
        p, v, a, t = oracle.integrate(p, v, a, t, s.dt, s.G, s.softening, nsteps=1)
        st = r.state()
        for k, x in zip(STATE_KEYS, (p, v, a, t)):
            assert np.array_equal(st[k], x), k


@live
@pytest.mark.parametrize("speed,nsteps", [(0.3, 100), (8.0, 30)])
def test_live_complete_step_bit_exact(oracle, speed, nsteps):
    """The COMPLETE integrate (collisions, merges, fractures, new ids, Math.random draws) == oracle World."""
    bodies = _make_golden().crowd(80, 3, 5.5, speed)
    r = jr.Reference(seed=42, collisions=True)
    w = oracle.World(seed=42, max_bodies=300)
    for (m, p, v, ty, t) in bodies:
        r.add(m, p, v, ty, t)
        w.add(m, p, v, ty, t)
    for step in range(nsteps):
        r.integrate(0.008)
        w.integrate(0.008, 0.5, 0.5, True, 1)
        a, b = r.state(), w.state()
        for k in a:
            assert np.array_equal(a[k], b[k]), (step, k)
    c = w.counters()
    assert (c["flash"], c["explosion"]) == (r.visuals.flash, r.visuals.explosion)
    assert r.physics.contactPairs.size == w.n_contacts
    assert (c["merge"] if speed < 1 else c["fracture"]) > 0


_EDGE_CASES = {
    # name: (bodies (mass, pos, vel, type, temp), steps, softening, collisions, maxBodies)
    "empty": ([], 2, 0.5, True, 300),
    "single": ([(1.0, (0, 0, 0), (1, 0, 0), "rock", 60.0)], 3, 0.5, True, 300),
    "coincident": ([(1.0, (0, 0, 0), (0, 0, 0), "rock", 0.0), (2.0, (0, 0, 0), (0, 0, 0), "ice", 0.0)], 3, 0.5, True, 300),
    "coincident_zero_softening_nan": ([(1.0, (0, 0, 0), (0, 0, 0), "rock", 0.0), (2.0, (0, 0, 0), (0, 0, 0), "ice", 0.0),
                                       (1.0, (3, 0, 0), (0, 0, 0), "rock", 0.0)], 2, 0.0, False, 300),      # G / pow(0, 1.5) = Infinity, 0 * Infinity = NaN
    "massless_pair": ([(0.0, (0, 0, 0), (0.1, 0, 0), "rock", 0.0), (0.0, (0, 0, 0), (-0.1, 0, 0), "rock", 0.0),
                       (1.0, (5, 0, 0), (0, 0, 0), "metal", 0.0)], 3, 0.5, True, 300),                       # invMass = 0 (idx:212): x / 0 in the impulse
    "massless_on_massive": ([(0.0, (0, 0, 0), (1, 0, 0), "rock", 0.0), (5.0, (0.5, 0, 0), (-1, 0, 0), "metal", 10.0)], 5, 0.5, True, 300),
    "thermal_floor": ([(1.0, (0, 0, 0), (0, 0, 0), "rock", 0.11), (1.0, (9, 0, 0), (0, 0, 0), "rock", 0.1)], 60, 0.5, False, 300),  # T < 0.1 -> 0 (idx:229)
    "fracture": ([(3.0, (-0.4, 0, 0), (40, 0, 0), "rock", 0.0), (3.0, (0.4, 0, 0), (-40, 0, 0), "metal", 0.0)], 4, 0.5, True, 300),
    "fracture_below_min_fragment_mass": ([(0.15, (-0.1, 0, 0), (40, 0, 0), "rock", 0.0), (0.15, (0.1, 0, 0), (-40, 0, 0), "ice", 0.0)], 4, 0.5, True, 300),
    "pile_of_four_merging": ([(1.0, (0, 0, 0), (0, 0, 0), "rock", 0.0), (1.0, (0.3, 0, 0), (-1, 0, 0), "rock", 0.0),
                              (1.0, (0.1, 0.2, 0), (0, -1, 0), "ice", 0.0), (1.0, (-0.2, 0.1, 0.1), (1, 0, 0), "metal", 0.0)], 80, 0.5, True, 300),
    "fracture_into_a_full_list": ([(3.0, (-0.4, 0, 0), (40, 0, 0), "rock", 0.0), (3.0, (0.4, 0, 0), (-40, 0, 0), "metal", 0.0)] +
                                  [(1.0, (20.0 + 5 * k, 0, 0), (0, 0, 0), "ice", 0.0) for k in range(5)], 4, 0.5, True, 7),   # addBody shifts the oldest out (idx:240-242)
}


@live
@pytest.mark.parametrize("name", sorted(_EDGE_CASES))
def test_live_edge_cases_bit_exact(oracle, name):
    """Empty and single-body lists, coincident bodies, NaN/Infinity propagation (the page signals nothing),
    massless bodies, the thermal floor, fractures that fall below minFragmentMass, a full body list: page == oracle."""
    bodies, steps, eps, collisions, cap = _EDGE_CASES[name]
    r = jr.Reference(seed=3, collisions=collisions)
    r.CONFIG.softening, r.CONFIG.maxBodies = eps, cap
    w = oracle.World(seed=3, max_bodies=cap)
    for (m, p, v, ty, t) in bodies:
        r.add(m, p, v, ty, t)
        w.add(m, p, v, ty, t)
    sizes = {len(bodies)}
    for step in range(steps):
        r.integrate(0.008)
        w.integrate(0.008, 0.5, eps, collisions, 1)
        a, b = r.state(), w.state()
        sizes.add(len(a["temp"]))
        for k in a:
            assert np.array_equal(a[k], b[k], equal_nan=True), (name, step, k)
    assert r.Math.draws >= len(bodies) and r.physics.contactPairs.size == w.n_contacts
    if name == "coincident_zero_softening_nan":
        assert np.isnan(a["acc"]).any()
    if name in ("fracture", "pile_of_four_merging"):
        assert len(sizes) > 1
    if name == "fracture_into_a_full_list":   # two bodies out, the fragments in, the oldest bodies pushed out: full again
        assert sizes == {cap} and w.counters()["fracture"] == 1 and not np.array_equal(a["pos_mass"][0, :3], [20.0, 0, 0])


@live
@pytest.mark.parametrize("seed", range(8))
def test_live_random_small_worlds_bit_exact(oracle, seed):
    """Random little worlds (2-14 bodies in a box a few radii wide, speeds from gentle to shattering, random CONFIG
    thresholds): whatever mixture of bounces, merges, fractures and list overflows they produce, page == oracle."""
    rng = np.random.default_rng(1000 + seed)
    n, cap = int(rng.integers(2, 15)), int(rng.choice([6, 12, 300]))
    speed = float(rng.choice([0.2, 2.0, 12.0, 40.0]))
    r = jr.Reference(seed=seed, collisions=True)
    c = r.CONFIG
    c.maxBodies, c.fractureThreshold, c.minFragmentMass, c.mergeTime = cap, float(rng.choice([5.0, 25.0])), float(rng.choice([0.05, 0.2])), float(rng.choice([0.02, 0.5]))
    w = oracle.World(seed=seed, max_bodies=cap)
    w.config(c.fractureThreshold, c.minFragmentMass, c.mergeTime, c.heatDecay)
    for _ in range(n):
        m, p, v = float(rng.uniform(0.1, 4.0)), tuple(rng.uniform(-2.0, 2.0, 3)), tuple(rng.normal(0, speed, 3))
        ty, t = jr.TYPES[int(rng.integers(0, 3))], float(rng.uniform(0, 80))
        r.add(m, p, v, ty, t)
        w.add(m, p, v, ty, t)
    for step in range(60):
        r.integrate(0.008)
        w.integrate(0.008, c.G, c.softening, True, 1)
        a, b = r.state(), w.state()
        assert len(a["temp"]) == len(b["temp"]), step
        for k in a:
            assert np.array_equal(a[k], b[k], equal_nan=True), (step, k)
    assert r.physics.contactPairs.size == w.n_contacts and (r.visuals.flash, r.visuals.explosion) == (w.counters()["flash"], w.counters()["explosion"])


@live
def test_live_collision_scenario_bit_exact(oracle):
    """The page's second built-in scene, resetScenario('collision') (idx:760-763): two mass-100 bodies fall
    together and merge; 10,000 complete steps, page == oracle at every checkpoint."""
    r = jr.Reference(seed=8, collisions=True)
    r.resetScenario("collision")
    s0 = r.state()
    w = oracle.World(seed=8, max_bodies=300)
    for i in range(2):
        w.add(s0["pos_mass"][i, 3], s0["pos_mass"][i, :3], s0["vel"][i, :3], jr.TYPES[s0["type"][i]], s0["temp"][i])
    r.Math.state = 99
    w.seed(99)
    sizes = []
    for _ in range(40):
        r.integrate(0.008, 250)
        w.integrate(0.008, 0.5, 0.5, True, 250)
        a, b = r.state(), w.state()
        sizes.append(len(a["temp"]))
        assert len(a["temp"]) == len(b["temp"])
        for k in ("pos_mass", "vel", "acc", "temp", "radius", "type"):
            assert np.array_equal(a[k], b[k]), k
    assert sizes[0] == 2 and sizes[-1] == 1 and w.counters()["merge"] == 1
    assert a["pos_mass"][0, 3] == 200.0                       # mass conserved by mergeBodies (idx:393)


@live
def test_live_reset_scenario_equals_mirror(nb):
    """The page's scene generator (idx:744-759) and the host mirror's, same Math.random stream."""
    ph = nb.physics
    for kind in ("galaxy", "collision"):
        r = jr.Reference(seed=77)
        r.resetScenario(kind)
        ph.seed_random(77)
        ps = ph.PhysicsSystem(precision="f64", collisions=True)
        ph.resetScenario(ps, kind)
        ph.seed_random(None)
        st = r.state()
        assert len(ps.bodies) == len(st["temp"]) == (151 if kind == "galaxy" else 2)
        for i, b in enumerate(ps.bodies):
            assert (b.pos.x, b.pos.y, b.pos.z, b.mass) == tuple(st["pos_mass"][i])
            assert (b.vel.x, b.vel.y, b.vel.z) == tuple(st["vel"][i, :3])
            assert (b.radius, b.temperature, b.type) == (st["radius"][i], st["temp"][i], jr.TYPES[st["type"][i]])


@live
def test_live_max_bodies_shift():
    r = jr.Reference()
    r.CONFIG.maxBodies = 3
    for k in range(5):
        r.add(1.0 + k, (10.0 * k, 0, 0), (0, 0, 0))
    assert [b.mass for b in r.physics.bodies] == [3.0, 4.0, 5.0]      # idx:240-242 drops the oldest


@live
def test_live_committed_vectors_rederive(tmp_path, nb):
    """The committed ref_js fixtures are what the reference produces today (the small ones are regenerated)."""
    mg = _make_golden()
    mg.HERE = str(tmp_path)
    mg.gravity_fixture("ref_js_plummer256_gravity", nb.scenes.plummer(256), [1, 2, 10])
    mg.collision_fixture("ref_js_crowd_fracture", mg.crowd(120, 11, 7.0, 8.0), seed=99, nsteps=40)
    for name in ("ref_js_plummer256_gravity", "ref_js_crowd_fracture"):
        a, b = np.load(tmp_path / (name + ".npz")), np.load(os.path.join(GOLD, name + ".npz"))
        assert sorted(a.files) == sorted(b.files)
        for k in a.files:
            assert np.array_equal(a[k], b[k]), (name, k)


@live
def test_live_visual_update_ribbons_bit_exact(oracle):
    """VisualSystem.update (idx:500-597) executed from the page: the history unshift/pop and the ribbon vertex
    loop equal oracle/visual_oracle.c bit for bit, every frame, through merges (bodies appearing and leaving)."""
    r = jr.Reference(seed=5, collisions=True)
    r.resetScenario("galaxy")
    r.Math.state = 1234
    vis = jr.ReferenceVisuals(r)
    chains, sizes = {}, set()
    for frame in range(60):
        r.integrate(0.008, 2)
        vis.update(0.016)
        st, H = r.state(), vis.histories()
        sizes.add(len(st["temp"]))
        assert np.array_equal(vis.ribbons(), oracle.trail_ribbons(H, st["radius"], (0.0, 80.0, 150.0))), frame
        for i, ident in enumerate(st["id"]):           # Body.history against history_init/history_push, per identity
            pos = st["pos_mass"][i:i + 1]
            chains[ident] = oracle.history_push(chains[ident] if ident in chains else oracle.history_init(pos, 80), pos)
            assert np.array_equal(chains[ident][0], H[i]), (frame, i)
    assert len(sizes) > 1                               # the body list did change


def test_subset_closures_function_expressions_and_objects():
    # closures rebind enclosing variables (nonlocal), `this` of a function assigned to a member is that object,
    # object literals carry methods, getters and shorthand properties, missing members are undefined
    js = """
    function f(obj, xs) {
        let t = 0, calls = 0;
        function bump(k = 2) { calls++; t += k; }
        xs.forEach(x => { t += x; });
        bump(); bump(5);
        obj.twice = function twice(a) { return this.base * 2 + a; };
        const g = x => { return x + 1; };
        const shorthand = 7;
        return { t, calls, shorthand, plus: g(1), get total() { return t + calls; }, inc() { t++; }, missing: obj.nope || 'undefined' };
    }"""
    obj = jr.JSObject({"base": 10})
    r = _run(js, "f", obj, jr.JSArray([1, 2, 3]))
    assert (r.t, r.calls, r.shorthand, r.plus, r.missing) == (13, 2, 7, 2, "undefined")
    assert obj.twice(1) == 21 and r.total == 15
    r.inc()
    assert r.total == 16 and r.undefined_member is None


def test_subset_typed_arrays():
    F32 = jr.typed_array("float32")
    a = F32(8)
    a[1] = 0.1
    assert a.length == 8 and a.subarray(0, 4).length == 4 and isinstance(a[1], float) and a[1] == float(np.float32(0.1))
    a.subarray(4, 8)[0] = 2.0                             # a view, like TypedArray.prototype.subarray
    assert a[4] == 2.0


@live
def test_live_planet_shader_text_equals_restatement(oracle):
    """planetFragmentShader (GLSL, idx:99-202) executed from its text, one binary32 operation per GLSL operation
    (oracle/glsl_reference.py), against visual_oracle.c's planet_fragment / snoise: every bit of every channel."""
    from oracle import glsl_reference as gr
    sh, cfg = gr.PlanetShader(), jr.Reference().CONFIG
    f = _make_golden().fragments(200, 11)
    for k in range(200):
        m = cfg.materials[jr.TYPES[f["type"][k]]]
        a = sh.fragment(f["vpos"][k], f["vnormal"][k], f["vview"][k], m.color1.rgb, m.color2.rgb, f["temperature"][k], f["seed"][k])
        b = oracle.planet_fragment(f["vpos"][k], f["vnormal"][k], f["vview"][k], f["type"][k], f["temperature"][k], f["seed"][k])
        assert np.array_equal(a, b), (k, a, b)
        assert sh.snoise(f["vpos"][k] * np.float32(2.0)) == oracle.snoise(f["vpos"][k] * np.float32(2.0))
    # the shader declares the six uniforms and four varyings the renderer feeds (idx:100-110)
    assert all(("uniform %s;" % u) in " ".join(sh.glsl.split()) for u in ("vec3 color1", "vec3 color2", "float temperature", "float seed", "vec3 sunPosition"))


@live
def test_live_page_shader_pipeline_matches_oracle_render(oracle):
    """planetVertexShader (idx:83-97) + planetFragmentShader executed from their text with three.js's matrices
    (modelMatrix = T*S, viewMatrix = lookAt inverse, normalMatrix = inverse transpose of modelView) against the
    pixels of oracle.render, which ray-casts the sphere instead of rasterising it.  The varyings then differ in
    their last bits and the noise amplifies that, so the gate is statistical; the control shows what a wrong
    convention costs (the page's vNormal is a VIEW-space normal lit by a WORLD-space light: the oracle keeps that)."""
    from oracle import glsl_reference as gr
    vs, fs, cfg = gr.PlanetVertexShader(), gr.PlanetShader(), jr.Reference().CONFIG
    W, H, fov = 64, 36, 45.0
    cam = np.array([0.0, 80.0, 150.0])                         # camera.position, idx:705; looking at the origin
    z = cam / np.linalg.norm(cam)
    x = np.cross([0.0, 1.0, 0.0], z)
    x /= np.linalg.norm(x)
    y = np.cross(z, x)
    Rv = np.stack([x, y, z])
    view = np.eye(4)
    view[:3, :3], view[:3, 3] = Rv, -Rv @ cam
    th, asp = np.tan(np.radians(fov) / 2), W / H
    rng = np.random.default_rng(0)
    good, control = [], []
    for scene in range(6):
        c, r = np.float32(rng.uniform(-15, 15, 3)).astype(np.float64), float(np.float32(rng.uniform(4, 12)))
        ty, temp, seed = int(rng.integers(0, 3)), float(rng.choice([0, 30, 80])), float(np.float32(rng.uniform(0, 100)))
        rgba, ids = oracle.render(np.array([[*c, 1.0]]), [r], [temp], [ty], [seed], W, H, camera=tuple(cam), target=(0.0, 0.0, 0.0), fov_deg=fov)
        model = np.eye(4)
        model[:3, :3] *= r
        model[:3, 3] = c
        ys, xs = np.nonzero(ids == 0)
        m = cfg.materials[jr.TYPES[ty]]
        for k in rng.choice(len(ys), size=min(25, len(ys)), replace=False):
            px, py = xs[k], ys[k]
            d = x * (((2 * (px + 0.5)) / W - 1) * th * asp) + y * ((1 - (2 * (py + 0.5)) / H) * th) - z
            oc = cam - c
            a, b, cc = d @ d, oc @ d, oc @ oc - r * r
            n = (cam + ((-b - np.sqrt(b * b - a * cc)) / a) * d - c) / r            # the unit-sphere vertex under this pixel
            for nmat, sink in ((Rv / r, good), (np.eye(3) / r, control)):
                var = vs.vertex(n, n, model, view, np.eye(4), nmat)
                col = fs.fragment(var["vPosition"], var["vNormal"], var["vViewPosition"], m.color1.rgb, m.color2.rgb, temp, seed)
                sink.append(float(np.abs(col - rgba[py, px]).max()))
    good, control = np.array(good), np.array(control)
    assert len(good) >= 60 and np.median(good) < 2e-4 and np.percentile(good, 90) < 5e-3
    assert np.median(control) > 100 * np.median(good)


def test_glsl_subset_rejects_control_flow():
    from oracle import glsl_reference as gr
    assert "def f(x):" in gr.transliterate("float f(float x) { float y = x * 2.0; y *= x; return y - 1.0; }")
    for bad in ("float f(float x) { if (x > 0.0) return x; return 0.0; }", "float f(float x) { for (int i = 0; i < 3; i++) x += 1.0; return x; }",
                "vec3 f(vec3 x) { x.xy = x.yx; return x; }", "float f(float x) { return 1.0; }\nfloat f(float x) { return 2.0; }"):
        with pytest.raises(SyntaxError):
            gr.transliterate(bad)


def test_subset_block_arrows_and_computed_bounds():
    js = """
    function f(xs) {
        const out = [];
        xs.forEach(x => { const y = x * 2; if (y > 2) { out.push(y); } });
        for (let i = 0; i < out.length - 1; i++) { out[i] = out[i] + out[i + 1]; }
        return out;
    }"""
    assert list(_run(js, "f", jr.JSArray([1, 2, 3, 4]))) == [10, 14, 8]


# ---------------------------------------------------------------- fixtures: run anywhere

def test_oracle_equals_reference_vectors_ribbons(oracle):
    g = np.load(os.path.join(GOLD, "ref_js_page_galaxy_ribbons.npz"))
    rib = oracle.trail_ribbons(g["history"], g["radius"], g["camera"], trail_length=int(g["trail_length"]))
    assert rib.dtype == np.float32 and np.array_equal(rib, g["ribbons"])

@pytest.mark.parametrize("name", ["ref_js_galaxy151_gravity", "ref_js_plummer256_gravity"])
def test_oracle_equals_reference_vectors_gravity(oracle, name):
    g = np.load(os.path.join(GOLD, name + ".npz"))
    p, v, a, t = g["pos_mass0"], g["vel0"], np.zeros_like(g["vel0"]), g["temp0"]
    done = 0
    for k, c in enumerate(g["checkpoints"]):
        p, v, a, t = oracle.integrate(p, v, a, t, float(g["dt"]), float(g["G"]), float(g["softening"]), nsteps=int(c) - done)
        done = int(c)
        for key, x in zip(STATE_KEYS, (p, v, a, t)):
            assert np.array_equal(x, g[key][k]), (name, int(c), key)


def test_oracle_equals_shader_text_vectors(oracle):
    g = np.load(os.path.join(GOLD, "ref_glsl_planet_fragments.npz"))
    for k in range(len(g["seed"])):
        b = oracle.planet_fragment(g["vpos"][k], g["vnormal"][k], g["vview"][k], g["type"][k], g["temperature"][k], g["seed"][k])
        assert np.array_equal(b, g["rgba"][k]), k
        assert oracle.snoise(g["vpos"][k]) == g["snoise"][k], k
    assert np.isfinite(g["rgba"]).all() and (g["rgba"][:, 3] == 1).all() and g["rgba"][:, :3].std() > 0.05


def _world_from_fixture(oracle, g):
    w = oracle.World(seed=int(g["seed"]), max_bodies=300)
    for i in range(len(g["mass0"])):
        w.add(float(g["mass0"][i]), g["pos0"][i], g["vel0"][i], jr.TYPES[int(g["type0"][i])], float(g["temp0"][i]))
    if int(g["reseed"]) >= 0:
        w.seed(int(g["reseed"]))
    return w


@pytest.mark.parametrize("name", ["ref_js_crowd_merge", "ref_js_crowd_fracture", "ref_js_page_galaxy_collisions"])
def test_oracle_world_equals_reference_vectors_collisions(oracle, name):
    g = np.load(os.path.join(GOLD, name + ".npz"))
    w = _world_from_fixture(oracle, g)
    w.integrate(float(g["dt"]), float(g["G"]), float(g["softening"]), True, int(g["nsteps"]))
    st = w.state()
    keys = STATE_KEYS + ("radius", "type") + (("id",) if int(g["reseed"]) < 0 else ())   # ids of a re-seeded run: identities only
    for k in keys:
        assert np.array_equal(st[k], g[k]), (name, k)
    c = w.counters()
    assert (c["flash"], c["explosion"], w.n_contacts) == (int(g["flashes"]), int(g["explosions"]), int(g["n_contacts"]))


@pytest.mark.parametrize("name", ["ref_js_crowd_merge", "ref_js_crowd_fracture", "ref_js_page_galaxy_collisions"])
def test_host_mirror_equals_reference_vectors_collisions(nb, oracle, name):
    """nbody_b200.physics.PhysicsSystem (candidate pairs + sequential response) against the reference's vectors,
    the device stood in for by oracle functions (as in tests/test_collisions_cpu.py)."""
    from .test_collisions_cpu import OracleDevice, _state_equal
    g = np.load(os.path.join(GOLD, name + ".npz"))
    ph = nb.physics
    ph.seed_random(int(g["seed"]))
    ps = ph.PhysicsSystem(precision="f64", collisions=True, contact_margin=0.5)
    ps._engine_obj = OracleDevice(oracle)
    for i in range(len(g["mass0"])):
        ps.addBody(ph.Body(float(g["mass0"][i]), ph.V3(*g["pos0"][i]), ph.V3(*g["vel0"][i]), jr.TYPES[int(g["type0"][i])], float(g["temp0"][i])))
    if int(g["reseed"]) >= 0:
        ph.seed_random(int(g["reseed"]))
    for _ in range(int(g["nsteps"])):
        ps.integrate(float(g["dt"]))
    ph.seed_random(None)
    _state_equal(ps, {k: g[k] for k in STATE_KEYS + ("radius",)})
    assert len(ps.contactPairs) == int(g["n_contacts"])

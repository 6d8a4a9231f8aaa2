"""Generates tests/golden/ref_js_*.npz by executing the REFERENCE'S OWN source text.

Runs only where /root/reference/index.html exists (the build container).  The page's
`CONFIG`, `Body`, `PhysicsSystem` and `resetScenario` are sliced out of the page at run
time and executed through oracle/js_reference.py (a JavaScript-subset transliterator; no
JS engine exists here).  Unlike make_golden.py's vectors, these are outputs of the
reference's statements, so they pin the C oracle (and through it the CUDA path) against
the reference.  The fixtures travel to the GPU box; the reference does not.

    python tests/golden/make_reference_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import nbody_b200 as nb  # noqa: E402
from oracle import js_reference as jr  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
STATE_KEYS = ("pos_mass", "vel", "acc", "temp")


def crowd(n, seed, box, speed):
    """Same recipe as tests/test_collisions_cpu.py::_crowd (numpy Generator, fixed seed)."""
    rng = np.random.default_rng(seed)
    return [(float(rng.uniform(0.3, 3.0)), tuple(rng.uniform(-box, box, 3)), tuple(rng.normal(0, speed, 3)),
             jr.TYPES[int(rng.integers(0, 3))], float(rng.uniform(0, 5))) for _ in range(n)]


def gravity_fixture(name, scene, checkpoints):
    """Gravity-only parity runs: idx:256 is a no-op, everything else is the page's integrate."""
    r = jr.Reference(collisions=False)
    r.CONFIG.G, r.CONFIG.softening = scene.G, scene.softening
    for i in range(scene.n):
        r.add(scene.pos_mass[i, 3], scene.pos_mass[i, :3], scene.vel[i, :3], "rock", scene.temp[i])
    out = {k: [] for k in STATE_KEYS}
    done = 0
    for c in checkpoints:
        r.integrate(scene.dt, c - done)
        done = c
        st = r.state()
        for k in STATE_KEYS:
            out[k].append(st[k])
    np.savez_compressed(os.path.join(HERE, name + ".npz"), pos_mass0=scene.pos_mass, vel0=scene.vel, temp0=scene.temp,
                        dt=scene.dt, G=scene.G, softening=scene.softening, checkpoints=np.array(checkpoints),
                        **{k: np.stack(v) for k, v in out.items()})
    print(name, scene.n, checkpoints)


def bodies_arrays(bodies):
    return dict(mass0=np.array([b[0] for b in bodies]), pos0=np.array([b[1] for b in bodies]),
                vel0=np.array([b[2] for b in bodies]), type0=np.array([jr.TYPES.index(b[3]) for b in bodies], dtype=np.int32),
                temp0=np.array([b[4] for b in bodies]))


def collision_fixture(name, bodies, seed, nsteps, dt=0.008, reseed=None, reference=None):
    """The COMPLETE integrate(dt) (idx:247-262 with resolveCollisions, merge, fracture)."""
    r = reference or jr.Reference(seed=seed, collisions=True)
    if reference is None:
        for (m, p, v, ty, t) in bodies:
            r.add(m, p, v, ty, t)
    if reseed is not None:
        r.Math.state = reseed
    draws0 = r.Math.draws
    r.integrate(dt, nsteps)
    st = r.state()
    np.savez_compressed(os.path.join(HERE, name + ".npz"), seed=seed, reseed=-1 if reseed is None else reseed, nsteps=nsteps, dt=dt,
                        G=r.CONFIG.G, softening=r.CONFIG.softening, n_contacts=r.physics.contactPairs.size,
                        flashes=r.visuals.flash, explosions=r.visuals.explosion, draws=r.Math.draws - draws0,
                        **bodies_arrays(bodies), **st)
    print(name, len(bodies), "->", len(r.physics.bodies), "bodies;", r.visuals.flash, "flashes,", r.visuals.explosion, "fragments")


def ribbons_fixture(name, frames, keep_every=5):
    """animate() (idx:873-883): subSteps x integrate(dt/subSteps) then visuals.update(bodies, dt), complete step
    with collisions, the page's own default scene.  Stores every `keep_every`-th body's history and ribbon
    vertices (the Float32Array the page hands to WebGL, idx:583-589) after the last frame."""
    r = jr.Reference(seed=5, collisions=True)
    r.resetScenario("galaxy")
    r.Math.state = 1234
    vis = jr.ReferenceVisuals(r)
    c = r.CONFIG
    for _ in range(frames):
        r.integrate(c.dt / c.subSteps, c.subSteps)
        vis.update(c.dt)
    st = r.state()
    keep = np.arange(0, len(st["temp"]), keep_every)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), frames=frames, camera=np.array([0.0, 80.0, 150.0]),
                        n_bodies=len(st["temp"]), kept=keep, radius=st["radius"][keep], history=vis.histories()[keep],
                        ribbons=vis.ribbons()[keep], trail_length=c.trailLength)
    print(name, frames, "frames,", len(st["temp"]), "bodies,", len(keep), "kept")


def fragments(count, seed):
    """Varyings and uniforms of `count` fragments on random spheres seen from the page's camera (0, 80, 150), idx:705."""
    rng = np.random.default_rng(seed)
    out = dict(vpos=[], vnormal=[], vview=[], type=[], temperature=[], seed=[])
    for _ in range(count):
        nrm = rng.normal(0, 1, 3)
        nrm /= np.linalg.norm(nrm)
        vpos = (rng.uniform(-80, 80, 3) + rng.uniform(0.4, 4.0) * nrm).astype(np.float32)
        out["vpos"].append(vpos)
        out["vnormal"].append((nrm * rng.uniform(0.5, 1.5)).astype(np.float32))        # un-normalised, as interpolation leaves it
        out["vview"].append((np.array([0.0, 80.0, 150.0]) - vpos).astype(np.float32))
        out["type"].append(int(rng.integers(0, 3)))
        out["temperature"].append(np.float32(rng.choice([0.0, 10.0, 30.0, 49.0, 60.0, 120.0, 1000.0])))
        out["seed"].append(np.float32(rng.uniform(0, 100)))                             # Math.random() * 100, idx:492
    return {k: np.array(v) for k, v in out.items()}


def shader_fixture(name, count=256):
    """gl_FragColor of the page's planetFragmentShader (idx:99-202), executed from its GLSL text (oracle/glsl_reference.py)."""
    from oracle import glsl_reference as gr
    sh, cfg = gr.PlanetShader(), jr.Reference().CONFIG
    f = fragments(count, 7)
    rgba = np.stack([sh.fragment(f["vpos"][k], f["vnormal"][k], f["vview"][k], cfg.materials[jr.TYPES[f["type"][k]]].color1.rgb,
                                 cfg.materials[jr.TYPES[f["type"][k]]].color2.rgb, f["temperature"][k], f["seed"][k]) for k in range(count)])
    noise = np.array([sh.snoise(f["vpos"][k]) for k in range(count)], dtype=np.float32)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), rgba=rgba, snoise=noise, **f)
    print(name, count, "fragments")


if __name__ == "__main__":
    if not jr.available():
        sys.exit("needs /root/reference/index.html")
    gravity_fixture("ref_js_galaxy151_gravity", nb.scenes.galaxy_default(), [1, 2, 100, 1000])   # BASELINE configs[0]
    gravity_fixture("ref_js_plummer256_gravity", nb.scenes.plummer(256), [1, 2, 10])             # configs[1] recipe, reduced n
    collision_fixture("ref_js_crowd_merge", crowd(120, 11, 7.0, 0.3), seed=99, nsteps=120)
    collision_fixture("ref_js_crowd_fracture", crowd(120, 11, 7.0, 8.0), seed=99, nsteps=40)
    # the page's own default scene from its own generator (idx:744-759), then 200 complete steps (100 frames)
    r = jr.Reference(seed=5, collisions=True)
    r.resetScenario("galaxy")
    s0 = r.state()
    page = [(s0["pos_mass"][i, 3], tuple(s0["pos_mass"][i, :3]), tuple(s0["vel"][i, :3]), jr.TYPES[s0["type"][i]], s0["temp"][i])
            for i in range(len(s0["temp"]))]
    collision_fixture("ref_js_page_galaxy_collisions", page, seed=5, nsteps=200, reseed=1234, reference=r)
    ribbons_fixture("ref_js_page_galaxy_ribbons", frames=100)
    shader_fixture("ref_glsl_planet_fragments")

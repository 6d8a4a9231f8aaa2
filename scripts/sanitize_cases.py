#!/usr/bin/env python
"""Small-N runs of every kernel family, for compute-sanitizer (scripts/gpu_suite.sh sanitize <tool>): the one-cluster kernel
(DSMEM + cluster barriers), the all-SM cooperative kernel (grid barrier), the launch-per-step sweeps (TMA ring + mbarriers, last-CTA
partial-sum pass), the fp64 and fp32 pairs-once passes (lane rotation, persistent scheduler, 64-bit reductions), contacts, the
history ring + ribbons and the renderer.  Every case is checked against the oracle, so a clean sanitizer log describes runs that
also computed the right thing."""
import hashlib
import os
import sys

import numpy as np

REPEAT = int(os.environ.get("CASES_REPEAT", "3"))   # every case this many times: the results must hash the same (no data race)

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nbody_b200 as nb  # noqa: E402
from oracle import oracle  # noqa: E402


def err(a, b):
    d = np.linalg.norm(np.asarray(a, np.float64)[:, :3] - b[:, :3], axis=1).max()
    return d / max(np.linalg.norm(b[:, :3], axis=1).max(), 1e-300)


def step_case(name, n, precision, env, nsteps=2, general=False, scene="plummer"):
    old = {k: os.environ.get(k) for k in env}
    os.environ.update({k: str(v) for k, v in env.items()})
    try:
        s = nb.scenes.SCENES[scene](n)
        dt = np.float32 if precision == "f32" else np.float64
        pm, vel, temp = s.astype(dt)
        if general:
            pm[:, 3] *= np.random.default_rng(n).uniform(0.5, 2.0, s.n).astype(dt)
        p, v, a, t = oracle.integrate(pm.astype(np.float64), vel.astype(np.float64), np.zeros((s.n, 4)), temp.astype(np.float64), s.dt, s.G,
                                      s.softening, nsteps=nsteps)
        hashes = set()
        for _ in range(REPEAT):
            with nb.Engine(s.n, precision=precision) as e:
                e.upload(pm, vel, None, temp)
                e.step(s.dt, s.G, s.softening, nsteps)
                o = e.download(temp=True)
                info = e.info()
            hashes.add(hashlib.blake2b(b"".join(np.ascontiguousarray(o[k]).tobytes() for k in ("pos_mass", "vel", "acc", "temp")), digest_size=8).hexdigest())
        assert len(hashes) == 1, (name, "results differ between repetitions", hashes)
        tol = 1e-11 if precision == "f64" else 5e-5
        worst = max(err(o["pos_mass"], p), err(o["vel"], v), err(o["acc"], a))
        print(f"{name:44s} n={s.n:6d} {precision} err {worst:.2e} x{REPEAT} {next(iter(hashes))}  sym32={info['f32_sym_last_pass']} warp={info['f32_sym_warp_units']} sym64={info['f64_symmetric']} uniform={info['uniform_mass']}", flush=True)
        assert worst < tol, (name, worst)
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


def main():
    per_step = {"NBODY_SMALL": 0, "NBODY_MEDIUM": 0}
    for precision in ("f64", "f32"):
        step_case("one cluster (small.cuh)", 151, precision, {}, scene="galaxy_default")
        step_case("all SMs, cooperative (medium.cuh)", 700, precision, {})
        step_case("all SMs, cooperative (medium.cuh)", 3001, precision, {}, general=True)
        step_case("launch per step, one-sided sweep", 151, precision, per_step, scene="galaxy_default")
        step_case("launch per step, one-sided sweep", 5000, precision, dict(per_step, NBODY_F32_SYM=0, NBODY_F64_SYM=0), general=True)
        step_case("launch per step, one-sided equal-mass sweep", 3001, precision, dict(per_step, NBODY_F32_SYM=0, NBODY_F64_SYM=0))
    step_case("fp64 pairs-once (force_f64_sym.cuh)", 5000, "f64", per_step, general=True)
    for R in (4, 8):
        for general in (False, True):
            step_case(f"fp32 pairs-once R={R} (force_f32_sym.cuh)", 5000, "f32", dict(per_step, NBODY_F32_SYM_MIN=1, NBODY_F32_SYM_R=R, NBODY_F32_SYMW=0), nsteps=3, general=general)
    step_case("fp32 pairs-once, outliers", 3002, "f32", dict(per_step, NBODY_F32_SYM_MIN=1, NBODY_F32_SYMW=0), scene="two_galaxies")
    wide = dict(per_step, NBODY_F32_SYMW_MIN=1, NBODY_F32_SYMW_MAX=1 << 30, NBODY_SYM_COOP=0)
    for R in (2, 4, 8):
        for general in (False, True):
            step_case(f"fp32 pairs-once, warp units R={R} (force_f32_symw.cuh)", 5000, "f32", dict(wide, NBODY_F32_SYMW_R=R), nsteps=3, general=general)
    step_case("fp32 pairs-once, warp units, default shape", 9001, "f32", dict(per_step, NBODY_SYM_COOP=0), nsteps=3, general=True)
    for R in (2, 4, 8):
        for general in (False, True):
            step_case(f"one cooperative launch, warp units R={R} (sym_coop.cuh)", 5000, "f32", dict(wide, NBODY_SYM_COOP=1, NBODY_F32_SYMW_R=R), nsteps=3, general=general)
    step_case("one cooperative launch, default shape", 9001, "f32", per_step, nsteps=4, general=True)

    # the rows next to the hot path: contacts (flagged and full scan), history ring + ribbons, renderer
    s = nb.scenes.galaxy_default()
    radius = 1.5 * (3 * s.pos_mass[:, 3] / (4 * np.pi)) ** (1 / 3)
    for precision in ("f64", "f32"):
        dt = np.float32 if precision == "f32" else np.float64
        pm, vel, temp = s.astype(dt)
        with nb.Engine(s.n, precision=precision) as e:
            e.upload(pm, vel, None, temp)
            e.set_radius(radius.astype(dt))
            e.force(s.G, s.softening)
            pairs = e.contacts(0.0)
            want = oracle.overlap_pairs(pm.astype(np.float64), radius.astype(dt).astype(np.float64), 0.0, mt=False)
            assert np.array_equal(pairs, want) and len(pairs) > 0
            e.history_enable(8)
            for _ in range(3):
                e.step(s.dt, s.G, s.softening, 2)
                e.history_push()
            e.trail_ribbons((0, 80, 150))
            e.render(64, 36, ids=True)
        print(f"contacts / history / ribbons / render                 n={s.n:6d} {precision} {len(pairs)} pairs", flush=True)
    # the page's complete step on the device (page_world.cu): collisions, merges and fractures, bit-identical to the oracle at G = 0
    rng = np.random.default_rng(3)
    w = oracle.World(seed=0xC0FFEE, max_bodies=300)
    for _ in range(70):
        w.add(float(rng.uniform(0.3, 3.0)), rng.uniform(-5, 5, 3), rng.normal(0, 8.0, 3), ("rock", "metal", "ice")[int(rng.integers(0, 3))], float(rng.uniform(0, 5)))
    st = w.state()
    with nb.Engine(1024, precision="f64") as e:
        e.upload(st["pos_mass"], st["vel"], st["acc"], st["temp"])
        e.world_enable(seed=w.rng, max_bodies=300)
        e.world_set_bodies(st["id"], st["type"], st["radius"])
        for _ in range(10):
            w.integrate(0.008, 0.0, 0.5, collisions=True, nsteps=3)
            e.world_step(0.008, 0.0, 0.5, 3, collisions=True)
        info = e.world_sync()
        o = e.download(temp=True)
    want = w.state()
    assert info["n"] == len(w) and np.array_equal(o["pos_mass"], want["pos_mass"]) and np.array_equal(o["vel"], want["vel"]) and np.array_equal(info["id"], want["id"])
    print(f"complete step with collisions on the device          n={info['n']:6d} f64 {info['counters']}", flush=True)
    print("sanitize cases ok")


if __name__ == "__main__":
    main()

"""Times nbody_world_step with and without the collision stage (default scene, 2 and 20 steps per launch); builder's lab."""
import sys, time, numpy as np
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nbody_b200 as nb
from oracle import oracle
oracle.build()
s = nb.scenes.galaxy_default()
TYPES = ("rock", "metal", "ice")
w = oracle.World(seed=0x5EED0001, max_bodies=300)
rng = np.random.default_rng(1)
for i in range(s.n):
    w.add(s.pos_mass[i, 3], s.pos_mass[i, :3], s.vel[i, :3], "metal" if i == 0 else TYPES[int(rng.integers(0, 3))], s.temp[i])
st = w.state()
for coll in (0, 1):
    for nsteps in (2, 20):
        e = nb.Engine(1024, precision="f64")
        e.upload(st["pos_mass"], st["vel"], st["acc"], st["temp"])
        e.world_enable(seed=w.rng, max_bodies=300)
        e.world_set_bodies(st["id"], st["type"], st["radius"])
        e.world_step(s.dt, s.G, s.softening, nsteps, collisions=bool(coll)); e.world_sync()
        t0 = time.perf_counter()
        reps = 200
        for _ in range(reps):
            e.world_step(s.dt, s.G, s.softening, nsteps, collisions=bool(coll))
        info = e.world_sync()
        us = (time.perf_counter() - t0) / reps * 1e6
        print(f"collisions={coll} nsteps={nsteps}: {us:.1f} us per launch, {us/nsteps:.2f} us per step, n={info['n']} {info['counters']}")
        e.close()

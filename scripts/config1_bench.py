#!/usr/bin/env python
"""BASELINE configs[0]: the page's default scene (N=151, G=0.5, eps=0.5), 1000 x integrate(0.008) = 500 display frames
of animate() (idx:873-883), gravity only.  Three timings, fp64:
  cpu      the C restatement of the JS step, one thread (the reference is single-threaded; Node/V8 unavailable)
  resident 500 x nbody_step(dt, G, eps, 2) with the state left on the device, one sync at the end
  per_frame the JS shim's flow: every frame upload(host arrays) -> step(.., 2) -> download(host arrays)
  step_host the same frame as ONE C-ABI call (nbody_step_host: arrays in, 2 x integrate, arrays out, one synchronise)
and the final states compared (1000 chaotic steps: 1e-9 norm-relative, tests/test_gpu_parity.py)."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nbody_b200 as nb  # noqa: E402
from oracle import oracle  # noqa: E402  (the timed CPU baseline leg and the checker)

s = nb.scenes.galaxy_default()
FRAMES, SUB = 500, 2

t0 = time.perf_counter()
p, v, a, t = oracle.integrate(s.pos_mass, s.vel, np.zeros_like(s.vel), s.temp, s.dt, s.G, s.softening, nsteps=FRAMES * SUB)
cpu_s = time.perf_counter() - t0

out = {}
for prec in ("f64", "f32"):
    dt = np.float64 if prec == "f64" else np.float32
    pm0, vel0, temp0 = s.astype(dt)
    with nb.Engine(s.n, precision=prec) as e:
        e.upload(pm0, vel0, None, temp0)
        e.step(s.dt, s.G, s.softening, SUB)
        e.sync()                                            # warm-up
        best_res, best_frame = 1e9, 1e9
        for _ in range(3):
            e.upload(pm0, vel0, None, temp0)
            e.sync()
            t0 = time.perf_counter()
            for _ in range(FRAMES):
                e.step(s.dt, s.G, s.softening, SUB)
            e.sync()
            best_res = min(best_res, time.perf_counter() - t0)
        res_state = e.download(temp=True)
        for _ in range(3):
            pm, vel, acc, temp = pm0.copy(), vel0.copy(), np.zeros_like(vel0), temp0.copy()
            t0 = time.perf_counter()
            for _ in range(FRAMES):
                e.upload(pm, vel, acc, temp)
                e.step(s.dt, s.G, s.softening, SUB)
                d = e.download(temp=True)
                pm, vel, acc, temp = d["pos_mass"], d["vel"], d["acc"], d["temp"]
            best_frame = min(best_frame, time.perf_counter() - t0)
        best_host = 1e9
        for _ in range(3):
            hp, hv, ha, ht = pm0.copy(), vel0.copy(), np.zeros_like(vel0), temp0.copy()
            t0 = time.perf_counter()
            for _ in range(FRAMES):
                e.step_host(hp, hv, ha, ht, s.dt, s.G, s.softening, SUB)
            best_host = min(best_host, time.perf_counter() - t0)
        err = lambda x, r: float(np.linalg.norm(x[:, :3].astype(np.float64) - r[:, :3], axis=1).max() / np.linalg.norm(r[:, :3], axis=1).max())
        out[prec] = {"resident_us_per_step": best_res / (FRAMES * SUB) * 1e6, "per_frame_us": best_frame / FRAMES * 1e6,
                     "per_frame_us_per_step": best_frame / (FRAMES * SUB) * 1e6, "step_host_frame_us": best_host / FRAMES * 1e6,
                     "step_host_equals_resident": bool(np.array_equal(hp, res_state["pos_mass"])),
                     "pos_err_resident_vs_cpu": err(res_state["pos_mass"], p), "pos_err_per_frame_vs_cpu": err(pm, p),
                     "per_frame_equals_resident": bool(np.array_equal(pm, res_state["pos_mass"]))}
# the COMPLETE reference step (resolveCollisions, merges, fractures included) on the device: csrc/page_world.cu, fp64
TYPES = ("rock", "metal", "ice")
w = oracle.World(seed=0x5EED0001, max_bodies=300)
rng = np.random.default_rng(1)
for i in range(s.n):
    w.add(s.pos_mass[i, 3], s.pos_mass[i, :3], s.vel[i, :3], "metal" if i == 0 else TYPES[int(rng.integers(0, 3))], s.temp[i])
st0 = w.state()
t0 = time.perf_counter()
w.integrate(s.dt, s.G, s.softening, collisions=True, nsteps=FRAMES * SUB)
cpu_full_s = time.perf_counter() - t0
world = {}
with nb.Engine(1024, precision="f64") as e:
    best = 1e9
    for _ in range(3):
        e.upload(st0["pos_mass"], st0["vel"], st0["acc"], st0["temp"])
        e.world_enable(seed=0x5EED0001 + 151, max_bodies=300)
        e.world_set_bodies(st0["id"], st0["type"], st0["radius"])
        e.world_step(s.dt, s.G, s.softening, SUB, collisions=True)
        e.world_sync()
        t0 = time.perf_counter()
        for _ in range(FRAMES):
            e.world_step(s.dt, s.G, s.softening, SUB, collisions=True)
        info = e.world_sync()
        best = min(best, time.perf_counter() - t0)
    best_dl = 1e9
    for _ in range(3):
        t0 = time.perf_counter()
        for _ in range(FRAMES):
            e.world_step(s.dt, s.G, s.softening, SUB, collisions=True)
            e.world_sync()
            e.download(temp=True)
        best_dl = min(best_dl, time.perf_counter() - t0)
    best_fr, fr = 1e9, None
    for _ in range(3):
        t0 = time.perf_counter()
        for _ in range(FRAMES):
            fr = e.world_frame(s.dt, s.G, s.softening, SUB, collisions=True, out=fr)
        best_fr = min(best_fr, time.perf_counter() - t0)
    world = {"resident_frame_us": best / FRAMES * 1e6, "frame_with_sync_and_download_us": best_dl / FRAMES * 1e6,
             "world_frame_us": best_fr / FRAMES * 1e6,
             "world_frame": "nbody_world_frame: the step kernel stores the frame's record into pinned host memory; one launch, one synchronise, "
                            "host arrays (pos_mass, vel, acc, temp, id, type, radius) filled every frame",
             "bodies_left": info["n"], "counters": info["counters"]}
print(json.dumps({"config": "BASELINE configs[0]: default scene N=151, 1000 x integrate(0.008), gravity only",
                  "complete_step_with_collisions": {"cpu_us_per_frame": cpu_full_s / FRAMES * 1e6, "cpu_bodies_left": len(w),
                                                    "cpu_counters": w.counters(), "gpu": world},
                  "cpu": {"kind": "port", "cores": 1, "us_per_step": cpu_s / (FRAMES * SUB) * 1e6,
                          "g_interactions_per_s": s.n * (s.n - 1) * FRAMES * SUB / cpu_s * 1e-9},
                  "gpu": out}))

#!/usr/bin/env python
"""Where the end-to-end step (bench.py `e2e`) spends its time: upload / step / download of the
public call, host arrays pinned, N = 1,048,576 fp32 by default."""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nbody_b200 as nb  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=1048576)
ap.add_argument("--reps", type=int, default=3)
a = ap.parse_args()
scene = nb.scenes.uniform_cube(a.n)
pm, vel, _ = scene.astype(np.float32)
host = {k: torch.empty((a.n, 4), dtype=torch.float32).pin_memory().numpy() for k in ("pos_mass", "vel", "acc")}
host["pos_mass"][:], host["vel"][:], host["acc"][:] = pm, vel, 0
rows = []
with nb.Engine(a.n, precision="f32") as e:
    for it in range(1 + a.reps):
        t0 = time.perf_counter()
        e.upload(host["pos_mass"], host["vel"], host["acc"])
        t1 = time.perf_counter()
        e.step(scene.dt, scene.G, scene.softening, 1)
        e.sync()
        t2 = time.perf_counter()
        e.download(pos_mass=True, vel=True, acc=True, out=host)
        t3 = time.perf_counter()
        if it:
            rows.append({"upload_ms": (t1 - t0) * 1e3, "step_ms": (t2 - t1) * 1e3, "download_ms": (t3 - t2) * 1e3})
print(json.dumps({"n": a.n, "uniform": os.environ.get("NBODY_UNIFORM", "1"), "reps": rows}))

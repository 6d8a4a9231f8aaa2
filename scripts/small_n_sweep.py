#!/usr/bin/env python
"""Per-step latency of nbody_step at the reference's own sizes (N <= 300, idx:57) and a little
above, both precisions, state resident: k steps in one call, wall clock around call + sync."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nbody_b200 as nb  # noqa: E402

for prec in ("f64", "f32"):
    for n in (151, 300, 512, 768, 1024, 2048, 3072, 4096, 6144, 8192):
        s = nb.scenes.galaxy_default() if n == 151 else nb.scenes.plummer(n)
        pm, vel, _ = s.astype(np.float32 if prec == "f32" else np.float64)
        with nb.Engine(max(n, 1024), precision=prec) as e:  # (k = 2000 steps per call; fewer for the larger systems)
            e.upload(pm, vel)
            e.step(s.dt, s.G, s.softening, 10)
            e.sync()
            best = {}
            for k in (2, 2000):                   # animate()'s two sub-steps per call, and a long run
                reps = 200 if k == 2 else 3
                t0 = time.perf_counter()
                for _ in range(reps):
                    e.step(s.dt, s.G, s.softening, k)
                    e.sync()
                best[k] = (time.perf_counter() - t0) / reps / k * 1e6
            info = e.info()
        print(json.dumps({"precision": prec, "n": n, "us_per_step_k2000": best[2000], "us_per_step_k2_with_sync": best[2],
                          "j_chunks": info["j_chunks"], "lib": os.path.basename(os.environ.get("NBODY_LIB", "default"))}), flush=True)

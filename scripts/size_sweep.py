#!/usr/bin/env python
"""Step time vs N for both precisions (state resident, nbody_step(k) + sync, wall clock):
shows the launch-latency regime of the reference's own N=151 scene, the fp64
reference-matching mode at config 2's N=16,384, and the fp32 kernel's ramp to its plateau."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nbody_b200 as nb  # noqa: E402

rows = []
for prec, sizes in (("f64", [151, 1024, 4096, 16384, 65536]), ("f32", [151, 1024, 4096, 16384, 65536, 131072, 262144, 524288])):
    for n in sizes:
        s = nb.scenes.galaxy_default() if n == 151 else nb.scenes.plummer(n)
        pm, vel, _ = s.astype(np.float32 if prec == "f32" else np.float64)
        with nb.Engine(s.n, precision=prec) as e:
            e.upload(pm, vel)
            e.step(s.dt, s.G, s.softening, 3)
            e.sync()
            inter = float(s.n) ** 2
            k = int(max(3, min(2000, 2e12 / inter if prec == "f32" else 1e11 / inter)))
            t0 = time.perf_counter()
            e.step(s.dt, s.G, s.softening, k)
            e.sync()
            dt = (time.perf_counter() - t0) / k
            info = e.info()
        rows.append({"precision": prec, "n": s.n, "steps": k, "us_per_step": dt * 1e6, "g_int_per_s": inter / dt * 1e-9,
                     "j_chunks": info["j_chunks"], "n_padded": info["n_padded"]})
        print(json.dumps(rows[-1]), flush=True)

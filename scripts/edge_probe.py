#!/usr/bin/env python
"""What the device does where the page produces NaN: two DISTINCT bodies at one point with softening 0
(idx:280: G / Math.pow(0, 1.5) = Infinity, Infinity * 0 = NaN, and the NaN spreads to every body at the next step)."""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nbody_b200 as nb  # noqa: E402

out = []
for prec in ("f64", "f32"):
    for n in (3, 700, 6000):                                  # cluster kernel, all-SM kernel, launch-per-step kernels
        dt = np.float64 if prec == "f64" else np.float32
        rng = np.random.default_rng(1)
        pm = np.zeros((n, 4), dtype=dt)
        pm[:, :3] = rng.uniform(-10, 10, (n, 3))
        pm[:, 3] = rng.uniform(0.5, 2.0, n)
        pm[1, :3] = pm[0, :3]                                 # bodies 0 and 1 coincide
        for eps in (0.0, 0.5):
            with nb.Engine(n, precision=prec) as e:
                e.upload(pm, np.zeros((n, 4), dtype=dt))
                e.force(1.0, eps)
                a_force = e.download()["acc"]
                e.upload(pm, np.zeros((n, 4), dtype=dt))
                e.step(0.01, 1.0, eps, 2)
                a_step = e.download()["acc"]
            out.append({"precision": prec, "n": n, "softening": eps,
                        "force_nan_bodies": int(np.isnan(a_force[:, :3]).any(axis=1).sum()), "force_inf_bodies": int(np.isinf(a_force[:, :3]).any(axis=1).sum()),
                        "step2_nan_bodies": int(np.isnan(a_step[:, :3]).any(axis=1).sum())})
for r in out:
    print(json.dumps(r))

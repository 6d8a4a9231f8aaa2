#!/usr/bin/env python
"""fp64 pairs-once pass (exact hi/lo fixed-point accumulators) on 1..P GPUs of one process: device time of the force pass per step.
   python scripts/f64_multi_gpu.py [n]"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nbody_b200 as nb  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
s = nb.scenes.plummer(n)
for general in (False, True):
    pm = s.pos_mass.copy()
    if general:
        pm[:, 3] *= np.random.default_rng(1).uniform(0.5, 1.5, n)
    for g in [k for k in (1, 2, 4, 8) if k <= nb.device_count()]:
        for sym in (1, 0):
            os.environ["NBODY_F64_SYM"] = str(sym)
            with nb.Engine(n, precision="f64", n_gpus=g) as e:
                e.upload(pm, s.vel)
                e.step(s.dt, s.G, s.softening, 2)
                e.sync()
                e.enable_timing(True)
                e.read_timing()
                e.step(s.dt, s.G, s.softening, 10)
                e.sync()
                k, ms = e.read_timing()
                info = e.info()
            print(json.dumps({"n": n, "gpus": g, "general_masses": general, "pairs_once": bool(info["f64_symmetric"]), "force_ms": ms / k,
                              "G_int_per_s": float(n) ** 2 / (ms / k * 1e-3) * 1e-9}), flush=True)
os.environ.pop("NBODY_F64_SYM", None)

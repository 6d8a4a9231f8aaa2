#!/usr/bin/env python
"""One nbody_step call of a mid-size fp32 system through the cooperative launch (csrc/sym_coop.cuh), for the profiler:
   ncu --set full -k regex:sym_coop -c 1 python scripts/coop_probe.py [n] [nsteps]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nbody_b200 as nb  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
nsteps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
s = nb.scenes.plummer(n)
pm, vel, _ = s.astype(np.float32)
with nb.Engine(s.n, precision="f32") as e:
    e.upload(pm, vel)
    l0 = e.info()["kernel_launches"]
    e.step(s.dt, s.G, s.softening, nsteps)
    e.sync()
    info = e.info()
    print({"n": s.n, "nsteps": nsteps, "launches": info["kernel_launches"] - l0, "warp_units": info["f32_sym_warp_units"],
           "bodies_per_thread": info["f32_sym_bodies_per_thread"]})

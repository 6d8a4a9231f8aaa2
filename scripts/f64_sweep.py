#!/usr/bin/env python
"""fp64 force kernel: arithmetic variant (NBODY_F64_MATH 0/1/2, csrc/force_f64.cuh) x work-unit
count (NBODY_WANT_UNITS) at config 2's N=16,384 and two neighbours.  For each variant also
the force error against the CPU oracle at N=4096 (norm-relative)."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nbody_b200 as nb  # noqa: E402
from oracle import oracle  # noqa: E402  (error column only)


def err(a, b):
    return float(np.linalg.norm(a[:, :3] - b[:, :3], axis=1).max() / np.linalg.norm(b[:, :3], axis=1).max())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "f64_sweep.jsonl"))
    ap.add_argument("--sizes", default="4096,16384,65536")
    ap.add_argument("--units", default="0,28416,7104,3552,1776")
    ap.add_argument("--math", default="0,1,2")
    ap.add_argument("--tag", default="")
    a = ap.parse_args()
    s4 = nb.scenes.plummer(4096)
    ref4 = oracle.compute_gravity(s4.pos_mass, s4.G, s4.softening, mt=True)
    rows = []
    for math in [int(m) for m in a.math.split(",")]:
        os.environ["NBODY_F64_MATH"] = str(math)
        with nb.Engine(4096, precision="f64") as e:
            e.upload(s4.pos_mass, s4.vel)
            e.force(s4.G, s4.softening)
            e4 = err(e.download()["acc"], ref4)
        for units in [int(u) for u in a.units.split(",")]:
            if units:
                os.environ["NBODY_WANT_UNITS"] = str(units)
            else:
                os.environ.pop("NBODY_WANT_UNITS", None)
            for n in [int(x) for x in a.sizes.split(",")]:
                s = nb.scenes.plummer(n)
                with nb.Engine(n, precision="f64") as e:
                    e.upload(s.pos_mass, s.vel)
                    e.step(s.dt, s.G, s.softening, 3)
                    e.sync()
                    k = int(max(3, min(500, 4e10 / float(n) ** 2)))
                    t0 = time.perf_counter()
                    e.step(s.dt, s.G, s.softening, k)
                    e.sync()
                    dt = (time.perf_counter() - t0) / k
                    info = e.info()
                rows.append({"lib": os.path.basename(os.environ.get("NBODY_LIB", "default")), "tag": a.tag, "uniform_kernel": info["uniform_mass"], "pairs_once_k": info["f64_symmetric"],
                             "i_tile": info["i_tile"], "math": math, "want_units": units or "default", "n": n, "us_per_step": dt * 1e6,
                             "g_int_per_s": float(n) ** 2 / dt * 1e-9, "j_chunks": info["j_chunks"], "force_err_n4096": e4})
                print(json.dumps(rows[-1]), flush=True)
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    with open(a.out, "w") as f:
        for r in rows:
            f.write(json.dumps(r) + "\n")


if __name__ == "__main__":
    main()

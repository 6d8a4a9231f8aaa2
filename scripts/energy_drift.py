#!/usr/bin/env python
"""BASELINE.json configs[2]: N=131,072 fp32 two-galaxy collision on 1 B200, 10k steps of
integrate(dt) with the total energy (fp64, evaluated on the device) sampled every
--every steps.  Prints one JSON line: max over t of |E(t)-E(0)|/|E(0)|, momentum drift,
and the throughput of the run.   python scripts/energy_drift.py [--n N] [--steps S]"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nbody_b200 as nb  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=131072)
    ap.add_argument("--steps", type=int, default=10000)
    ap.add_argument("--every", type=int, default=500)
    ap.add_argument("--precision", default="f32")
    a = ap.parse_args()
    s = nb.scenes.two_galaxies(a.n)
    dt = np.float32 if a.precision == "f32" else np.float64
    pm, vel, _ = s.astype(dt)
    samples = []
    with nb.Engine(s.n, precision=a.precision) as e:
        e.upload(pm, vel)
        ke, pe, mom = e.energy(s.G, s.softening)
        e0, p0 = ke + pe, mom.copy()
        samples.append({"step": 0, "E": e0, "K": ke, "P": pe})
        t_step = 0.0
        done = 0
        while done < a.steps:
            k = min(a.every, a.steps - done)
            t0 = time.perf_counter()
            e.step(s.dt, s.G, s.softening, k)
            e.sync()
            t_step += time.perf_counter() - t0
            done += k
            ke, pe, mom = e.energy(s.G, s.softening)
            samples.append({"step": done, "E": ke + pe, "K": ke, "P": pe, "dP": float(np.abs(mom - p0).max())})
        info = e.info()
    drift = max(abs(x["E"] - e0) / abs(e0) for x in samples)
    out = {"scene": f"two_galaxies N={s.n} (configs[2])", "precision": a.precision, "steps": a.steps, "dt": s.dt, "G": s.G,
           "softening": s.softening, "E0": e0, "max_rel_energy_drift": drift,
           "final_rel_energy_error": abs(samples[-1]["E"] - e0) / abs(e0),
           "max_momentum_drift": max(x.get("dP", 0.0) for x in samples), "total_momentum_scale": float(np.abs(p0).max()),
           "seconds_stepping": t_step, "ms_per_step": 1e3 * t_step / a.steps,
           "g_interactions_per_s": float(s.n) ** 2 * a.steps / t_step * 1e-9, "j_chunks": info["j_chunks"],
           "samples": samples}
    print(json.dumps(out))


if __name__ == "__main__":
    main()

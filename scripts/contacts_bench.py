#!/usr/bin/env python
"""Measure the collision broad phase on one B200.

For each scene: the fp32 force sweep with and without nearest-neighbour tracking (the
cost of fusing the overlap search into the sweep), nbody_contacts with the flags
(sparse scan) and without (every body scanned), and the CPU oracle's overlap loop
(idx:301-313 restated; test infrastructure, timed as the CPU baseline) on a bounded
sample.  Pair tests are counted one-sided like interactions: N^2 per search.

    python scripts/contacts_bench.py [--out gpurun_out/contacts_bench.json]
"""
import argparse
import json
import math
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nbody_b200 as nb  # noqa: E402
from oracle import oracle  # noqa: E402  (CPU baseline leg only)


def timed_force(e, G, eps, reps):
    e.force(G, eps)
    e.sync()
    e.enable_timing(True)
    for _ in range(reps):
        e.force(G, eps)
    n, ms = e.read_timing()
    e.enable_timing(False)
    return ms / n


def wall(fn, reps):
    fn()
    t0 = time.perf_counter()
    for _ in range(reps):
        out = fn()
    return (time.perf_counter() - t0) / reps * 1e3, out


def run(name, scene, rad_scale, reps):
    pm, vel, _ = scene.astype(np.float32)
    n = scene.n
    rad = ((3 * pm[:, 3].astype(np.float64) / (4 * math.pi)) ** (1 / 3) * rad_scale).astype(np.float32)
    row = {"scene": name, "n": n, "radius_scale": rad_scale, "r_max": float(rad.max()), "r_median": float(np.median(rad))}
    with nb.Engine(n, precision="f32") as e:
        e.upload(pm, vel)
        row["force_ms"] = timed_force(e, scene.G, scene.softening, reps)
        e.set_radius(rad)
        row["force_near_ms"] = timed_force(e, scene.G, scene.softening, reps)
        row["near_overhead"] = row["force_near_ms"] / row["force_ms"] - 1.0
        if os.environ.get("CONTACTS_QUICK"):
            row["lib"] = os.environ.get("NBODY_LIB", "default")
            return row
        ms, pairs = wall(lambda: e.contacts(0.0), reps)
        st = e.contact_stats()
        row.update(contacts_flagged_ms=ms, pairs=int(len(pairs)), flagged=st["flagged"], used_flags=st["used_flags"])
        e.set_radius(rad)                                   # drops the flags: full scan
        ms_full, pairs_full = wall(lambda: e.contacts(0.0), max(1, reps // 2))
        assert np.array_equal(pairs, pairs_full)
        row["contacts_full_scan_ms"] = ms_full
        row["full_scan_gtests_per_s"] = n * n / (ms_full * 1e-3) / 1e9
        row["fused_gtests_per_s"] = n * n / ((row["force_near_ms"] - row["force_ms"] + ms) * 1e-3) / 1e9
    ns = min(n, 32768)
    pm64, r64 = pm[:ns].astype(np.float64), rad[:ns].astype(np.float64)
    t0 = time.perf_counter()
    oracle.overlap_pairs(pm64, r64, 0.0, mt=False)
    sec = time.perf_counter() - t0
    row["cpu_oracle"] = {"sample_n": ns, "seconds": sec, "threads": 1, "gtests_per_s": ns * (ns - 1) / sec / 1e9,
                         "note": "oracle restatement of idx:301-313, N(N-1) one-sided tests = N(N-1)/2 pairs"}
    return row


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "contacts_bench.json"))
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--big", action="store_true", help="also N=1,048,576")
    a = ap.parse_args()
    rows = [run("two_galaxies", nb.scenes.two_galaxies(131072), 4.0, a.reps)]
    if a.big:
        s = nb.scenes.uniform_cube(1048576)
        rows.append(run("uniform_cube", s, 20.0, 2))
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    with open(a.out, "w") as f:
        for r in rows:
            f.write(json.dumps(r) + "\n")
            print(json.dumps(r))


if __name__ == "__main__":
    main()

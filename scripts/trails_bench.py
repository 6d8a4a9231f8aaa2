#!/usr/bin/env python
"""Measure the renderer-feed kernels on one B200 against the HBM
roofline: history push (16 B read + 16 B written per body, fp32) and the ribbon builder
(depth x 16 B read + depth x 24 B written per body), with the CPU oracle's ribbon loop
(idx:563-592 restated) timed on a bounded sample.

    python scripts/trails_bench.py [--n 1048576] [--depth 80]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nbody_b200 as nb  # noqa: E402
from oracle import oracle  # noqa: E402  (CPU baseline leg only)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=1048576)
    ap.add_argument("--depth", type=int, default=80)
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "trails_bench.json"))
    a = ap.parse_args()
    import torch
    s = nb.scenes.uniform_cube(a.n)
    pm, vel, _ = s.astype(np.float32)
    rad = np.full(a.n, 0.1, np.float32)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm = float(peaks.get("hbm_gbs", 6552.0))
    st = torch.cuda.Stream()
    row = {"n": a.n, "depth": a.depth, "hbm_peak_gbs": hbm, "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback 6552"}
    with nb.Engine(a.n, precision="f32") as e:
        e.set_stream(st.cuda_stream)
        e.upload(pm, vel)
        e.set_radius(rad)
        e.history_enable(a.depth)
        e.history_push()
        e.step(s.dt, s.G, s.softening, 1)
        flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

        def timed(fn):
            ms = []
            with torch.cuda.stream(st):
                for _ in range(3):
                    fn()
                for _ in range(a.reps):
                    flush.zero_()                           # inputs evicted from the 126 MB L2
                    t0, t1 = torch.cuda.Event(True), torch.cuda.Event(True)
                    t0.record(st)
                    fn()
                    t1.record(st)
                    t1.synchronize()
                    ms.append(t0.elapsed_time(t1))
            return float(np.median(ms))
        push_ms = timed(e.history_push)
        rib_ms = timed(lambda: e.trail_ribbons((0, 80, 150), download=False))
        push_bytes = a.n * 32
        rib_bytes = a.n * a.depth * (16 + 24) + a.n * 4
        row["push"] = {"ms": push_ms, "algorithmic_bytes": push_bytes, "gbs": push_bytes / push_ms / 1e6, "frac": push_bytes / push_ms / 1e6 / hbm}
        row["ribbons"] = {"ms": rib_ms, "algorithmic_bytes": rib_bytes, "gbs": rib_bytes / rib_ms / 1e6, "frac": rib_bytes / rib_ms / 1e6 / hbm,
                          "vertices_per_s": a.n * a.depth * 2 / rib_ms * 1e3}
        ns = min(a.n, 65536)
        hist = e.history()[:ns].astype(np.float64)
    t0 = time.perf_counter()
    oracle.trail_ribbons(hist, rad[:ns].astype(np.float64), (0, 80, 150))
    sec = time.perf_counter() - t0
    row["cpu_oracle"] = {"sample_n": ns, "seconds": sec, "threads": 1, "vertices_per_s": ns * a.depth * 2 / sec}
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    with open(a.out, "w") as f:
        f.write(json.dumps(row) + "\n")
    print(json.dumps(row))


if __name__ == "__main__":
    main()

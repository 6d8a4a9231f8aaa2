#!/usr/bin/env python
"""Measure the headless body renderer on one B200: frames and shaded
pixels per second for the reference's default scene at 1080p and 4K, and for a dense
field that covers the whole frame, next to the CPU oracle (the same shader arithmetic in
C, all host threads) on a bounded frame.

    python scripts/render_bench.py [--out gpurun_out/render_bench.json]
"""
import argparse
import importlib.util
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nbody_b200 as nb  # noqa: E402
from oracle import oracle  # noqa: E402  (CPU baseline leg only)


def gpu_frame_ms(e, w, h, reps):
    import torch
    st = torch.cuda.Stream()
    e.set_stream(st.cuda_stream)
    ms = []
    with torch.cuda.stream(st):
        for _ in range(3):
            e.render(w, h, download=False)
        for _ in range(reps):
            t0, t1 = torch.cuda.Event(True), torch.cuda.Event(True)
            t0.record(st)
            e.render(w, h, download=False)
            t1.record(st)
            t1.synchronize()
            ms.append(t0.elapsed_time(t1))
    e.set_stream(0)
    return float(np.median(ms))


def run(name, pm, radius, temp, types, seed, sizes, reps, cpu_size):
    rows = []
    n = pm.shape[0]
    with nb.Engine(n, precision="f32") as e:
        e.upload(pm.astype(np.float32), np.zeros((n, 4), np.float32), None, temp.astype(np.float32))
        e.set_radius(radius.astype(np.float32))
        e.set_materials(types, seed)
        for (w, h) in sizes:
            _, ids = e.render(w, h, ids=True)
            covered = int((ids >= 0).sum())
            ms = gpu_frame_ms(e, w, h, reps)
            t0 = time.perf_counter()
            e.render(w, h)
            e2e = (time.perf_counter() - t0) * 1e3
            rows.append({"scene": name, "n": n, "width": w, "height": h, "covered_pixels": covered, "frame_ms": ms,
                         "frames_per_s": 1e3 / ms, "mpixels_per_s": w * h / ms / 1e3, "shaded_mpixels_per_s": covered / ms / 1e3,
                         "frame_with_download_ms": e2e, "d2h_bytes": w * h * 16})
    w, h = cpu_size
    t0 = time.perf_counter()
    _, ids = oracle.render(pm, radius, temp, types, seed, w, h)
    sec = time.perf_counter() - t0
    rows.append({"scene": name, "impl": "cpu_oracle", "n": n, "width": w, "height": h, "covered_pixels": int((ids >= 0).sum()),
                 "frame_ms": sec * 1e3, "mpixels_per_s": w * h / sec / 1e6, "threads": oracle.max_threads(),
                 "note": "brute force over bodies per pixel (O(pixels x n)); the restated shader arithmetic, OpenMP over rows"})
    return rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "render_bench.json"))
    ap.add_argument("--reps", type=int, default=20)
    a = ap.parse_args()
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(ROOT, "tests", "golden", "make_golden.py"))
    mg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mg)
    s = nb.scenes.galaxy_default()
    radius, temp, types, seed = mg.render_inputs(s)
    rows = run("galaxy_default", s.pos_mass, radius / 3.0, temp, types, seed, [(1920, 1080), (3840, 2160)], a.reps, (480, 270))
    rng = np.random.default_rng(1)
    n = 4096                                                    # a wall of spheres filling the frame: every pixel shaded
    pm = np.concatenate([rng.uniform(-160, 160, (n, 2)), rng.uniform(-40, 0, (n, 1)), np.ones((n, 1))], axis=1)
    rows += run("dense_wall", pm, np.full(n, 9.0), rng.uniform(0, 150, n), rng.integers(0, 3, n).astype(np.int32),
                rng.uniform(0, 100, n).astype(np.float32), [(1920, 1080), (3840, 2160)], a.reps, (240, 135))
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    with open(a.out, "w") as f:
        for r in rows:
            f.write(json.dumps(r) + "\n")
            print(json.dumps(r))


if __name__ == "__main__":
    main()

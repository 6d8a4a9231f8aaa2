#!/usr/bin/env python
"""Kernel-variant timings of the fp32 force pass on one GPU (builder's lab, not the bench):
one JSON line per (scene, n, variant) with the device time of the force pass per step."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nbody_b200 as nb  # noqa: E402


def run(scene, n, general, env, steps=3, label=""):
    old = {k: os.environ.get(k) for k in env}
    os.environ.update({k: str(v) for k, v in env.items()})
    try:
        s = nb.scenes.SCENES[scene](n)
        pm, vel, _ = s.astype(np.float32)
        if general:
            pm[:, 3] = (pm[:, 3].mean() * np.random.default_rng(1).uniform(0.5, 1.5, s.n)).astype(np.float32)
        with nb.Engine(s.n, precision="f32") as e:
            e.upload(pm, vel)
            e.step(s.dt, s.G, s.softening, 1)
            e.sync()
            e.enable_timing(True)
            e.read_timing()
            t0 = time.perf_counter()
            e.step(s.dt, s.G, s.softening, steps)
            e.sync()
            wall = (time.perf_counter() - t0) / steps
            k, ms = e.read_timing()
            info = e.info()
        ms_step = ms / max(k, 1)
        print(json.dumps({"scene": scene, "n": n, "general_masses": general, "variant": label, "env": env, "force_ms": ms_step,
                          "wall_ms": wall * 1e3, "G_int_per_s": float(n) ** 2 / (ms_step * 1e-3) * 1e-9,
                          "sym": info["f32_symmetric"], "sym_R": info["f32_sym_bodies_per_thread"], "sym_used": info["f32_sym_last_pass"],
                          "uniform": info["uniform_mass"], "j_chunks": info["j_chunks"]}), flush=True)
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


def run_wall(scene, n, general, env, steps=200, label=""):
    """Wall time per step of `steps` fused steps (no per-pass events: the one-launch kernels stay eligible), best of 3."""
    old = {k: os.environ.get(k) for k in env}
    os.environ.update({k: str(v) for k, v in env.items()})
    try:
        s = nb.scenes.SCENES[scene](n)
        pm, vel, _ = s.astype(np.float32)
        if general:
            pm[:, 3] = (pm[:, 3].mean() * np.random.default_rng(1).uniform(0.5, 1.5, s.n)).astype(np.float32)
        with nb.Engine(s.n, precision="f32") as e:
            e.upload(pm, vel)
            e.step(s.dt, s.G, s.softening, 3)
            e.sync()
            best = 1e9
            for _ in range(3):
                t0 = time.perf_counter()
                e.step(s.dt, s.G, s.softening, steps)
                e.sync()
                best = min(best, (time.perf_counter() - t0) / steps)
            info = e.info()
        print(json.dumps({"scene": scene, "n": s.n, "general_masses": general, "variant": label, "env": env, "us_per_step": best * 1e6,
                          "G_int_per_s": float(s.n) ** 2 / best * 1e-9, "sym": info["f32_symmetric"], "sym_R": info["f32_sym_bodies_per_thread"],
                          "warp_units": info["f32_sym_warp_units"], "sym_used": info["f32_sym_last_pass"], "uniform": info["uniform_mass"]}), flush=True)
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    N = int(os.environ.get("LAB_N", 1 << 20))
    tag = os.environ.get("LAB_TAG", "")
    variants = [("sym R4", {"NBODY_F32_SYM_R": 4}), ("sym R8", {"NBODY_F32_SYM_R": 8}), ("one-sided", {"NBODY_F32_SYM": 0})]
    if which in ("all", "1m"):
        for general in (False, True):
            for label, env in variants:
                run("uniform_cube", N, general, env, label=label)
        for run_len in (1, 2, 8, 16):
            run("uniform_cube", N, False, {"NBODY_F32_SYM_R": 4, "NBODY_F32_SYM_RUN": run_len}, label=f"sym R4 run={run_len}")
    if which in ("all", "mid"):
        for n in (32768, 65536, 131072, 262144):
            for label, env in variants:
                run("two_galaxies", n, False, env, steps=5, label=label)
        for n in (16384, 32768):
            for label, env in variants:
                e2 = dict(env)
                e2["NBODY_F32_SYM_MIN"] = 1
                run("plummer", n, True, e2, steps=5, label=label + " (forced)")
    if which == "onesided":      # a kernel-variant build (NBODY_LIB) of the one-sided sweep
        for general in (False, True):
            run("uniform_cube", N, general, {"NBODY_F32_SYM": 0}, label="one-sided " + tag)
    if which == "symx":          # the other bodies-per-thread choice of each mass variant
        run("uniform_cube", N, False, {"NBODY_F32_SYM_R": 4}, label="sym R4 " + tag)
        run("uniform_cube", N, True, {"NBODY_F32_SYM_R": 8}, label="sym R8 " + tag)
    if which == "ncu":           # one short run for the profiler: one warm step, one profiled
        run("uniform_cube", N, os.environ.get("LAB_GENERAL", "0") == "1", {}, steps=1, label="ncu " + tag)
    if which == "sym":           # ... and of the pairs-once pass, default bodies per thread
        for general in (False, True):
            run("uniform_cube", N, general, {}, label="sym " + tag)
    if which == "fine":          # the warp-unit pairs-once pass against what each size took before it
        wide = {"NBODY_F32_SYMW_MIN": 1, "NBODY_F32_SYMW_MAX": 1 << 30, "NBODY_SMALL": 0, "NBODY_MEDIUM": 0}
        sizes = [int(x) for x in os.environ.get("LAB_SIZES", "2048 4096 6144 8192 12288 16384 24576 32768 49152 65536 131072").split()]
        for n in sizes:
            for general in (False, True):
                steps = 200 if n <= 32768 else 20
                run_wall("plummer", n, general, {"NBODY_F32_SYMW": 0}, steps=steps, label="before")
                run_wall("plummer", n, general, {"NBODY_SYM_COOP": 0}, steps=steps, label="per step")
                run_wall("plummer", n, general, {}, steps=steps, label="default")
                for R in [int(x) for x in os.environ.get("LAB_RS", "2 4 8").split()]:
                    run_wall("plummer", n, general, dict(wide, NBODY_F32_SYMW_R=R), steps=steps, label=f"warp units R{R}")
    if which == "big":           # the warp-unit pass against the CTA-wide one at large N
        wide = {"NBODY_F32_SYMW_MIN": 1, "NBODY_F32_SYMW_MAX": 1 << 30}
        for n in [int(x) for x in os.environ.get("LAB_SIZES", "262144 1048576").split()]:
            for general in (False, True):
                run("uniform_cube", n, general, {"NBODY_F32_SYMW": 0}, label="CTA-wide " + tag)
                run("uniform_cube", n, general, {}, label="default " + tag)
                for R in [int(x) for x in os.environ.get("LAB_RS", "4 8").split()]:
                    run("uniform_cube", n, general, dict(wide, NBODY_F32_SYMW_R=R), label=f"warp units R{R} {tag}")


#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric on its config: G body-body interactions/s
of the all-pairs gravity + kick/drift step at N = 1,048,576 (fp32 uniform cube,
configs[3]), weak-sharded over --gpus B200s (one process per GPU under
torchrun), with the roofline fraction and the CPU baseline in the same line.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--n BODIES] [--impl reference]

A "step" is one reference integrate(dt) (/root/reference/index.html:247-262 minus
collisions) = N^2 one-sided interactions.  `value` is timed with CUDA events on
the launching stream with state resident in HBM; `e2e` goes through the C ABI
with HOST buffers (upload -> step -> download inside the timed region).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "all-pairs interactions/s at N=1M (N^2 per step)"
UNIT = "G interactions/s"
ISSUE_SLOTS = 13          # 12 FP32-pipe + 1 MUFU per interaction
FP32_LANES_PER_SM = 128


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d.get("hbm_gbs", 6650.0), "sm_max_mhz": d.get("sm_max_mhz", 1965.0), "src": "measured"}
    return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0, "src": "fallback"}


class ClockSampler:
    """nvidia-smi clocks/throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.rows, self.proc, self.idx = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.idx}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
        sm = [float(r[1]) for r in self.rows if len(r) >= 8 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 8 and r[2].replace(".", "").isdigit()]
        pw = [float(r[3]) for r in self.rows if len(r) >= 8 and r[3].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[k] for r in self.rows if len(r) >= 8 for k in range(4) if r[4 + k].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": reasons}


def cpu_baseline_sample(scene, sample_n: int, steps: int):
    """The oracle (C port of the reference's JS step, 1 thread like the reference's
    main-thread loop, idx:873-883) on a bounded sample of the same workload."""
    from oracle import oracle
    pm, vel = scene.pos_mass[:sample_n].copy(), scene.vel[:sample_n].copy()
    acc = np.zeros_like(vel)
    oracle.integrate(pm[:256], vel[:256], acc[:256], None, scene.dt, scene.G, scene.softening, nsteps=1)  # warm
    t0 = time.perf_counter()
    oracle.integrate(pm, vel, acc, None, scene.dt, scene.G, scene.softening, nsteps=steps)
    dt = time.perf_counter() - t0
    inter = float(sample_n) * (sample_n - 1) * steps     # one-sided; the loop evaluates half as pairs
    return {"value": inter / dt * 1e-9, "unit": UNIT, "cores": 1, "kind": "port",
            "sample": f"first {sample_n} bodies of the scene, {steps} x integrate(dt), N(N-1) interactions "
                      f"= {int(inter / 2)} pair evaluations/2 per the reference's i<j loop; {dt:.2f} s; host has {os.cpu_count()} cores",
            "seconds": dt}


def run_reference(args, rank):
    """--impl reference: the reference's own CPU path.  Its JS cannot run here (no
    JS engine in the image), so this is the C restatement in oracle/, one thread."""
    if rank != 0:
        return
    import nbody_b200 as nb
    scene = nb.scenes.SCENES[args.scene](args.n)
    sample_n = min(args.n, args.ref_sample)
    times = []
    from oracle import oracle
    pm, vel = scene.pos_mass[:sample_n].copy(), scene.vel[:sample_n].copy()
    acc = np.zeros_like(vel)
    for s in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        pm, vel, acc, _ = oracle.integrate(pm, vel, acc, None, scene.dt, scene.G, scene.softening, nsteps=1)
        if s >= args.warmup:
            times.append(time.perf_counter() - t0)
    ms = 1e3 * sum(times) / len(times)
    val = float(sample_n) * (sample_n - 1) / (ms * 1e-3) * 1e-9
    # courtesy line, NOT the reference's path (it is single-threaded): the same force loop, all-pairs
    # per body, OpenMP over i on every host core
    t0 = time.perf_counter()
    oracle.compute_gravity(pm, scene.G, scene.softening, mt=True)
    t_mt = time.perf_counter() - t0
    all_cores = {"value": float(sample_n) * (sample_n - 1) / t_mt * 1e-9, "unit": UNIT, "cores": oracle.max_threads(),
                 "note": "courtesy: OpenMP over i on all host cores, force evaluation only; not the reference's (single-threaded) path"}
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"{scene.name} N={args.n} (configs[{scene.meta['config'] - 1}]); each step = one integrate(dt) on the first "
                                   f"{sample_n} bodies (bounded sample)", "n": args.n, "sample_n": sample_n},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": 1, "kind": "port",
                             "sample": f"first {sample_n} bodies, {args.steps} steps, 1 thread (the reference is "
                                       f"single-threaded JS, idx:873-883; Node/V8 unavailable in this image: the C restatement, bit-identical to the "
                                       f"page's own source executed by oracle/js_reference.py); host has {os.cpu_count()} cores"},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "all_cores": all_cores}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--n", "--bodies", dest="n", type=int, default=1048576)   # --bodies: torchrun's argparse trips over --n
    ap.add_argument("--scene", default="uniform_cube", choices=["uniform_cube", "cold_collapse"],
                    help="uniform_cube = BASELINE configs[3] (the metric's workload); cold_collapse = configs[4], with --bodies 4194304 --gpus 8")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--ref-sample", type=int, default=16384)
    ap.add_argument("--cpu-sample", type=int, default=16384)
    ap.add_argument("--cpu-steps", type=int, default=4)    # ~11 s of one host core
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            sys.exit("launch with torchrun for --gpus > 1 (one process per GPU)")

    import torch
    import torch.distributed as dist
    import nbody_b200 as nb

    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    scene = nb.scenes.SCENES[args.scene](args.n)
    pm, vel, _ = scene.astype(np.float32)
    eng = nb.create_rank_engine(args.n, "f32")
    # a real (non-default) stream: the engine, the L2 flush and the timing events all run on it
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    assert stream.cuda_stream != 0
    eng.set_stream(stream.cuda_stream)
    eng.upload(pm, vel)
    info = eng.info()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")   # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def one_step():
        flush.zero_()
        eng.step(scene.dt, scene.G, scene.softening, 1)

    for _ in range(args.warmup):
        one_step()
    barrier()
    eng.enable_timing(True)
    eng.read_timing()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = eng.info()["kernel_launches"]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    for _ in range(args.steps):
        one_step()
    e1.record(stream)
    barrier()
    ms_total = e0.elapsed_time(e1)
    launches = eng.info()["kernel_launches"] - l0
    k_launches, k_ms = eng.read_timing()
    eng.enable_timing(False)
    clocks = sampler.stop() if rank == 0 else None

    # ---- e2e: host buffers through the public call, copies inside the timed region ----
    e2e = None
    if not args.no_e2e:
        lo, hi = info["i_begin"], info["i_end"]
        host = {k: torch.empty((args.n, 4), dtype=torch.float32).pin_memory().numpy() for k in ("pos_mass", "vel", "acc")}
        host["pos_mass"][:], host["vel"][:], host["acc"][:] = pm, vel, 0
        e2e_steps = max(1, min(args.steps, 3))
        for it in range(1 + e2e_steps):
            if it == 1:
                barrier()
                t0 = time.perf_counter()
            eng.upload(host["pos_mass"], host["vel"], host["acc"])
            eng.step(scene.dt, scene.G, scene.softening, 1)
            eng.download(pos_mass=True, vel=True, acc=True, out=host)
        barrier()
        e2e_s = (time.perf_counter() - t0) / e2e_steps
        h2d = args.n * 16 + 2 * (hi - lo) * 16
        d2h = args.n * 16 + 2 * (hi - lo) * 16

    def allmax(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def allsum(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    ms_total = allmax(ms_total)
    if not args.no_e2e:
        e2e_s = allmax(e2e_s)
        h2d, d2h = allsum(h2d), allsum(d2h)
    launches_all = allsum(launches)
    k_ms_max = allmax(k_ms / max(k_launches, 1))

    if rank == 0:
        pk = peaks()
        inter_step = float(args.n) ** 2
        ms_step = ms_total / args.steps
        value = inter_step / (ms_step * 1e-3) * 1e-9
        sms = info["sm_count"]
        peak_max = sms * FP32_LANES_PER_SM * pk["sm_max_mhz"] * 1e6 / ISSUE_SLOTS * 1e-9      # G int/s per GPU
        # dominant kernel: the force kernel; its launch processes (N/world) x N interactions
        k_inter = inter_step / world
        achieved = k_inter / (k_ms_max * 1e-3) * 1e-9
        traffic = traffic_source = None
        tp = os.path.join(ROOT, "profiles", "ncu_traffic.json")
        if os.path.exists(tp):                      # dram__bytes_read.sum + dram__bytes_write.sum of one ncu --set full capture
            t = json.load(open(tp))
            if t.get("n") == args.n and t.get("world") == world:
                traffic = t["dram_bytes_read"] + t["dram_bytes_write"]   # bytes per launch
                traffic_source = t["source"]
        roof = {"bound": "fp32_issue", "achieved": achieved, "peak": peak_max, "unit": UNIT + " per GPU",
                "frac": achieved / peak_max, "traffic": traffic, "traffic_source": traffic_source,
                "kernel": "force_f32_kernel", "kernel_ms": k_ms_max, "kernel_launches": k_launches,
                "peak_def": f"{sms} SMs x 128 FP32 lanes x {pk['sm_max_mhz']:.0f} MHz ({pk['src']} sm_max_mhz) / 13 issue slots per interaction",
                "frac_fp32_pipe_only": achieved / (peak_max * 13 / 12),
                "kernel_variant": ("equal-mass: 11 FP32 lane-ops + 1 MUFU per interaction (the per-pair mass multiply drops out; "
                                   "configs[3] has equal masses 1/N)" if info.get("uniform_mass") else
                                   "general: 12 FP32 lane-ops + 1 MUFU per interaction"),
                "frac_vs_own_op_count": achieved / (peak_max * 13 / (12 if info.get("uniform_mass") else 13)),
                "hbm": {"algorithmic_bytes_per_launch": 64.0 * args.n / world + 16.0 * args.n,
                        "achieved_gbs": (64.0 * args.n / world + 16.0 * args.n) / (k_ms_max * 1e-3) * 1e-9,
                        "peak_gbs": pk["hbm_gbs"], "note": "HBM is negligible for this path (SURVEY.md 8d)"}}
        if clocks and clocks.get("sm_mhz"):
            pc = sms * FP32_LANES_PER_SM * clocks["sm_mhz"] * 1e6 / ISSUE_SLOTS * 1e-9
            roof["frac_at_sampled_clock"] = achieved / pc
        line = {"metric": METRIC if args.n == 1048576 else METRIC.replace("N=1M", f"N={args.n}"), "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": f"{scene.name} N={args.n} fp32, G={scene.G:g} eps={scene.softening:g} dt={scene.dt:g} "
                                       f"(BASELINE configs[{scene.meta['config'] - 1}]); "
                                       "one step = integrate(dt) = N^2 interactions",
                           "n": args.n, "parallelism": f"i-range shards x{world}, positions exchanged by " + {0: "nothing (1 GPU)", 1: "ncclAllGather",
                                       2: "fused peer stores over NVLink + epoch barrier"}[info["exchange"]],
                           "l2": "flushed between steps (256 MiB write)", "j_chunks": info["j_chunks"], "i_tile": info["i_tile"],
                           "flop_convention": "20 flop/interaction", "tflops_20": value * 20e-3,
                           "equal_mass_kernel": bool(info.get("uniform_mass"))},
                "roofline": roof, "gpu_launches": int(launches_all), "clocks": clocks}
        if not args.no_e2e:
            line["e2e"] = {"value": inter_step / e2e_s * 1e-9, "unit": UNIT, "h2d_bytes_per_step": int(h2d),
                           "d2h_bytes_per_step": int(d2h), "ms_per_step": e2e_s * 1e3,
                           "api": "Engine.upload(host) -> Engine.step(dt,G,eps,1) -> Engine.download(host) (pinned host arrays)"}
        if not args.no_cpu and world == 1:
            line["cpu_baseline"] = cpu_baseline_sample(scene, min(args.n, args.cpu_sample), args.cpu_steps)
        print(json.dumps(line), flush=True)

    eng.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

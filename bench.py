#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric on its config: G body-body interactions/s
of the all-pairs gravity + kick/drift step at N = 1,048,576 (fp32 uniform cube,
configs[3]), shared by --gpus B200s (one process per GPU under torchrun), with
the roofline fraction, the CPU baseline and a correctness check of what was
timed in the same line.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--bodies N] [--scene S] [--general-masses]
                    [--impl reference] [--single-process] [--energy]

A "step" is one reference integrate(dt) (/root/reference/index.html:247-262 minus
collisions) = N^2 one-sided interactions (the reference's own i<j loop evaluates
N(N-1)/2 pairs and applies each to both bodies, idx:272-288; so does the fp32
production kernel, csrc/force_f32_sym.cuh).  `value` is timed with CUDA events on
the launching stream with state resident in HBM; `e2e` goes through the C ABI
with HOST buffers (upload -> step -> download inside the timed region).

What the line proves about itself (task 1):
  check.sampled_acc_err  accelerations of sampled bodies after the timed loop against the
                         oracle's sum over all N bodies at the same positions (gate 2e-5);
  check.state_hash       blake2b-64 of pos_mass after a fixed 3-step fused run from the scene:
                         the same on 1, 2, 4 and 8 GPUs (integer accumulation);
  fused                  the same K steps as ONE nbody_step(dt, G, eps, K) call.
The process exits non-zero when a check fails.
"""
from __future__ import annotations

import argparse
import hashlib
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
FP32_LANES_PER_SM = 128
ACC_GATE = 2e-5

# issue slots per ONE-SIDED interaction (one slot = one FP32 lane-op or one MUFU lane-op; a packed FFMA2/FADD2/FMUL2 holds the
# dispatch port for two cycles and does two lane-ops, profiles/r1_pipe_bench2.txt)
SLOTS = {
    "survey": 13.0,             # one-sided sweep, 12 FP32 + 1 MUFU
    "sweep_general": 13.0,      # csrc/force_f32.cuh
    "sweep_equal_mass": 12.0,   # ... without the per-pair mass multiply
    "pairs_once_general": 8.5,  # csrc/force_f32_sym.cuh: (16 FP32 + 1 MUFU) per PAIR = two interactions
    "pairs_once_equal_mass": 7.5,
}


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
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_min_mhz": min(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": reasons}


# ---------------------------------------------------------------------------------------------
# the CPU arm: the oracle (C restatement of the reference's JS step), ONE thread like the page's
# main-thread loop (idx:873-883), pinned to one core, several repetitions, the fastest one quoted
# ---------------------------------------------------------------------------------------------

class pinned_to_one_core:
    """os.sched_setaffinity to a single allowed core for the duration (the unpinned single-thread runs of round 1
    wandered between 0.077 and 0.118 G interactions/s on the same box)."""

    def __enter__(self):
        self.old = None
        try:
            self.old = os.sched_getaffinity(0)
            self.core = max(self.old)
            os.sched_setaffinity(0, {self.core})
        except (AttributeError, OSError):
            self.core = None
        return self

    def __exit__(self, *exc):
        if self.old is not None:
            try:
                os.sched_setaffinity(0, self.old)
            except OSError:
                pass


def cpu_step_times(scene, sample_n: int, reps: int, warm: int = 1):
    """`reps` timed single integrate(dt) calls on the first sample_n bodies of the scene (after `warm` untimed ones),
    pinned to one core.  Returns the list of seconds per step."""
    from oracle import oracle
    pm, vel = scene.pos_mass[:sample_n].copy(), scene.vel[:sample_n].copy()
    acc = np.zeros_like(vel)
    times = []
    with pinned_to_one_core() as pin:
        oracle.integrate(pm[:512], vel[:512], acc[:512], None, scene.dt, scene.G, scene.softening, nsteps=1)  # page in the library
        for s in range(warm + reps):
            t0 = time.perf_counter()
            pm, vel, acc, _ = oracle.integrate(pm, vel, acc, None, scene.dt, scene.G, scene.softening, nsteps=1)
            if s >= warm:
                times.append(time.perf_counter() - t0)
        core = pin.core
    return times, core


def cpu_record(times, sample_n, core):
    inter = float(sample_n) * (sample_n - 1)         # one-sided; the loop evaluates half as pairs
    rates = sorted(inter / t * 1e-9 for t in times)
    return {"value": rates[-1], "unit": UNIT, "cores": 1, "kind": "port",
            "min": rates[0], "median": rates[len(rates) // 2], "max": rates[-1], "repetitions": len(times),
            "pinned_core": core,
            "sample": f"first {sample_n} bodies of the scene, {len(times)} x one integrate(dt) = N(N-1) interactions "
                      f"= {int(inter / 2)} pair evaluations per the reference's i<j loop, 1 thread pinned to core {core}, the FASTEST repetition "
                      f"quoted ({min(times):.2f} s; slowest {max(times):.2f} s); the reference is single-threaded JS (idx:873-883), Node/V8 is not in "
                      f"this image: the C restatement, bit-identical to the page's own source executed by oracle/js_reference.py; host has "
                      f"{os.cpu_count()} cores"}


def run_reference(args, rank):
    """--impl reference: the reference's own CPU path.  Its JS cannot run here (no
    JS engine in the image), so this is the C restatement in oracle/, one thread."""
    if rank != 0:
        return
    import nbody_b200 as nb
    scene = make_scene(nb, args)
    sample_n = min(args.n, args.ref_sample)
    times, core = cpu_step_times(scene, sample_n, reps=args.steps, warm=args.warmup)
    rec = cpu_record(times, sample_n, core)
    from oracle import oracle
    pm = scene.pos_mass[:sample_n].copy()
    # courtesy line, NOT the reference's path (it is single-threaded): the same force loop, all-pairs
    # per body, OpenMP over i on every host core
    t0 = time.perf_counter()
    oracle.compute_gravity(pm, scene.G, scene.softening, mt=True)
    t_mt = time.perf_counter() - t0
    all_cores = {"value": float(sample_n) * (sample_n - 1) / t_mt * 1e-9, "unit": UNIT, "cores": oracle.max_threads(),
                 "note": "courtesy: OpenMP over i on all host cores, force evaluation only; not the reference's (single-threaded) path"}
    best = min(times)
    line = {"impl": "reference", "metric": metric_name(args), "value": rec["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * best, "ms_per_step_median": 1e3 * sorted(times)[len(times) // 2],
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"{scene.name} N={args.n} (configs[{scene.meta['config'] - 1}]); each step = one integrate(dt) on the first "
                                   f"{sample_n} bodies (bounded sample), fastest of {len(times)} steps", "n": args.n, "sample_n": sample_n},
            "cpu_baseline": rec,
            "e2e": {"value": rec["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "all_cores": all_cores}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------

def metric_name(args):
    return METRIC if args.n == 1048576 else METRIC.replace("N=1M", f"N={args.n}")


def make_scene(nb, args):
    scene = nb.scenes.SCENES[args.scene](args.n)
    if args.general_masses:
        # the reference's own mass recipe scaled to the scene (idx:758: Math.random()*2 + 0.5), seeded
        u = nb.scenes.splitmix64_uniform(0x5EED00AA, np.arange(scene.n))
        scene.pos_mass[:, 3] = scene.pos_mass[:, 3].mean() * (u * 2.0 + 0.5) / 1.5
        scene.name += " with random masses (idx:758)"
    return scene


def state_hash(a: np.ndarray) -> str:
    return hashlib.blake2b(np.ascontiguousarray(a).tobytes(), digest_size=8).hexdigest()


def kernel_model(info):
    if info.get("f32_sym_last_pass"):
        kernel = "force_f32_symw_kernel" if info.get("f32_sym_warp_units") else "force_f32_sym_kernel"   # same arithmetic per pair, warp-sized units
        return ("pairs_once_equal_mass" if info.get("uniform_mass") else "pairs_once_general"), kernel + " (+ f32_sym_finish_kernel)"
    return ("sweep_equal_mass" if info.get("uniform_mass") else "sweep_general"), "force_f32_kernel"


def run_energy(args, nb):
    """BASELINE.json configs[2]: N=131,072 fp32 two-galaxy collision on 1 B200, --steps steps of integrate(dt) with the
    total energy (fp64, evaluated on the device) sampled every --energy-every steps."""
    scene = make_scene(nb, args)
    pm, vel, _ = scene.astype(np.float32)
    samples = []
    with nb.Engine(scene.n, precision="f32") as e:
        e.upload(pm, vel)
        ke, pe, mom = e.energy(scene.G, scene.softening)
        e0, p0 = ke + pe, mom.copy()
        samples.append({"step": 0, "E": e0})
        t_step, done = 0.0, 0
        while done < args.steps:
            k = min(args.energy_every, args.steps - done)
            t0 = time.perf_counter()
            e.step(scene.dt, scene.G, scene.softening, k)
            e.sync()
            t_step += time.perf_counter() - t0
            done += k
            ke, pe, mom = e.energy(scene.G, scene.softening)
            samples.append({"step": done, "E": ke + pe, "dP": float(np.abs(mom - p0).max())})
        info = e.info()
    drift = max(abs(x["E"] - e0) / abs(e0) for x in samples)
    model, kname = kernel_model(info)
    line = {"metric": f"max relative energy drift over {args.steps} steps", "value": drift, "unit": "|E(t)-E(0)|/|E(0)|", "n_gpus": 1,
            "steps": args.steps, "warmup": 0, "ms_per_step": 1e3 * t_step / args.steps, "higher_is_better": False, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{scene.name} N={scene.n} fp32, G={scene.G:g} eps={scene.softening:g} dt={scene.dt:g} "
                                   f"(BASELINE configs[{scene.meta['config'] - 1}]); energy in fp64 on the device every {args.energy_every} steps",
                       "kernel": kname, "kernel_model": model},
            "E0": e0, "final_rel_energy_error": abs(samples[-1]["E"] - e0) / abs(e0),
            "max_momentum_drift": max(x.get("dP", 0.0) for x in samples), "total_momentum_scale": float(np.abs(p0).max()),
            "g_interactions_per_s": float(scene.n) ** 2 * args.steps / t_step * 1e-9, "samples": samples}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--n", "--bodies", dest="n", type=int, default=1048576)   # --bodies: torchrun's argparse trips over --n
    ap.add_argument("--scene", default="uniform_cube", choices=["uniform_cube", "cold_collapse", "two_galaxies", "plummer"],
                    help="uniform_cube = BASELINE configs[3] (the metric's workload); cold_collapse = configs[4], with --bodies 4194304 --gpus 8")
    ap.add_argument("--general-masses", action="store_true", help="random masses (the reference's recipe, idx:758): the general-mass kernels")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--single-process", action="store_true",
                    help="drive all --gpus devices from THIS process (nbody_create(n_gpus), north_star's N-API host model) instead of torchrun ranks")
    ap.add_argument("--energy", action="store_true", help="energy-drift run (configs[2]): --scene two_galaxies --bodies 131072 --steps 10000")
    ap.add_argument("--energy-every", type=int, default=500)
    ap.add_argument("--ref-sample", type=int, default=16384)
    ap.add_argument("--cpu-sample", type=int, default=16384)
    ap.add_argument("--cpu-reps", type=int, default=3)     # ~3 s each on one host core
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-check", action="store_true")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    single = args.single_process and world == 1
    if world != args.gpus and not single:
        if world == 1 and args.gpus > 1:
            sys.exit("launch with torchrun for --gpus > 1 (one process per GPU), or pass --single-process")

    import torch
    import torch.distributed as dist
    import nbody_b200 as nb

    if args.energy:
        run_energy(args, nb)
        return

    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n_dev = args.gpus if single else world       # devices that share the step

    scene = make_scene(nb, args)
    pm, vel, _ = scene.astype(np.float32)
    if single:
        eng = nb.Engine(args.n, precision="f32", n_gpus=args.gpus)
    else:
        eng = nb.create_rank_engine(args.n, "f32")
    # a real (non-default) stream: the engine (device 0 / the rank's device), the L2 flush and the timing events all run on it
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
        if single:
            eng.sync()

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

    for _ in range(args.warmup):
        flush.zero_()
        eng.step(scene.dt, scene.G, scene.softening, 1)
    barrier()
    eng.enable_timing(True)
    eng.read_timing()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = eng.info()["kernel_launches"]
    # one event pair per step AROUND THE STEP ONLY: the L2 flush between steps is outside every pair; the outer
    # pair (flushes included) is reported next to it
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    o0, o1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    o0.record(stream)
    for a, b in ev:
        flush.zero_()
        a.record(stream)
        eng.step(scene.dt, scene.G, scene.softening, 1)
        b.record(stream)
    o1.record(stream)
    barrier()
    ms_total = sum(a.elapsed_time(b) for a, b in ev)
    ms_outer = o0.elapsed_time(o1)
    launches = eng.info()["kernel_launches"] - l0
    k_launches, k_ms = eng.read_timing()
    eng.enable_timing(False)
    clocks = sampler.stop() if rank == 0 else None
    info = eng.info()       # (which force pass ran)

    # ---- check 1: what the timed loop left behind, against the oracle (sampled bodies, all N partners) ----
    check = None
    if not args.no_check:
        out = eng.download(pos_mass=True, vel=False, acc=True)
        lo, hi = (0, args.n) if single else (info["i_begin"], info["i_end"])
        rng = np.random.default_rng(1234 + rank)
        idx = np.unique(np.concatenate([[lo, hi - 1], rng.integers(lo, hi, 30)])) if hi > lo else np.zeros(0, np.int64)
        err = 0.0
        if len(idx):
            from oracle import oracle
            ref = oracle.acc_subset(out["pos_mass"].astype(np.float64), idx, scene.G, scene.softening)
            got = out["acc"][idx].astype(np.float64)
            err = float(np.linalg.norm(got[:, :3] - ref[:, :3], axis=1).max() / np.linalg.norm(ref[:, :3], axis=1).max())
            if not np.isfinite(err):
                err = float("inf")
        check = {"sampled_acc_err": allmax(err), "gate": ACC_GATE, "sampled_bodies_per_rank": int(len(idx)),
                 "what": "acc of sampled bodies after the timed loop vs the oracle's fp64 sum over all N bodies at the same positions "
                         "(norm-relative, SURVEY.md 8c)"}

    # ---- the same K steps as ONE fused call (positions cross HBM once per step, exchange fused into the epilogue) ----
    eng.upload(pm, vel)
    eng.step(scene.dt, scene.G, scene.softening, 1)      # warm
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    lf0 = eng.info()["kernel_launches"]
    flush.zero_()
    f0.record(stream)
    eng.step(scene.dt, scene.G, scene.softening, args.steps)
    f1.record(stream)
    barrier()
    fused_ms = allmax(f0.elapsed_time(f1))
    fused_launches = allsum(eng.info()["kernel_launches"] - lf0)

    # ---- check 2: a fixed 3-step fused run from the scene, hashed: the same bits on every GPU count ----
    if not args.no_check:
        eng.upload(pm, vel)
        eng.step(scene.dt, scene.G, scene.softening, 3)
        h = state_hash(eng.download(pos_mass=True, vel=False, acc=False)["pos_mass"])
        hs = [h]
        if world > 1:
            hs = [None] * world
            dist.all_gather_object(hs, h)
        check["state_hash"] = hs[0]
        check["state_hash_same_on_all_ranks"] = all(x == hs[0] for x in hs)
        check["state_hash_what"] = "blake2b-64 of pos_mass[n,4] float32 after upload(scene) + nbody_step(dt,G,eps,3)"
        check["ok"] = bool(check["sampled_acc_err"] < ACC_GATE and check["state_hash_same_on_all_ranks"])

    # ---- e2e: host buffers through the public call, copies inside the timed region ----
    if not args.no_e2e:
        lo, hi = (0, args.n) if single else (info["i_begin"], info["i_end"])
        host = {k: torch.empty((args.n, 4), dtype=torch.float32).pin_memory().numpy() for k in ("pos_mass", "vel", "acc")}
        host["pos_mass"][:], host["vel"][:], host["acc"][:] = pm, vel, 0
        e2e_steps = max(1, min(args.steps, 3))
        for it in range(1 + e2e_steps):
            if it == 1:
                barrier()
                t0 = time.perf_counter()
            eng.upload(host["pos_mass"], host["vel"], host["acc"])
            eng.step(scene.dt, scene.G, scene.softening, 1)
            eng.download(pos_mass=True, vel=True, acc=True, out=host, own_only=True)   # every rank reads back the bodies it owns
        barrier()
        e2e_s = allmax((time.perf_counter() - t0) / e2e_steps)
        h2d = allsum(eng.info()["upload_h2d_bytes"])      # counted by the library from the copies it issues
        d2h = allsum(3 * (hi - lo) * 16)

    ms_total = allmax(ms_total)
    ms_outer = allmax(ms_outer)
    launches_all = allsum(launches)
    k_ms_max = allmax(k_ms / max(k_launches, 1))

    if rank == 0:
        pk = peaks()
        inter_step = float(args.n) ** 2
        ms_step = ms_total / args.steps
        value = inter_step / (ms_step * 1e-3) * 1e-9
        sms = info["sm_count"]
        lanes = sms * FP32_LANES_PER_SM * pk["sm_max_mhz"] * 1e6 * 1e-9      # G FP32 lane-ops/s per GPU
        model, kname = kernel_model(info)
        peak_own = lanes / SLOTS[model]
        peak_survey = lanes / SLOTS["survey"]
        # dominant kernel: the force pass; on every GPU one launch processes 1/n_dev of the step's N^2 interactions
        k_inter = inter_step / n_dev
        achieved = k_inter / (k_ms_max * 1e-3) * 1e-9
        traffic = traffic_source = None
        tp = os.path.join(ROOT, "profiles", "ncu_traffic.json")
        if os.path.exists(tp):                      # dram__bytes_read.sum + dram__bytes_write.sum of one ncu --set full capture
            first = json.load(open(tp))
            for t in [first] + list(first.get("also", [])):   # one capture per kernel variant
                if t.get("n") == args.n and t.get("world") == n_dev and t.get("kernel_model", model) == model:
                    traffic = t["dram_bytes_read"] + t["dram_bytes_write"]   # bytes per launch
                    traffic_source = t["source"]
                    break
        alg_bytes = 64.0 * args.n / n_dev + 16.0 * args.n + (48.0 * args.n if model.startswith("pairs_once") else 0.0)
        roof = {"bound": "fp32_issue", "achieved": achieved, "peak": peak_own, "unit": UNIT + " per GPU",
                "frac": achieved / peak_own, "traffic": traffic, "traffic_source": traffic_source,
                "kernel": kname, "kernel_model": model, "kernel_ms": k_ms_max, "kernel_launches": k_launches,
                "issue_slots_per_interaction": SLOTS[model],
                "peak_def": f"{sms} SMs x 128 FP32 lanes x {pk['sm_max_mhz']:.0f} MHz ({pk['src']} sm_max_mhz) / {SLOTS[model]} issue slots per one-sided "
                            f"interaction of THIS kernel ({model}: FP32 lane-ops + MUFU lane-ops it executes per interaction)",
                "peak_survey_R": peak_survey,
                "frac_vs_survey_R": achieved / peak_survey,
                "survey_R_def": "SURVEY.md 8d / BASELINE.md: the same lanes / 13 slots (12 FP32 + 1 MUFU per one-sided interaction); a pairs-once "
                                "kernel executes fewer lane-ops per interaction than that model, so this ratio can exceed 1",
                "hbm": {"algorithmic_bytes_per_launch": alg_bytes,
                        "achieved_gbs": alg_bytes / (k_ms_max * 1e-3) * 1e-9,
                        "peak_gbs": pk["hbm_gbs"], "note": "HBM is negligible for this path (SURVEY.md 8d)"}}
        if clocks and clocks.get("sm_mhz"):
            roof["frac_at_sampled_clock"] = achieved / (peak_own * clocks["sm_mhz"] / pk["sm_max_mhz"])
        exch = {0: "nothing (1 GPU)", 1: "ncclAllGather", 2: "fused peer stores over NVLink + epoch barrier"}[info["exchange"]]
        if model.startswith("pairs_once"):
            par = (f"pairs-once unit table shared by cost x{n_dev}; each GPU accumulates fixed-point partial sums for all bodies, owners add the "
                   f"peers' accumulators ({'peer loads over NVLink' if info['exchange'] == 2 else 'ncclAllReduce int64' if n_dev > 1 else '1 GPU'}); "
                   f"i-range shards x{n_dev} own the integrator; positions exchanged by {exch}")
        else:
            par = f"i-range shards x{n_dev}, positions exchanged by {exch}"
        line = {"metric": metric_name(args), "value": value, "unit": UNIT, "n_gpus": n_dev, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_step, "ms_per_step_with_l2_flush": ms_outer / args.steps, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": f"{scene.name} N={args.n} fp32, G={scene.G:g} eps={scene.softening:g} dt={scene.dt:g} "
                                       f"(BASELINE configs[{scene.meta['config'] - 1}]); "
                                       "one step = integrate(dt) = N^2 interactions",
                           "n": args.n, "parallelism": par, "process_model": "one process, all devices" if single else "one process per GPU (torchrun)",
                           "l2": "flushed between steps (256 MiB write), outside the per-step event pairs",
                           "timed_call": "nbody_step(dt,G,eps,1) per step: kick1_drift + force pass (+ finish) in mode kick-2",
                           "pairs_once_tiles_per_unit": info.get("f32_symmetric", 0), "bodies_per_thread": info.get("f32_sym_bodies_per_thread", 0),
                           "pairs_once_warp_units": bool(info.get("f32_sym_warp_units", 0)),
                           "flop_convention": "20 flop/interaction", "tflops_20": value * 20e-3,
                           "equal_mass_kernel": bool(info.get("uniform_mass"))},
                "roofline": roof, "gpu_launches": int(launches_all), "clocks": clocks,
                "fused": {"value": inter_step * args.steps / (fused_ms * 1e-3) * 1e-9, "unit": UNIT, "ms_per_step": fused_ms / args.steps,
                          "gpu_launches": int(fused_launches),
                          "what": f"ONE nbody_step(dt,G,eps,{args.steps}) call: every step but the last runs the epilogue in mode kick-2 + next kick-1 + "
                                  "drift + position exchange; no L2 flush between its steps"}}
        if check is not None:
            line["check"] = check
        if not args.no_e2e:
            line["e2e"] = {"value": inter_step / e2e_s * 1e-9, "unit": UNIT, "h2d_bytes_per_step": int(h2d),
                           "d2h_bytes_per_step": int(d2h), "ms_per_step": e2e_s * 1e3,
                           "api": "Engine.upload(host) -> Engine.step(dt,G,eps,1) -> Engine.download(host, own_only) (pinned host arrays; every rank uploads "
                                  "and reads back the slice it owns, positions reach the peers over NVLink)"}
        if not args.no_cpu and n_dev == 1:
            times, core = cpu_step_times(scene, min(args.n, args.cpu_sample), reps=args.cpu_reps)
            line["cpu_baseline"] = cpu_record(times, min(args.n, args.cpu_sample), core)
        print(json.dumps(line), flush=True)

    eng.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if check is not None and not check.get("ok", True):
        sys.exit(f"bench.py: check failed: {check}")


if __name__ == "__main__":
    main()

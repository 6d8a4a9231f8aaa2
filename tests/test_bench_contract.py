"""bench.py's reference arm (`--impl reference`) keeps the driver's JSON contract: one line, BASELINE.json's metric and
unit, `impl`, a `cpu_baseline` describing the run and an `e2e` that repeats the value with zero copies; under torchrun
only rank 0 works and prints, the other ranks exit 0 silently.  (The GPU arm needs a device: tests/test_gpu_*.py and the
driver's own run cover it.)"""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(env_extra, *args):
    env = dict(os.environ, **env_extra)
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "2", "--warmup", "1",
                           "--ref-sample", "1024", *args], capture_output=True, text=True, env=env, cwd=ROOT, timeout=300)


def test_reference_arm_line():
    r = _run({})
    assert r.returncode == 0, r.stderr
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    base = json.load(open(os.path.join(ROOT, "BASELINE.json")))
    assert d["impl"] == "reference" and d["higher_is_better"] is True and d["vs_baseline"] is None
    assert d["unit"] == "G interactions/s" and "N=1M" in d["metric"] and "interactions/s" in base["metric"]
    assert d["steps"] == 2 and d["warmup"] == 1 and d["n_gpus"] == 1 and d["value"] > 0 and d["ms_per_step"] > 0
    assert d["config"]["n"] == 1048576 and d["config"]["sample_n"] == 1024 and "uniform_cube" in d["config"]["workload"]
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] == 1 and cb["value"] == d["value"] and "1024 bodies" in cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    # value = N(N-1) one-sided interactions per step / time
    assert abs(d["value"] - 1024 * 1023 / (d["ms_per_step"] * 1e-3) * 1e-9) < 1e-9 * d["value"] + 1e-12


def test_reference_arm_other_ranks_stay_silent():
    r = _run({"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"}, "--gpus", "2")
    assert r.returncode == 0 and r.stdout.strip() == ""


@pytest.mark.gpu
def test_gpu_arm_line_reduced_size():
    """The product arm at a reduced body count (the default run is the driver's job): every key of the contract is there
    and consistent -- value = n^2 / time, roofline achieved/peak/frac, e2e through host arrays with the bytes counted,
    launches counted, clocks sampled, the CPU baseline labelled."""
    n = 131072
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--bodies", str(n), "--steps", "3", "--warmup", "3", "--cpu-sample", "2048",
                        "--cpu-reps", "2"], capture_output=True, text=True, cwd=ROOT, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip().startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["unit"] == "G interactions/s" and d["n_gpus"] == 1 and d["steps"] == 3 and d["warmup"] == 3 and d["dtype"] == "f32"
    assert d["higher_is_better"] is True and d["vs_baseline"] is None and d["data"] == "synthetic" and "workload" in d["config"]
    assert abs(d["value"] - n * n / (d["ms_per_step"] * 1e-3) * 1e-9) < 1e-6 * d["value"]
    roof = d["roofline"]
    assert abs(roof["frac"] - roof["achieved"] / roof["peak"]) < 1e-9 and 0.5 < roof["frac"] < 1.05 and roof["kernel_launches"] == 3
    assert roof["kernel_model"].startswith("pairs_once") and abs(roof["frac_vs_survey_R"] - roof["achieved"] / roof["peak_survey_R"]) < 1e-9
    assert d["gpu_launches"] >= 2 * 3
    # the line checks what it timed: sampled accelerations against the oracle, a state hash, and the fused call
    ck = d["check"]
    assert ck["ok"] is True and ck["sampled_acc_err"] < ck["gate"] == 2e-5 and len(ck["state_hash"]) == 16 and ck["state_hash_same_on_all_ranks"]
    assert d["fused"]["value"] > 0.8 * d["value"] and d["fused"]["gpu_launches"] >= 3
    assert d["ms_per_step_with_l2_flush"] >= d["ms_per_step"]
    cb = d["cpu_baseline"]
    assert cb["repetitions"] == 2 and cb["min"] <= cb["median"] <= cb["max"] == cb["value"]
    e = d["e2e"]
    assert e["h2d_bytes_per_step"] == e["d2h_bytes_per_step"] == 3 * n * 16 and 0 < e["value"] <= d["value"] * 1.02
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] == 1 and d["cpu_baseline"]["value"] > 0
    assert set(d["clocks"]) >= {"sm_mhz", "sm_max_mhz", "reasons"}

"""The N-API addon (n-body-sim_b200/addon/nbody_napi.cc) executed end to end.
Node.js is absent from this image, so tests/napi_mock_host.cc stands in for the
Node process: it provides the napi_* functions, loads nbody_napi.node and drives
it the way addon/physics_shim.js does.  The results must equal the oracle's."""
import json
import os
import subprocess

import numpy as np
import pytest

from .conftest import ROOT, norm_rel_err

ADDON = os.path.join(ROOT, "n-body-sim_b200", "addon", "nbody_napi.node")
HOST = os.path.join(ROOT, "tests", "_build", "napi_mock_host")


def _fixture_bytes(s, acc, temp):
    n = s.n
    return np.concatenate([[n, s.dt, s.G, s.softening], s.pos_mass.ravel(), s.vel.ravel(), acc.ravel(), temp]).astype(np.float64)


def _split(buf, n):
    return (buf[4:4 + 4 * n].reshape(n, 4), buf[4 + 4 * n:4 + 8 * n].reshape(n, 4),
            buf[4 + 8 * n:4 + 12 * n].reshape(n, 4), buf[4 + 12 * n:4 + 13 * n])


def test_addon_exports_and_imports():
    """CPU-side: the addon exports Node's entry symbol and imports only declared napi_* names."""
    assert os.path.exists(ADDON), "run __graft_entry__.build() first"
    syms = subprocess.run(["nm", "-D", ADDON], capture_output=True, text=True, check=True).stdout
    assert " T napi_register_module_v1" in syms
    hdr = open(os.path.join(ROOT, "n-body-sim_b200", "addon", "napi_min.h")).read()
    for line in syms.splitlines():
        parts = line.split()
        if len(parts) == 2 and parts[0] == "U" and parts[1].startswith("napi_"):
            assert parts[1] + "(" in hdr, parts[1]


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["step", "split", "host"])
@pytest.mark.parametrize("prec,tol", [("f64", 1e-12), ("f32", 2e-5)])
def test_addon_matches_oracle(nb, oracle, tmp_path, mode, prec, tol):
    s = nb.scenes.galaxy_default()                    # the reference's own default scene
    acc0 = np.zeros_like(s.vel)
    fin, fout = tmp_path / "in.bin", tmp_path / "out.bin"
    _fixture_bytes(s, acc0, s.temp).tofile(fin)
    nsteps = 2                                        # animate(): subSteps = 2, idx:877-879
    r = subprocess.run([HOST, ADDON, str(fin), str(fout), str(nsteps), mode, prec], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    meta = json.loads(r.stdout.strip().splitlines()[-1])
    assert meta["n"] == s.n and meta["kernel_launches"] > 0
    pm, vel, acc, temp = _split(np.fromfile(fout, dtype=np.float64), s.n)
    p, v, a, t = oracle.integrate(s.pos_mass, s.vel, acc0, s.temp, s.dt, s.G, s.softening, nsteps=nsteps)
    assert norm_rel_err(pm, p) < tol and norm_rel_err(vel, v) < tol and norm_rel_err(acc, a) < tol
    if prec == "f64":
        assert np.array_equal(temp, t)
        ke, pe, _ = oracle.energy(p, v, s.G, s.softening)
        assert abs(meta["kinetic"] - ke) < 1e-10 * abs(ke) and abs(meta["potential"] - pe) < 1e-10 * abs(pe)


@pytest.mark.gpu
def test_addon_world_entry_points(nb, oracle, tmp_path):
    """worldEnable / worldSetBodies / worldStep / worldSync through the addon glue: the complete step on the device with
    radii so small that nothing touches equals the gravity-only contract; wrong calls throw."""
    s = nb.scenes.galaxy_default()
    acc0 = np.zeros_like(s.vel)
    fin, fout = tmp_path / "in.bin", tmp_path / "out.bin"
    _fixture_bytes(s, acc0, s.temp).tofile(fin)
    r = subprocess.run([HOST, ADDON, str(fin), str(fout), "4", "world", "f64"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    pm, vel, acc, temp = _split(np.fromfile(fout, dtype=np.float64), s.n)
    p, v, a, t = oracle.integrate(s.pos_mass, s.vel, acc0, s.temp, s.dt, s.G, s.softening, nsteps=4)
    assert norm_rel_err(pm, p) < 1e-12 and norm_rel_err(vel, v) < 1e-12 and norm_rel_err(acc, a) < 1e-12
    assert np.array_equal(temp, t)


@pytest.mark.gpu
@pytest.mark.parametrize("prec", ["f64", "f32"])
def test_addon_collision_and_renderer_feed(nb, oracle, tmp_path, prec):
    """setRadius / contacts / history / trailRibbons / snapshot through the addon glue."""
    s = nb.scenes.galaxy_default()
    acc0 = np.zeros_like(s.vel)
    fin, fout = tmp_path / "in.bin", tmp_path / "out.bin"
    _fixture_bytes(s, acc0, s.temp).tofile(fin)
    nsteps, depth = 12, 8
    r = subprocess.run([HOST, ADDON, str(fin), str(fout), str(nsteps), "collide", prec], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    dt = np.float32 if prec == "f32" else np.float64
    pm, vel, acc, temp = _split(np.fromfile(fout, dtype=np.float64), s.n)
    snap = nb.snapshot.read(str(fout) + ".nbs")
    radius = snap["radius"]                      # the radii the host computed (its cbrt may differ from numpy's by an ulp)
    np.testing.assert_allclose(radius, 1.5 * np.cbrt(3.0 * s.pos_mass[:, 3].astype(dt).astype(np.float64) / (4.0 * np.pi)), rtol=1e-6)
    pairs = np.fromfile(str(fout) + ".pairs", dtype=np.uint32).reshape(-1, 2)
    ref = oracle.overlap_pairs(pm, radius.astype(np.float64), 0.0, mt=False)   # kick-2 does not move bodies
    assert len(ref) > 0 and np.array_equal(pairs, ref)
    assert snap["step"] == nsteps and snap["dtype"] == np.dtype(dt)
    assert np.array_equal(snap["pos_mass"].astype(np.float64), pm) and np.array_equal(snap["vel"].astype(np.float64), vel)
    hist = np.fromfile(str(fout) + ".history", dtype=np.float64).reshape(s.n, depth, 3)
    assert np.array_equal(hist[:, 0], pm[:, :3])                               # newest sample = current position
    rib = np.fromfile(str(fout) + ".ribbons", dtype=np.float32).reshape(s.n, depth, 6)
    want = oracle.trail_ribbons(hist, radius.astype(np.float64), (0, 80, 150))
    if prec == "f64":
        assert np.array_equal(rib, want)
    else:
        assert np.abs(rib - want).max() < 2e-3 * float(radius.max())
    frame = np.fromfile(str(fout) + ".frame", dtype=np.float32).reshape(54, 96, 4)
    k = np.arange(s.n)
    ref, _ = oracle.render(pm, radius, temp, (k % 3).astype(np.int32), k.astype(np.float32) * np.float32(0.37), 96, 54)
    assert np.array_equal(frame, ref) and (frame[..., 3] > 0).sum() > 50

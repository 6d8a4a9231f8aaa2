"""CPU tests for the renderer feed and snapshot format: the
NBODYSN1 wire format (host module) and pins for the oracle's restatement of the
ribbon-trail loop (idx:563-592; the reference has no tests, so geometric known answers)."""
import struct

import numpy as np
import pytest


def test_snapshot_roundtrip_and_layout(nb, tmp_path):
    s = nb.scenes.galaxy_default()
    rad = np.linspace(0.5, 3.0, s.n)
    for dtype in (np.float64, np.float32):
        p = str(tmp_path / f"g_{np.dtype(dtype).name}.nbs")
        nb.snapshot.write(p, s.pos_mass, s.vel, None, s.temp, rad, step=42, dt=s.dt, G=s.G, softening=s.softening, dtype=dtype)
        raw = open(p, "rb").read()
        esz = np.dtype(dtype).itemsize
        assert len(raw) == 64 + (12 * s.n + 2 * s.n) * esz
        assert raw[:8] == b"NBODYSN1"
        assert struct.unpack_from("<QQ", raw, 8) == (s.n, 42)
        assert struct.unpack_from("<ddd", raw, 24) == (s.dt, s.G, s.softening)
        assert struct.unpack_from("<II", raw, 48) == (0 if dtype == np.float32 else 1, 3)
        # first array starts right after the header, in the host layout x,y,z,m
        np.testing.assert_array_equal(np.frombuffer(raw, dtype, 4, 64), s.pos_mass[0].astype(dtype))
        d = nb.snapshot.read(p)
        assert (d["n"], d["step"], d["dt"], d["G"], d["softening"]) == (s.n, 42, s.dt, s.G, s.softening)
        np.testing.assert_array_equal(d["pos_mass"], s.pos_mass.astype(dtype))
        np.testing.assert_array_equal(d["vel"], s.vel.astype(dtype))
        assert not d["acc"].any()
        np.testing.assert_array_equal(d["temp"], s.temp.astype(dtype))
        np.testing.assert_array_equal(d["radius"], rad.astype(dtype))


def test_snapshot_optional_arrays_and_malformed(nb, tmp_path):
    s = nb.scenes.galaxy_default()
    p = str(tmp_path / "a.nbs")
    nb.snapshot.write(p, s.pos_mass, s.vel, s.vel * 2)
    d = nb.snapshot.read(p)
    assert d["temp"] is None and d["radius"] is None
    np.testing.assert_array_equal(d["acc"], s.vel * 2)
    raw = open(p, "rb").read()
    for bad in (raw[:-8], raw + b"\0", b"XBODYSN1" + raw[8:], raw[:32]):
        q = str(tmp_path / "bad.nbs")
        open(q, "wb").write(bad)
        with pytest.raises(ValueError):
            nb.snapshot.read(q)
    with pytest.raises(ValueError):
        nb.snapshot.write(p, s.pos_mass, s.vel[:5])
    with pytest.raises(ValueError):
        nb.snapshot.write(p, s.pos_mass.astype(np.float16), s.vel)


def test_history_push_semantics(oracle):
    pos = np.arange(12.0).reshape(3, 4)
    h = oracle.history_init(pos, depth=5)
    assert h.shape == (3, 5, 3) and (h == pos[:, None, :3]).all()            # idx:540
    for k in range(1, 8):
        h = oracle.history_push(h, pos + k)
    # newest first, the oldest dropped (idx:564-565)
    np.testing.assert_array_equal(h[:, :, 0], pos[:, None, 0] + np.array([7, 6, 5, 4, 3.0]))


def test_ribbon_geometry_known_answers(oracle):
    rng = np.random.default_rng(2)
    n, depth = 6, 80
    start = rng.uniform(-50, 50, (n, 3))
    vel = rng.normal(0, 1.0, (n, 3))
    hist = start[:, None, :] - vel[:, None, :] * np.arange(depth)[None, :, None] * 0.1   # straight-line motion
    radius = rng.uniform(0.5, 3.0, n)
    cam = np.array([0.0, 80.0, 150.0])                                                   # camera.position, idx:705
    rib = oracle.trail_ribbons(hist, radius, cam).astype(np.float64)
    a, b = rib[:, :, :3], rib[:, :, 3:]
    mid, perp = 0.5 * (a + b), 0.5 * (a - b)
    np.testing.assert_allclose(mid, hist, rtol=0, atol=2e-5)                              # vertex pair straddles history[i]
    width = radius[:, None] * 0.8 * np.maximum(0, 1 - np.arange(depth) / (depth - 1))    # idx:569-570
    np.testing.assert_allclose(np.linalg.norm(perp, axis=2), width, rtol=0, atol=3e-5)
    assert (np.linalg.norm(perp[:, -1], axis=1) < 1e-6).all()                             # the tail tapers to zero
    # perpendicular to the direction of motion and to the view direction (camera-facing ribbon)
    u = perp[:, :-1] / np.linalg.norm(perp[:, :-1], axis=2, keepdims=True)
    d = vel / np.linalg.norm(vel, axis=1, keepdims=True)
    assert np.abs((u * d[:, None, :]).sum(2)).max() < 1e-4
    to_cam = cam - hist[:, :-1]
    to_cam /= np.linalg.norm(to_cam, axis=2, keepdims=True)
    assert np.abs((u * to_cam).sum(2)).max() < 1e-4


def test_ribbon_degenerate_history_has_no_nan(oracle):
    hist = oracle.history_init(np.array([[1.0, 2.0, 3.0, 0.0]]), depth=80)                # a body just created
    rib = oracle.trail_ribbons(hist, [2.0], [0, 80, 150])
    assert np.isfinite(rib).all()
    np.testing.assert_array_equal(rib[0, :, :3], np.float32([[1, 2, 3]] * 80))            # length() || 1: zero width

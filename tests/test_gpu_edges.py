"""Edge cases of the rows next to the hot path through the C ABI: smallest
and ragged body counts, extreme ring depths, degenerate geometry, multi-GPU gathers."""
import os

import numpy as np
import pytest

from .conftest import norm_rel_err

pytestmark = pytest.mark.gpu


def _rad(m):
    return (3 * np.asarray(m, np.float64) / (4 * np.pi)) ** (1 / 3)


@pytest.mark.parametrize("precision", ["f32", "f64"])
@pytest.mark.parametrize("n", [1, 2, 33, 513])
def test_tiny_and_ragged_counts(nb, oracle, precision, n):
    """n = 1 (no pair exists), n = 2, and counts that fill neither a CTA nor a tile."""
    dt = np.float32 if precision == "f32" else np.float64
    rng = np.random.default_rng(n)
    pm = np.concatenate([rng.uniform(-3, 3, (n, 3)), rng.uniform(0.5, 2.0, (n, 1))], axis=1).astype(dt)
    vel = np.zeros((n, 4), dt)
    rad = _rad(pm[:, 3]).astype(dt)
    with nb.Engine(n, precision=precision) as e:
        e.upload(pm, vel, None, np.full(n, 60.0, dt))
        e.set_radius(rad)
        e.force(0.5, 0.5)
        got = e.contacts(0.1)
        ref = oracle.overlap_pairs(pm.astype(np.float64), rad.astype(np.float64), 0.1, mt=False)
        assert np.array_equal(got, ref)
        if n == 1:
            assert len(got) == 0
        e.history_enable(2)                                   # the smallest ring
        e.history_push()
        e.kick1_drift(0.01)
        e.history_push()
        out = e.download(temp=True)
        h = e.history()
        assert h.shape == (n, 2, 3) and np.array_equal(h[:, 0], out["pos_mass"][:, :3]) and np.array_equal(h[:, 1], pm[:, :3])
        rib = e.trail_ribbons((0, 80, 150))
        want = oracle.trail_ribbons(h.astype(np.float64), rad.astype(np.float64), (0, 80, 150))
        if precision == "f64":
            assert np.array_equal(rib, want)
        else:
            assert np.abs(rib - want).max() < 2e-3 * float(rad.max())
        img, ids = e.render(80, 45, ids=True)
    ref_img, ref_ids = oracle.render(out["pos_mass"], rad, out["temp"], np.zeros(n, np.int32), np.zeros(n), 80, 45)
    assert np.array_equal(ids, ref_ids) and np.array_equal(img, ref_img)


def test_deep_ring_and_wraparound(nb, oracle):
    """depth 256 (the maximum) in fp64: 128 KB of shared memory per CTA, three wrap-arounds."""
    s = nb.scenes.galaxy_default()
    rad = _rad(s.pos_mass[:, 3])
    with nb.Engine(s.n, precision="f64") as e:
        for bad in (1, 257, -3):
            with pytest.raises(nb.NBodyError):
                e.history_enable(bad)
        e.upload(s.pos_mass, s.vel)
        e.set_radius(rad)
        e.history_enable(256)
        for k in range(600):
            e.kick1_drift(0.004)
            e.history_push()
        h = e.history()
        # acc = 0 throughout (no force pass): every push is the previous position plus vel*dt, rounded step by step
        p = s.pos_mass[:, :3].copy()
        track = []
        for k in range(600):
            p = p + s.vel[:, :3] * 0.004
            track.append(p.copy())
        want = np.stack(track[::-1][:256], axis=1)
        assert np.array_equal(h, want)
        assert np.array_equal(e.trail_ribbons((5, 6, 7)), oracle.trail_ribbons(want, rad, (5, 6, 7)))
        e.history_enable(0)                                   # frees the ring
        with pytest.raises(nb.NBodyError):
            e.history_push()


def test_coincident_and_zero_radius_bodies(nb, oracle):
    """Coincident centres (distSq = 0), a zero radius, touching exactly: the strict < of idx:313 decides."""
    pm = np.array([[0, 0, 0, 1.0], [0, 0, 0, 1.0], [3, 0, 0, 1.0], [5, 0, 0, 1.0], [9, 0, 0, 1.0]])
    rad = np.array([1.0, 0.0, 1.0, 1.0, 0.0])              # bodies 2 and 3 touch exactly: 2 == 1 + 1 -> no contact
    with nb.Engine(5, precision="f64") as e:
        e.upload(pm, np.zeros((5, 4)))
        e.set_radius(rad)
        got = e.contacts(0.0)
        assert np.array_equal(got, oracle.overlap_pairs(pm, rad, 0.0, mt=False))
        assert got.tolist() == [[0, 1]]
        assert e.contacts(1e-9).tolist() == [[0, 1], [2, 3]]
        with pytest.raises(nb.NBodyError):
            e.set_radius(np.array([1.0, -1.0, 1.0, 1.0, 1.0]))
        img, ids = e.render(64, 36, ids=True)                 # zero-radius bodies are never drawn
        assert not np.isin(ids, [1, 4]).any() and np.isfinite(img).all()


def test_multi_gpu_render_and_trails(nb, oracle):
    """Two devices: temperatures are sharded and gathered for the frame; positions are everywhere."""
    if nb.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    s = nb.scenes.galaxy_default()
    pm, vel, _ = s.astype(np.float32)
    rad = (_rad(pm[:, 3]) * 3).astype(np.float32)
    temp = np.linspace(0, 300, s.n).astype(np.float32)
    out = {}
    os.environ["NBODY_SMALL"] = "0"      # one GPU would take the one-cluster kernel (small.cuh): same step, other sum order
    try:
        for g in (1, 2):
            with nb.Engine(s.n, precision="f32", n_gpus=g) as e:
                e.upload(pm, vel, None, temp)
                e.set_radius(rad)
                e.history_enable(8)
                for _ in range(3):
                    e.step(s.dt, s.G, s.softening, 2)
                    e.history_push()
                out[g] = (e.render(160, 90), e.history(), e.trail_ribbons((0, 80, 150)), e.download(temp=True))
    finally:
        del os.environ["NBODY_SMALL"]
    for a, b in zip(out[1][:3], out[2][:3]):
        assert np.array_equal(a, b)
    d = out[2][3]
    ref, _ = oracle.render(d["pos_mass"], rad, d["temp"], np.zeros(s.n, np.int32), np.zeros(s.n), 160, 90)
    assert np.array_equal(out[2][0], ref)


# ---- the one-cluster multi-step kernel for the reference's own sizes (csrc/small.cuh) ----------

@pytest.mark.parametrize("precision,tol", [("f64", 1e-12), ("f32", 3e-5)])
@pytest.mark.parametrize("n", [1, 2, 7, 8, 9, 151, 300, 512, 513])
def test_small_system_kernel(nb, oracle, precision, tol, n):
    """N <= 512 on one GPU: nbody_step is one launch of one thread-block cluster that runs all the
    steps (N = 513 takes the cooperative all-SM kernel, medium.cuh).  Against the oracle's integrate() and against
    the launch-per-step path, cluster shares from 1 body per CTA (and idle CTAs) to 64; thermal decay
    bit-exact; a second call continues from the stored acc."""
    dt = np.float64 if precision == "f64" else np.float32
    rng = np.random.default_rng(100 + n)
    pm = np.concatenate([rng.normal(0, 4, (n, 3)), rng.uniform(0.5, 2.5, (n, 1))], axis=1).astype(dt)
    vel = np.zeros((n, 4), dt)
    vel[:, :3] = rng.normal(0, 0.5, (n, 3))
    temp = rng.uniform(0, 50, n).astype(dt)
    G, eps, h = 0.5, 0.5, 0.008

    def run(small):
        os.environ["NBODY_SMALL"] = "1" if small else "0"
        try:
            with nb.Engine(600, precision=precision) as e:
                e.upload(pm, vel, None, temp)
                l0 = e.info()["kernel_launches"]
                e.step(h, G, eps, 5)
                launches = e.info()["kernel_launches"] - l0
                e.step(h, G, eps, 1)
                return e.download(temp=True), launches
        finally:
            del os.environ["NBODY_SMALL"]

    out, launches = run(True)
    assert launches == 1                                      # (513 bodies: the all-SM form of the same idea, medium.cuh)
    p, v, a, t = oracle.integrate(pm, vel, np.zeros_like(vel), temp, h, G, eps, nsteps=6)
    scale = max(1.0, np.abs(p[:, :3]).max())
    assert np.abs(out["pos_mass"][:, :3] - p[:, :3]).max() < tol * scale
    assert np.abs(out["vel"][:, :3] - v[:, :3]).max() < 10 * tol * max(1.0, np.abs(v).max())
    assert np.abs(out["acc"][:, :3] - a[:, :3]).max() < 10 * tol * max(1e-3, np.abs(a).max())
    assert np.array_equal(out["pos_mass"][:, 3], pm[:, 3])
    if precision == "f64":
        np.testing.assert_array_equal(out["temp"], t)
    ref, _ = run(False)
    for k in ("pos_mass", "vel", "acc"):
        assert np.abs(out[k] - ref[k]).max() < 10 * tol * max(1.0, np.abs(ref[k]).max()), k
    np.testing.assert_array_equal(out["temp"], ref["temp"])


@pytest.mark.parametrize("precision", ["f64", "f32"])
def test_small_system_kernel_zero_softening(nb, oracle, precision):
    """eps = 0: the self pair is selected away (fp64) or clamped (fp32), no NaN."""
    dt = np.float64 if precision == "f64" else np.float32
    s = nb.scenes.galaxy_default()
    pm, vel, _ = s.astype(dt)
    with nb.Engine(s.n, precision=precision) as e:
        e.upload(pm, vel)
        e.step(0.002, s.G, 0.0, 3)
        out = e.download()
    p, v, a, _ = oracle.integrate(pm, vel, np.zeros_like(vel), None, 0.002, s.G, 0.0, nsteps=3)
    assert np.isfinite(out["acc"]).all()
    tol = 1e-12 if precision == "f64" else 2e-5
    assert norm_rel_err(out["pos_mass"], p) < tol and norm_rel_err(out["acc"], a) < 10 * tol


# ---- 513 .. 4,096 bodies: all steps of a call in one cooperative launch over all SMs (csrc/medium.cuh) ----------

@pytest.mark.parametrize("precision,tol", [("f64", 1e-12), ("f32", 3e-5)])
@pytest.mark.parametrize("n,eps", [(513, 0.5), (600, 0.0), (1000, 0.05), (2049, 0.5), (4096, 0.05)])
def test_medium_system_kernel(nb, oracle, precision, tol, n, eps):
    """One launch runs every step of the call, positions exchanged through the context's two position
    buffers with one grid barrier per step.  Against the oracle's integrate() and against the
    launch-per-step path (NBODY_MEDIUM=0); odd and even step counts (the current buffer flips), a second
    call continues from the stored acc, thermal decay bit-exact, a force pass afterwards reads the
    right buffer, two runs agree bit for bit."""
    dt = np.float64 if precision == "f64" else np.float32
    rng = np.random.default_rng(200 + n)
    pm = np.concatenate([rng.normal(0, 6, (n, 3)), rng.uniform(0.5, 2.5, (n, 1))], axis=1).astype(dt)
    vel = np.zeros((n, 4), dt)
    vel[:, :3] = rng.normal(0, 0.3, (n, 3))
    temp = rng.uniform(0, 50, n).astype(dt)
    G, h = 0.5, 0.004

    def run(medium, cap=n):
        os.environ["NBODY_MEDIUM"] = "1" if medium else "0"
        try:
            with nb.Engine(cap, precision=precision) as e:
                e.upload(pm[:cap] if cap < n else pm, vel, None, temp)
                l0 = e.info()["kernel_launches"]
                e.step(h, G, eps, 3)
                launches = e.info()["kernel_launches"] - l0
                e.step(h, G, eps, 2)
                out = e.download(temp=True)
                e.force(G, eps)
                return out, launches, e.download()["acc"]
        finally:
            del os.environ["NBODY_MEDIUM"]

    out, launches, acc_again = run(True)
    assert launches == 1
    p, v, a, t = oracle.integrate(pm, vel, np.zeros_like(vel), temp, h, G, eps, nsteps=5)
    scale = max(1.0, np.abs(p[:, :3]).max())
    assert np.abs(out["pos_mass"][:, :3] - p[:, :3]).max() < tol * scale
    assert np.abs(out["vel"][:, :3] - v[:, :3]).max() < 10 * tol * max(1.0, np.abs(v).max())
    assert np.abs(out["acc"][:, :3] - a[:, :3]).max() < 10 * tol * max(1e-3, np.abs(a).max())
    assert np.array_equal(out["pos_mass"][:, 3], pm[:, 3])
    if precision == "f64":
        np.testing.assert_array_equal(out["temp"], t)
    assert np.abs(acc_again[:, :3] - a[:, :3]).max() < 10 * tol * max(1e-3, np.abs(a).max())   # same positions, other kernel
    ref, ref_launches, _ = run(False)
    assert ref_launches >= 4                                   # kick1_drift + 3 force passes (two kernels each in the fp64 pairs-once pass)
    for k in ("pos_mass", "vel", "acc"):
        assert np.abs(out[k] - ref[k]).max() < 10 * tol * max(1.0, np.abs(ref[k]).max()), k
    np.testing.assert_array_equal(out["temp"], ref["temp"])
    again, _, _ = run(True)
    for k in ("pos_mass", "vel", "acc", "temp"):
        np.testing.assert_array_equal(out[k], again[k])


@pytest.mark.parametrize("precision,n_max", [("f32", 5120), ("f64", 4096)])
def test_medium_system_kernel_upper_bound(nb, oracle, precision, n_max):
    """The largest system of each precision against the oracle; one body more takes the launch-per-step kernels in fp64 and the
    cooperative pairs-once launch in fp32 (csrc/sym_coop.cuh: one launch as well, plus the memset-free clean-up)."""
    dt = np.float32 if precision == "f32" else np.float64
    for n in (n_max, n_max + 1):
        s = nb.scenes.plummer(n)
        pm, vel, _ = s.astype(dt)
        with nb.Engine(s.n, precision=precision) as e:
            e.upload(pm, vel)
            l0 = e.info()["kernel_launches"]
            e.step(s.dt, s.G, s.softening, 3)
            launches = e.info()["kernel_launches"] - l0
            out = e.download()
        assert (launches == 1) if n == n_max or precision == "f32" else (launches >= 4)
        if n == n_max or precision == "f32":
            p, v, a, _ = oracle.integrate(pm, vel, np.zeros_like(vel), None, s.dt, s.G, s.softening, nsteps=3)
            tol = 1e-12 if precision == "f64" else 1e-6
            assert norm_rel_err(out["pos_mass"], p) < tol and norm_rel_err(out["acc"], a) < (1e-12 if precision == "f64" else 2e-5)

"""Small-system I/O through the C ABI (csrc/integrate.cuh pack_state/unpack_state; nbody_step_host): systems of up to
4,096 bodies on one device travel as ONE pinned record each way, the upload returns without a synchronise (the equal-mass
detection is read when first needed), and nbody_step_host is the JS shim's frame -- arrays in, nsteps x integrate(dt)
(idx:247-262, 877-879), arrays out -- in one call.  Everything must equal the separate-copies path bit for bit."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _state(nb, n, dtype, seed=0):
    s = nb.scenes.galaxy_default() if n == 151 else nb.scenes.plummer(n)
    pm, vel, temp = s.astype(dtype)
    rng = np.random.default_rng(seed)
    acc = rng.normal(0, 1e-3, (n, 4)).astype(dtype)
    acc[:, 3] = 0
    temp = rng.uniform(0, 200, n).astype(dtype)
    return s, pm, vel, acc, temp


def _engine(nb, n, prec, small_io):
    old = os.environ.get("NBODY_SMALL_IO")
    os.environ["NBODY_SMALL_IO"] = "1" if small_io else "0"
    try:
        return nb.Engine(n, precision=prec)
    finally:
        if old is None:
            del os.environ["NBODY_SMALL_IO"]
        else:
            os.environ["NBODY_SMALL_IO"] = old


@pytest.mark.parametrize("prec", ["f64", "f32"])
@pytest.mark.parametrize("n", [1, 2, 151, 300, 513, 4096, 4097])
def test_upload_download_roundtrip_both_paths(nb, prec, n):
    dtype = np.float64 if prec == "f64" else np.float32
    _, pm, vel, acc, temp = _state(nb, n, dtype)
    for small_io in (True, False):
        with _engine(nb, n, prec, small_io) as e:
            e.upload(pm, vel, acc, temp)
            out = e.download(temp=True)
            for k, x in (("pos_mass", pm), ("vel", vel), ("acc", acc), ("temp", temp)):
                assert np.array_equal(out[k], x), (small_io, k)
            e.upload(pm, vel)                                   # acc = 0 like a new Body (idx:217), no temperatures
            out = e.download(temp=True)
            assert np.array_equal(out["pos_mass"], pm) and np.array_equal(out["vel"], vel)
            assert not out["acc"].any() and not out["temp"].any()
            part = e.download(pos_mass=False, vel=True, acc=False)
            assert set(part) == {"vel"} and np.array_equal(part["vel"], vel)


@pytest.mark.parametrize("prec", ["f64", "f32"])
@pytest.mark.parametrize("n,nsteps", [(151, 2), (151, 7), (700, 2), (3000, 3)])
def test_steps_identical_with_and_without_small_io(nb, prec, n, nsteps):
    """The cluster kernel (n <= 512), the all-SM kernel and the launch-per-step kernels all start from the same device
    arrays whichever way they were filled."""
    dtype = np.float64 if prec == "f64" else np.float32
    s, pm, vel, acc, temp = _state(nb, n, dtype)
    outs = []
    for small_io in (True, False):
        with _engine(nb, n, prec, small_io) as e:
            e.upload(pm, vel, acc, temp)
            e.step(s.dt, s.G, s.softening, nsteps)
            outs.append(e.download(temp=True))
    for k in outs[0]:
        assert np.array_equal(outs[0][k], outs[1][k]), k


@pytest.mark.parametrize("prec", ["f64", "f32"])
@pytest.mark.parametrize("n", [151, 300, 2048, 5000])
def test_step_host_equals_three_calls(nb, prec, n):
    dtype = np.float64 if prec == "f64" else np.float32
    s, pm, vel, acc, temp = _state(nb, n, dtype)
    with nb.Engine(n, precision=prec) as e:
        e.upload(pm, vel, acc, temp)
        e.step(s.dt, s.G, s.softening, 2)
        ref = e.download(temp=True)
    with nb.Engine(n, precision=prec) as e:
        a = [x.copy() for x in (pm, vel, acc, temp)]
        e.step_host(*a, s.dt, s.G, s.softening, 2)
        for x, k in zip(a, ("pos_mass", "vel", "acc", "temp")):
            assert np.array_equal(x, ref[k]), k
        dev = e.download(temp=True)                            # the state stayed on the device as well
        for k in ref:
            assert np.array_equal(dev[k], ref[k]), k
        # frames chain: feeding the arrays back equals stepping the resident state
        e.step_host(*a, s.dt, s.G, s.softening, 2)
        with nb.Engine(n, precision=prec) as e2:
            e2.upload(pm, vel, acc, temp)
            e2.step(s.dt, s.G, s.softening, 2)
            e2.step(s.dt, s.G, s.softening, 2)
            two = e2.download(temp=True)
        for x, k in zip(a, ("pos_mass", "vel", "acc", "temp")):
            assert np.array_equal(x, two[k]), k
        # without temperatures
        b = [x.copy() for x in (pm, vel, acc)]
        e.step_host(*b, None, s.dt, s.G, s.softening, 2)
        assert np.array_equal(b[0], ref["pos_mass"]) and np.array_equal(b[1], ref["vel"])
        with pytest.raises(ValueError):
            e.step_host(pm.astype(np.float32 if prec == "f64" else np.float64), vel, acc, None, s.dt, s.G, s.softening, 1)
        with pytest.raises(nb.NBodyError):
            e.step_host(*[x.copy() for x in (pm, vel, acc)], None, s.dt, s.G, -1.0, 1)


def test_deferred_mass_detection(nb, oracle):
    """The small-system upload returns before the equal-mass count is back; info() and the first large-kernel force
    launch wait for it.  Equal masses -> the equal-mass sweep, a different body list -> the general one, back to back."""
    n = 3000                                                    # one-record I/O, launch-per-step force kernel
    s = nb.scenes.plummer(n)
    pm, vel, _ = s.astype(np.float32)
    with nb.Engine(n, precision="f32") as e:
        e.upload(pm, vel)
        e.force(s.G, s.softening)                               # resolves the detection itself
        acc_equal = e.download()["acc"]
        assert e.info()["uniform_mass"] == 1
        pm2 = pm.copy()
        pm2[:, 3] *= np.linspace(0.5, 1.5, n, dtype=np.float32)
        e.upload(pm, vel)                                       # two uploads back to back: the second one counts
        e.upload(pm2, vel)
        assert e.info()["uniform_mass"] == 0
        e.force(s.G, s.softening)
        acc_general = e.download()["acc"]
        pm3 = pm.copy()
        pm3[7, 3] *= 3.0
        e.upload(pm3, vel)
        e.force(s.G, s.softening)
        assert e.info()["uniform_mass"] == 2                    # equal-mass sweep + one exception
        acc_outlier = e.download()["acc"]
    for acc, p in ((acc_equal, pm), (acc_general, pm2), (acc_outlier, pm3)):
        ref = oracle.compute_gravity(p.astype(np.float64), s.G, s.softening, mt=True)
        err = np.linalg.norm(acc[:, :3] - ref[:, :3], axis=1).max() / np.linalg.norm(ref[:, :3], axis=1).max()
        assert err < 2e-5, err

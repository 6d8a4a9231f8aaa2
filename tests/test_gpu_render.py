"""GPU tests of the headless body renderer through the C ABI: the
planet shaders (idx:83-202) as CUDA kernels against the oracle's C restatement.  Both
sides evaluate one IEEE-rounded fp32 operation per GLSL operation, so the bar is
bit-exact images and identical id buffers."""
import importlib.util
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _inputs(nb):
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(GOLD, "make_golden.py"))
    mg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mg)
    s = nb.scenes.galaxy_default()
    return (s,) + mg.render_inputs(s)


@pytest.mark.parametrize("precision", ["f32", "f64"])
def test_golden_image_bit_exact(nb, precision):
    s, radius, temp, types, seed = _inputs(nb)
    dt = np.float32 if precision == "f32" else np.float64
    g = np.load(os.path.join(GOLD, "render_galaxy_96x54.npz"))
    with nb.Engine(s.n, precision=precision) as e:
        e.upload(s.pos_mass.astype(dt), s.vel.astype(dt), None, temp.astype(dt))
        with pytest.raises(nb.NBodyError):
            e.render(96, 54)                                  # no radii yet
        e.set_radius(radius.astype(dt))
        e.set_materials(types, seed)
        rgba, ids = e.render(96, 54, ids=True)
        assert e.render(96, 54, download=False) is None
    assert np.array_equal(ids, g["ids"])
    assert np.array_equal(rgba, g["rgba"])


@pytest.mark.parametrize("w,h,cam,fov", [(640, 360, (0.0, 80.0, 150.0), 45.0), (333, 211, (-60.0, 25.0, 40.0), 70.0)])
def test_default_scene_vs_oracle(nb, oracle, w, h, cam, fov):
    s, radius, temp, types, seed = _inputs(nb)
    with nb.Engine(s.n, precision="f32") as e:
        pm, vel, _ = s.astype(np.float32)
        e.upload(pm, vel, None, temp.astype(np.float32))
        e.set_radius(radius.astype(np.float32))
        e.set_materials(["rock", "metal", "ice"][0:0] or types, seed)
        e.step(s.dt, s.G, s.softening, 4)                      # render the state the engine holds, after a frame or two
        out = e.download(temp=True)
        rgba, ids = e.render(w, h, camera=cam, fov_deg=fov, ids=True)
    ref, ref_ids = oracle.render(out["pos_mass"], radius, out["temp"], types, seed, w, h, camera=cam, fov_deg=fov)
    assert np.array_equal(ids, ref_ids)
    assert np.array_equal(rgba, ref)
    assert (ids >= 0).sum() > 1000


def test_many_small_bodies_and_camera_inside(nb, oracle):
    """20,000 bodies around the camera, most a pixel or less, a few huge ones, one enclosing the camera."""
    rng = np.random.default_rng(5)
    n = 20000
    pm = np.concatenate([rng.normal(0, 120, (n, 3)), np.ones((n, 1))], axis=1).astype(np.float32)
    radius = rng.uniform(0.05, 1.5, n).astype(np.float32)
    radius[:5] = 40.0
    pm[7, :3] = (0.0, 80.0, 150.0)
    radius[7] = 3.0                                            # the camera sits inside this one: no front hit
    temp = rng.uniform(0, 200, n).astype(np.float32)
    types = rng.integers(0, 3, n).astype(np.int32)
    seed = rng.uniform(0, 100, n).astype(np.float32)
    with nb.Engine(n, precision="f32") as e:
        e.upload(pm, np.zeros((n, 4), np.float32), None, temp)
        e.set_radius(radius)
        e.set_materials(types, seed)
        rgba, ids = e.render(480, 270, ids=True)
    ref, ref_ids = oracle.render(pm, radius, temp, types, seed, 480, 270)
    assert np.array_equal(ids, ref_ids)
    assert np.array_equal(rgba, ref)
    assert len(np.unique(ids)) > 2000 and 7 not in ids


def test_render_errors_and_defaults(nb):
    from nbody_b200 import _lib
    s, radius, temp, types, seed = _inputs(nb)
    with nb.Engine(s.n, precision="f64") as e:
        e.upload(s.pos_mass, s.vel)
        e.set_radius(radius)
        a = e.render(64, 36)                                   # no materials, no temperatures: rock, seed 0, T = 0
        assert np.isfinite(a).all() and (a[..., 3] > 0).any()
        for bad in ((0, 10), (10, -1)):
            with pytest.raises(nb.NBodyError) as ei:
                e.render(*bad)
            assert ei.value.status == _lib.NBODY_EINVAL
        with pytest.raises(nb.NBodyError):
            e.render(64, 36, camera=(0, 0, 0), target=(0, 0, 0))
        with pytest.raises(nb.NBodyError):
            e.set_materials(np.full(s.n, 3), None)
